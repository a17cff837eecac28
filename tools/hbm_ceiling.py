import torch, time
n = 64*2162688
raw = torch.randint(0, 16384, (n,), dtype=torch.int16, device="cuda")
out = torch.empty(n, dtype=torch.float32, device="cuda")
src32 = torch.empty(n, dtype=torch.float32, device="cuda")
def t(f, iters=40):
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/iters*1e3
us = t(lambda: out.copy_(raw)); print("int16->f32 convert  %.1f us  %.0f GB/s (6 B/el)" % (us, n*6/us/1e3))
us = t(lambda: out.zero_()); print("write only 4B/el    %.1f us  %.0f GB/s" % (us, n*4/us/1e3))
us = t(lambda: out.copy_(src32)); print("copy f32 (8 B/el)   %.1f us  %.0f GB/s" % (us, n*8/us/1e3))
h = torch.empty(n//2, dtype=torch.float32, device="cuda")
us = t(lambda: h.copy_(src32[:n//2])); print("copy f32 half       %.1f us  %.0f GB/s" % (us, n*4/us/1e3))
us = t(lambda: raw.sum(dtype=torch.int64)); print("read only 2B/el     %.1f us  %.0f GB/s" % (us, n*2/us/1e3))
us = t(lambda: src32.sum()); print("read only 4B/el     %.1f us  %.0f GB/s" % (us, n*4/us/1e3))
