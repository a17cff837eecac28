# quick bit-exactness check of every variant library against the main one (which passed the oracle tests) on cm mode 4, calib + image
import sys, glob, ctypes as C, numpy as np, torch
sys.path.insert(0, 'psana-ray_b200')
from psana_ray import _native as N, synthetic as syn
spec = syn.EPIX10K2M
consts = syn.make_constants(spec, seed=0, cfg_variant="mixed")
raw = syn.make_raw_frames(spec, consts, 9, seed=1)
d_raw = torch.from_numpy(raw.view(np.int16)).cuda()
iX, iY = syn.make_pixel_index_arrays(spec)
outs = {}
for path in [N.LIB_PATH] + sorted(glob.glob('psana-ray_b200/lib/variants/*.so')):
    lib = N.load(path); ctx = C.c_void_p()
    assert lib.psr_create(C.byref(ctx), 0, spec.nseg, spec.rows, spec.cols, 0) == 0
    p = lambda x: None if x is None else np.ascontiguousarray(x).ctypes.data_as(C.c_void_p)
    assert lib.psr_set_constants(ctx, p(consts["peds"]), p(consts["gain"]), None, p(consts["cfg"]), p(consts["mask"]), 7) == 0
    lib.psr_set_common_mode(ctx, 4, C.c_float(30.0), 10, 0)
    assert lib.psr_set_geometry(ctx, p(iX), p(iY)) == 0
    H, W = C.c_int(), C.c_int(); lib.psr_image_shape(ctx, C.byref(H), C.byref(W))
    res = []
    for kind, n in ((0, spec.npix), (1, H.value * W.value)):
        o = torch.full((9, n), float('nan'), dtype=torch.float32, device='cuda')
        assert lib.psr_calib_batch(ctx, d_raw.data_ptr(), 9, o.data_ptr(), kind, torch.cuda.current_stream().cuda_stream) == 0, lib.psr_last_error()
        torch.cuda.synchronize(); res.append(o.cpu().numpy().view(np.uint32))
    outs[path] = res
    lib.psr_destroy(ctx)
ref = outs[N.LIB_PATH]
for k, v in outs.items():
    print(k.split('/')[-1], 'calib', np.array_equal(v[0], ref[0]), 'image', np.array_equal(v[1], ref[1]))
