#!/usr/bin/env python
"""Kernel-only sweep over tuning variants of libpsana_ray_cuda (built by build.build_variant).

    python tools/kbench.py --workload calib --libs lib/variants/*.so --fpb 16,32,64 --iters 50

One process, one set of device buffers; each variant library is loaded separately through ctypes.
Timing: CUDA events on the launching stream around `iters` back-to-back launches after 5 warm-ups;
inputs (277 MB raw + 554 MB out for epix10k2M x 64) are larger than L2.
"""
import argparse, ctypes as C, glob, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "psana-ray_b200"))
import numpy as np
import torch
from psana_ray import _native as N, synthetic as syn

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="calib", choices=["calib", "cm", "jungfrau", "image_nocm", "image"])
ap.add_argument("--libs", nargs="+", default=[N.LIB_PATH])
ap.add_argument("--fpb", default="0")
ap.add_argument("--iters", type=int, default=50)
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--out", default=None)
ap.add_argument("--cm-mode", type=int, default=4)
ap.add_argument("--det", default=None)
ap.add_argument("--npix-min", type=int, default=10)
ap.add_argument("--gap", type=int, default=10, help="panel gap of the synthetic geometry (image width = f(gap): 10 -> W % 4 == 2, 12 -> W % 4 == 0)")
ap.add_argument("--geo", default="default", choices=["default", "flat", "rot90", "stack", "col1gap", "col2nogap", "col2pad"], help="flat: all panels unrotated in a 2-column stack; rot90: all rotated by 90 degrees")
ap.add_argument("--opt", action="append", default=[], help="name=value for psr_set_option, e.g. image_gather=1")
a = ap.parse_args()

det = a.det or ("jungfrau4M" if a.workload == "jungfrau" else "epix10k2M")
spec = syn.detector_spec(det)
consts = syn.make_constants(spec, seed=0)
raw = syn.make_raw_frames(spec, consts, 8, seed=1)
raw = np.concatenate([raw] * max(1, a.batch // 8))[:a.batch]
d_raw = torch.from_numpy(raw.view(np.int16)).cuda()
geo = syn.make_pixel_index_arrays(spec, gap=a.gap) if a.workload in ("image_nocm", "image") else None
if geo is not None and a.geo != "default":
    R, Cc = spec.rows, spec.cols
    r, c = np.meshgrid(np.arange(R), np.arange(Cc), indexing="ij")
    iX = np.empty(spec.shape, dtype=np.int64); iY = np.empty(spec.shape, dtype=np.int64)
    for sg in range(spec.nseg):
        pr, pc = divmod(sg, 2)
        if a.geo == "stack":                      # image == calib layout: one column of panels, no gaps
            iX[sg], iY[sg] = r + sg * R, c
        elif a.geo == "col1gap":                  # one column of panels, gap rows between them
            iX[sg], iY[sg] = r + sg * (R + a.gap), c
        elif a.geo == "col2nogap":                # two columns, no gaps at all
            iX[sg], iY[sg] = r + pr * R, c + pc * Cc
        elif a.geo == "col2pad":                  # two columns, gap only between the columns (rows contiguous vertically)
            iX[sg], iY[sg] = r + pr * R, c + pc * (Cc + a.gap)
        elif a.geo == "flat":
            iX[sg], iY[sg] = r + pr * (R + a.gap), c + pc * (Cc + a.gap)
        else:
            iX[sg], iY[sg] = c + pr * (Cc + a.gap), (R - 1 - r) + pc * (R + a.gap)
    geo = (iX.astype(np.int32), iY.astype(np.int32))
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
libs = []
for pat in a.libs:
    libs += sorted(glob.glob(pat)) or [pat]
rows = []
for path in libs:
    lib = N.load(path)
    ctx = C.c_void_p()
    kind = 0 if spec.kind == "epix10ka" else 1
    assert lib.psr_create(C.byref(ctx), kind, spec.nseg, spec.rows, spec.cols, 0) == 0, lib.psr_last_error()
    p = lambda x: None if x is None else np.ascontiguousarray(x).ctypes.data_as(C.c_void_p)
    assert lib.psr_set_constants(ctx, p(consts["peds"]), p(consts.get("gain")), p(consts.get("offs")), p(consts.get("cfg")), p(consts.get("mask")), spec.nmodes) == 0, lib.psr_last_error()
    out_kind = 0
    nout = spec.npix
    if a.workload in ("cm", "image"):
        lib.psr_set_common_mode(ctx, a.cm_mode, C.c_float(100.0), a.npix_min, 0)
    if geo is not None:
        assert lib.psr_set_geometry(ctx, p(geo[0]), p(geo[1])) == 0
        H, W = C.c_int(), C.c_int(); lib.psr_image_shape(ctx, C.byref(H), C.byref(W)); nout = H.value * W.value; out_kind = 1
    for kv in a.opt:
        k, v = kv.split("="); assert lib.psr_set_option(ctx, k.encode(), int(v)) == 0, lib.psr_last_error()
    d_out = torch.empty((a.batch, nout), dtype=torch.float32, device="cuda")
    alg = (2 * spec.npix + 4 * nout) * a.batch
    st = torch.cuda.current_stream().cuda_stream
    for fpb in [int(x) for x in a.fpb.split(",")]:
        if fpb > 0:
            lib.psr_set_option(ctx, b"frames_per_block", fpb)
        for _ in range(5):
            assert lib.psr_calib_batch(ctx, d_raw.data_ptr(), a.batch, d_out.data_ptr(), out_kind, st) == 0, lib.psr_last_error()
        torch.cuda.synchronize()
        best = 1e9; tot = 0.0
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.iters):
                lib.psr_calib_batch(ctx, d_raw.data_ptr(), a.batch, d_out.data_ptr(), out_kind, st)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.iters
            best = min(best, ms); tot += ms
        row = {"lib": os.path.basename(path), "workload": a.workload, "fpb": fpb, "ms_best": best, "ms_avg": tot / 3,
               "fps": a.batch / (tot / 3) * 1e3, "gbs": alg / (tot / 3) / 1e6, "frac": alg / (tot / 3) / 1e6 / peak}
        rows.append(row)
        print(f"{row['lib']:60s} fpb={fpb:3d}  {row['ms_avg']*1e3:8.1f} us  {row['fps']:10.0f} fps  {row['gbs']:7.0f} GB/s  frac {row['frac']:.3f}", flush=True)
    lib.psr_destroy(ctx)
    del d_out
if a.out:
    with open(a.out, "w") as fh:
        for r in rows:
            fh.write(json.dumps(r) + "\n")
