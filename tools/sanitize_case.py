#!/usr/bin/env python
"""Smallest run that touches every kernel once -- the target for compute-sanitizer (one tool per gpurun call):
    compute-sanitizer --tool racecheck python tools/sanitize_case.py
    compute-sanitizer --tool memcheck  python tools/sanitize_case.py
(Round 2: compute-sanitizer is closed on the GPU pool; the same case runs plain as a smoke of every kernel, and the parity
tests' NaN-prefilled outputs and exact comparisons stand in for memcheck.)
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "psana-ray_b200"))
import numpy as np, torch
from psana_ray import synthetic as syn
from psana_ray.calib import GpuCalibrator

for det, nseg in (("epix10k2M", 6), ("jungfrau4M", 1)):          # 6 segments: quad 1 is rotated by 90 degrees (the row-owner image epilogue)
    spec = syn.detector_spec(det).with_nseg(nseg)
    consts = syn.make_constants(spec, seed=1, cfg_variant="mixed")
    raw = syn.make_raw_frames(spec, consts, 2, seed=2)
    d = torch.from_numpy(raw.view(np.int16)).cuda()
    cal = GpuCalibrator(spec)
    cal.set_constants(consts["peds"], consts.get("gain"), consts.get("offs"), consts.get("cfg"), consts.get("mask"))
    cal.set_geometry(*syn.make_pixel_index_arrays(spec))
    cal.decode(d)
    for cm in (None, (7, 4, 50.0, 10), (7, 7, 50.0, 10)):
        cal.set_common_mode(cm)
        cal.calib(d)
        cal.calib(d, image=True)
    out = torch.empty(cal.out_shape(1, True), dtype=torch.float32, device="cuda")
    cal.calib_image_noalloc(d[:1], out)
    h = torch.from_numpy(raw.view(np.int16)).pin_memory()
    ho = torch.empty((2,) + spec.shape, dtype=torch.float32).pin_memory()
    cal.set_common_mode(None)
    cal.calib_host(h, h_out=ho)
    torch.cuda.synchronize()
    cal.close()
print("sanitize_case done")
