"""psr_calib_image_noalloc: the image-mode hot call for callers that must not allocate device memory
(a consumer on a GPU whose memory another job already owns, or a loop that cannot take the stall
an allocation causes).  Its scratch is an array inside the library, so this test checks what such a
caller relies on: the results equal psr_calib_batch(PSR_OUT_IMAGE) -- and the oracle -- bit for bit,
also on a second call that finds the first call's data still in the scratch; the device's free
memory does not move across the calls; and oversized requests are refused.

Collected last on purpose (file name): a failure here cannot hide another test's result."""
import numpy as np
import pytest

import kat
from oracle import psana_calib as oc
from psana_ray import synthetic as syn

pytestmark = pytest.mark.gpu


def _setup(torch, spec, cmpars, seed):
    from psana_ray.calib import GpuCalibrator
    consts = syn.make_constants(spec, seed=seed, cfg_variant="mixed")
    cal = GpuCalibrator(spec, device=0)
    cal.set_constants(consts["peds"], consts.get("gain"), consts.get("offs"), consts.get("cfg"), consts.get("mask"))
    cal.set_common_mode(cmpars)
    cal.set_geometry(*syn.make_pixel_index_arrays(spec))
    return cal, consts


@pytest.mark.parametrize("cmpars", [(7, 4, 50.0, 10), (7, 7, 50.0, 10), (7, 2, 50.0, 10), None])
def test_noalloc_matches_regular_entry_point_on_two_calls_in_a_row(cmpars):
    import torch
    spec = syn.EPIX10K2M                        # full detector: 3 frames fill 26 of the 32 MiB
    cal, consts = _setup(torch, spec, cmpars, seed=123)
    raw = syn.make_raw_frames(spec, consts, 5, seed=124)
    d = torch.from_numpy(raw.view(np.int16)).cuda()
    ref = cal.calib(d, image=True)             # existing entry point (context-allocated scratch)
    # call 1: three frames; call 2 straight after: two other frames, scratch still holds call 1's data
    out1 = torch.full(cal.out_shape(3, True), float("nan"), dtype=torch.float32, device="cuda")
    out2 = torch.full(cal.out_shape(2, True), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]          # cudaMemGetInfo: the DEVICE's free bytes, whoever allocates
    cal.calib_image_noalloc(d[:3], out1)
    torch.cuda.synchronize()
    free_mid = torch.cuda.mem_get_info()[0]
    cal.calib_image_noalloc(d[3:], out2)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    # what the name promises: neither call took device memory (the context, its tables and the built-in array were all
    # in place when psr_create / set_* returned)
    assert free_mid == free_before and free_after == free_before, (free_before, free_mid, free_after)
    assert torch.equal(out1.view(torch.int32), ref[:3].view(torch.int32))      # tolerance: 0 ulp, as everywhere in tests/
    assert torch.equal(out2.view(torch.int32), ref[3:].view(torch.int32))
    # and against the oracle itself, not only against the sibling entry point
    iX, iY = syn.make_pixel_index_arrays(spec)
    got = np.concatenate([out1.cpu().numpy(), out2.cpu().numpy()])
    for f in (0, 4):
        kat.assert_same_bits(got[f], oc.det_image(spec.kind, raw[f], consts, iX, iY, cmpars), f"frame {f} vs oracle, cmpars {cmpars}")
    # and once more with the same input: idempotent
    out1b = torch.empty_like(out1)
    cal.calib_image_noalloc(d[:3], out1b)
    assert torch.equal(out1b.view(torch.int32), out1.view(torch.int32))


def test_noalloc_rejects_requests_beyond_the_built_in_scratch():
    import torch
    from psana_ray import _native as N
    spec = syn.EPIX10K2M
    cal, consts = _setup(torch, spec, (7, 4, 50.0, 10), seed=125)
    limit = N.PSR_NOALLOC_SCRATCH_BYTES // (4 * spec.npix)
    assert limit == 3
    raw = torch.zeros((limit + 1,) + spec.shape, dtype=torch.int16, device="cuda")
    out = torch.empty(cal.out_shape(limit + 1, True), dtype=torch.float32, device="cuda")
    with pytest.raises(N.PsrError) as ei:
        cal.calib_image_noalloc(raw, out)
    assert "at most 3 frames" in str(ei.value)
    cal.calib_image_noalloc(raw[:limit], out[:limit])            # exactly at the limit is accepted
    torch.cuda.synchronize()


@pytest.mark.parametrize("cmpars", [(7, 4, 50.0, 10), (7, 3, 50.0, 10)])
def test_noalloc_jungfrau_two_frames_fit_three_do_not(cmpars):
    import torch
    from psana_ray import _native as N
    spec = syn.JUNGFRAU4M                       # 16 MiB per calibrated frame: two fill the 32 MiB array exactly
    assert N.PSR_NOALLOC_SCRATCH_BYTES // (4 * spec.npix) == 2
    cal, consts = _setup(torch, spec, cmpars, seed=126)
    raw = syn.make_raw_frames(spec, consts, 3, seed=127)
    d = torch.from_numpy(raw.view(np.int16)).cuda()
    ref = cal.calib(d, image=True)
    out = torch.full(cal.out_shape(2, True), float("nan"), dtype=torch.float32, device="cuda")
    cal.calib_image_noalloc(d[:2], out)
    assert torch.equal(out.view(torch.int32), ref[:2].view(torch.int32))
    out1 = torch.full(cal.out_shape(1, True), float("nan"), dtype=torch.float32, device="cuda")
    cal.calib_image_noalloc(d[2:], out1)
    assert torch.equal(out1.view(torch.int32), ref[2:].view(torch.int32))
    iX, iY = syn.make_pixel_index_arrays(spec)
    kat.assert_same_bits(out1.cpu().numpy()[0], oc.det_image(spec.kind, raw[2], consts, iX, iY, cmpars), "jungfrau frame 2 vs oracle")
    with pytest.raises(N.PsrError) as ei:
        cal.calib_image_noalloc(d, torch.empty(cal.out_shape(3, True), dtype=torch.float32, device="cuda"))
    assert "at most 2 frames" in str(ei.value)
