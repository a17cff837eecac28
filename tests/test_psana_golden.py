"""Golden vectors dumped from REAL psana runs (``tools/dump_psana_golden.py``, run where psana is installed) against
the oracle and against the GPU path.  None can be produced offline, so with no ``tests/golden/psana_*.npz`` present
these tests skip; the rehearsal test below runs the whole kit against ``tests/fake_psana`` instead, so the day a real
dump lands the path to a pinned oracle is one command ("Pinning the oracle").

Tolerances are north_star's: gain decode and masks bit-exact, calibrated pixels <= 1e-5 relative, medians exact
(a wrong median moves every eligible pixel of its bank / row / column by more than the tolerance)."""
import glob
import os
import sys

import numpy as np
import pytest

from oracle import psana_calib as oc

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "psana_*.npz")))
RTOL = 1e-5


def _load(path):
    z = np.load(path, allow_pickle=False)
    d = {k: z[k] for k in z.files}
    cm = d["cmpars_used"] if d["cmpars_used"].size else d["cmpars_run"]
    d["cmpars"] = None if cm.size < 3 or int(cm[0]) == 0 else ((int(cm[0]), int(cm[1]), float(cm[2])) + ((int(cm[3]),) if cm.size > 3 else ()))
    d["consts"] = {k: d[s] for k, s in (("peds", "peds"), ("gain", "gain"), ("offs", "offs"), ("cfg", "cfg"), ("mask", "calib_mask")) if s in d}
    d["kind"] = str(d["kind"])
    return d


def _close(got, exp, what):
    got = np.asarray(got, dtype=np.float32); exp = np.asarray(exp, dtype=np.float32)
    assert got.shape == exp.shape, (what, got.shape, exp.shape)
    # masked pixels are exact zeros on both sides: mask parity is bit-exact
    np.testing.assert_array_equal(got == 0, exp == 0, err_msg=f"{what}: zero (masked) pixels differ")
    tol = RTOL * np.maximum(np.abs(exp), 1e-3)
    bad = np.abs(got.astype(np.float64) - exp) > tol
    assert not bad.any(), f"{what}: {int(bad.sum())} of {bad.size} pixels beyond {RTOL} relative; first {np.argwhere(bad)[0]}: {got[bad][0]} vs {exp[bad][0]}"


def check_oracle(d):
    for f in range(d["raw"].shape[0]):
        _close(oc.det_calib(d["kind"], d["raw"][f], d["consts"], d["cmpars"]), d["calib"][f], f"oracle calib, event {f}")
        if "image" in d:
            _close(oc.img_from_pixel_arrays(d["iX"], d["iY"], d["calib"][f]), d["image"][f], f"oracle image assembly, event {f}")


def check_gpu(d):
    import torch
    from psana_ray.calib import GpuCalibrator
    from psana_ray import synthetic as syn
    spec = syn.detector_spec(str(d["detector_name"])).with_nseg(d["raw"].shape[1])
    c = d["consts"]
    with GpuCalibrator(spec, device=0) as cal:
        cal.set_constants(c["peds"], c.get("gain"), c.get("offs"), c.get("cfg"), c.get("mask"))
        cal.set_common_mode(d["cmpars"])
        d_raw = torch.from_numpy(d["raw"].view(np.int16)).cuda()
        got = cal.calib(d_raw).cpu().numpy()
        for f in range(got.shape[0]):
            _close(got[f], d["calib"][f], f"GPU calib, event {f}")
        if "image" in d:
            cal.set_geometry(d["iX"], d["iY"])
            img = cal.calib(d_raw, image=True).cpu().numpy()
            for f in range(img.shape[0]):
                _close(img[f], d["image"][f], f"GPU image, event {f}")


@pytest.mark.skipif(not GOLDEN, reason="no tests/golden/psana_*.npz: dump one with tools/dump_psana_golden.py on a box with psana")
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_matches_psana_dump(path):
    check_oracle(_load(path))


@pytest.mark.gpu
@pytest.mark.skipif(not GOLDEN, reason="no tests/golden/psana_*.npz: dump one with tools/dump_psana_golden.py on a box with psana")
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_gpu_matches_psana_dump(path):
    check_gpu(_load(path))


def _rehearsal_dump(tmp_path, monkeypatch, det, cm):
    """The dump tool run against the psana stub: same code path as on a psana box."""
    fake = os.path.join(HERE, "fake_psana")
    monkeypatch.syspath_prepend(fake)
    monkeypatch.syspath_prepend(os.path.join(os.path.dirname(HERE), "tools"))
    monkeypatch.setenv("PSANA_FAKE_NSEG", "2"); monkeypatch.setenv("PSANA_FAKE_NUM_EVENTS", "3"); monkeypatch.setenv("PSANA_FAKE_CFG", "mixed")
    for m in [k for k in sys.modules if k.split(".")[0] in ("psana", "Detector")]:
        monkeypatch.delitem(sys.modules, m)
    import dump_psana_golden
    out = str(tmp_path / "psana_rehearsal.npz")
    dump_psana_golden.main(["--exp", "mfxl1038923", "--run", "58", "--detector_name", det, "--events", "2", "--cmpars", cm, "--out", out])
    for m in [k for k in sys.modules if k.split(".")[0] in ("psana", "Detector")]:
        sys.modules.pop(m, None)
    return _load(out)


@pytest.mark.parametrize("det,cm", [("epix10k2M", "7,6,50,10"), ("jungfrau4M", "7,3,50,10")])
def test_golden_kit_rehearsal_oracle(tmp_path, monkeypatch, det, cm):
    d = _rehearsal_dump(tmp_path, monkeypatch, det, cm)
    assert d["raw"].shape[0] == 2 and d["cmpars"] == tuple(float(v) if i == 2 else int(v) for i, v in enumerate(cm.split(",")))
    assert (d["calib_mask"] >= d["bad_pixel_mask"]).all() and (d["calib_mask"] != d["bad_pixel_mask"]).any()
    check_oracle(d)
    # the check has teeth: a median off by one pixel's worth is caught
    d["calib"][0, 0, 5:60, 3:40] += 0.01
    with pytest.raises(AssertionError):
        check_oracle(d)


@pytest.mark.gpu
@pytest.mark.parametrize("det,cm", [("epix10k2M", "7,6,50,10"), ("jungfrau4M", "7,3,50,10")])
def test_golden_kit_rehearsal_gpu(tmp_path, monkeypatch, det, cm):
    check_gpu(_rehearsal_dump(tmp_path, monkeypatch, det, cm))
