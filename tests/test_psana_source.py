"""The real-data adapter: ``PsanaSource`` and ``psana-ray-producer --source psana`` against
``tests/fake_psana``, a stub of the psana calls they make (psana itself cannot be installed offline).

Reference call sites mirrored: PsanaWrapperSmd(exp, run, detector_name) (ref: producer.py:150-154),
create_bad_pixel_mask() (:81), iter_events (:88) -- here raw frames + per-run constants, calibrated on the GPU."""
import os
import subprocess
import sys
import time

import numpy as np
import pytest

import kat
from oracle import psana_calib as oc

HERE = os.path.dirname(os.path.abspath(__file__))
FAKE = os.path.join(HERE, "fake_psana")
PKG = os.path.join(os.path.dirname(HERE), "psana-ray_b200")
EXP, RUN = "mfxl1038923", 58


@pytest.fixture
def fake_psana(monkeypatch):
    monkeypatch.syspath_prepend(FAKE)
    monkeypatch.setenv("PSANA_FAKE_NSEG", str(NSEG))
    for m in [k for k in sys.modules if k == "psana" or k.startswith("psana.") or k == "Detector" or k.startswith("Detector.")]:
        monkeypatch.delitem(sys.modules, m)
    yield
    for m in [k for k in sys.modules if k == "psana" or k.startswith("psana.") or k == "Detector" or k.startswith("Detector.")]:
        sys.modules.pop(m, None)


NSEG = 2                                                                 # fewer segments than the real detectors: same code, seconds instead of minutes


def _synthetic(det, nev, cfg="aml"):
    from psana_ray.event_source import SyntheticSource
    return SyntheticSource(EXP, RUN, det, num_events=nev, cfg_variant=cfg, nseg=NSEG)


@pytest.mark.parametrize("det,cfg", [("epix10k2M", "aml"), ("epix10k2M", "mixed"), ("jungfrau4M", "aml")])
def test_psana_source_constants_masks_geometry(fake_psana, monkeypatch, det, cfg):
    monkeypatch.setenv("PSANA_FAKE_NUM_EVENTS", "6")
    monkeypatch.setenv("PSANA_FAKE_CFG", cfg)
    monkeypatch.setenv("PSANA_FAKE_CM", "7,4,50,10")
    from psana_ray.event_source import PsanaSource
    src = PsanaSource(EXP, RUN, det)
    ref = _synthetic(det, 6, cfg)
    c = src.constants()
    rc = ref.constants()
    assert src.spec.shape == ref.spec.shape
    np.testing.assert_array_equal(c["peds"], rc["peds"])
    np.testing.assert_array_equal(c["gain"], rc["gain"])
    if det.startswith("jungfrau"):
        np.testing.assert_array_equal(c["offs"], rc["offs"])
        assert "cfg" not in c
    else:
        # built from the detector's configuration through psana's own helper: bits 4..2 of the synthetic bytes
        np.testing.assert_array_equal(c["cfg"], rc["cfg"] & 0x1C)
    # the calibration mask is the pixel-status mask, NOT the producer's bad-pixel mask (ref: producer.py:81,92-93)
    bad = src.create_bad_pixel_mask()
    np.testing.assert_array_equal(bad, rc["mask"])
    assert c["mask"].dtype == np.uint8 and (c["mask"] >= bad).all() and (c["mask"] != bad).any()
    assert c["mask"][:, 0, :].any() and not bad[:, 0, :].any()           # edges: good for det.calib, dropped by the producer mask
    assert src.common_mode() == (7, 4, 50.0, 10)
    gx, gy = src.geometry()
    np.testing.assert_array_equal(gx, ref.geometry()[0]); np.testing.assert_array_equal(gy, ref.geometry()[1])
    # options
    assert PsanaSource(EXP, RUN, det, common_mode=(7, 2, 30.0)).common_mode() == (7, 2, 30.0)
    np.testing.assert_array_equal(PsanaSource(EXP, RUN, det, calib_mask="total").constants()["mask"], rc["mask"])
    assert "mask" not in PsanaSource(EXP, RUN, det, calib_mask="none").constants()


def test_psana_source_without_common_mode_constant(fake_psana, monkeypatch):
    monkeypatch.delenv("PSANA_FAKE_CM", raising=False)
    from psana_ray.event_source import PsanaSource
    assert PsanaSource(EXP, RUN, "epix10k2M").common_mode() is None


def test_epix10ka_cfg_fallback_builds_from_config_object(fake_psana, monkeypatch):
    """psana releases without cbits_config_epix10ka_any: the bytes come from the configuration object through
    epix10ka_config_bytes -- and agree with psana's helper on the ASIC -> quadrant layout of the stub."""
    monkeypatch.setenv("PSANA_FAKE_CFG", "mixed")
    from Detector import UtilsEpix10ka as ue
    from psana_ray.event_source import PsanaSource
    want = PsanaSource(EXP, RUN, "epix10k2M").constants()["cfg"]
    monkeypatch.delattr(ue, "cbits_config_epix10ka_any")
    got = PsanaSource(EXP, RUN, "epix10k2M").constants()["cfg"]
    np.testing.assert_array_equal(got, want)
    monkeypatch.delattr(ue, "get_epix10ka_any_config_object")
    with pytest.raises(RuntimeError, match="epix10ka_cfg"):
        PsanaSource(EXP, RUN, "epix10k2M").constants()
    override = np.full((NSEG, 352, 384), 12, dtype=np.uint8)
    np.testing.assert_array_equal(PsanaSource(EXP, RUN, "epix10k2M", epix10ka_cfg=override).constants()["cfg"], override)


@pytest.mark.parametrize("size", [1, 3])
def test_psana_source_events_are_sharded_like_mpidatasource(fake_psana, monkeypatch, size):
    monkeypatch.setenv("PSANA_FAKE_NUM_EVENTS", "7")
    from psana_ray.event_source import PsanaSource
    ref = _synthetic("epix10k2M", 7)
    seen = []
    for rank in range(size):
        monkeypatch.setenv("RANK", str(rank)); monkeypatch.setenv("WORLD_SIZE", str(size))
        src = PsanaSource(EXP, RUN, "epix10k2M", rank=rank, size=size)
        mine = list(src.iter_raw_events())
        want = list(range(rank, 7, size))                                 # nevent % size == rank
        assert len(mine) == len(want)
        for (raw, energy), n in zip(mine, want):
            assert raw.dtype == np.uint16 and raw.shape == ref.spec.shape
            np.testing.assert_array_equal(raw, ref.raw_event(n))
            assert energy == ref.photon_energy(n)
        seen += want
    assert sorted(seen) == list(range(7))


# ---- the whole producer on a real psana-shaped run, on the GPU ---------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("det,calib,cm,cfg,badmask", [("epix10k2M", True, "7,4,50,10", "mixed", True),
                                                      ("epix10k2M", False, None, "aml", False),
                                                      ("jungfrau4M", True, "7,7,50,10", "aml", True)])
def test_producer_cli_source_psana_end_to_end(tmp_path, monkeypatch, det, calib, cm, cfg, badmask):
    from psana_ray import shared_queue as SQ
    from psana_ray.data_reader import DataReader
    root = str(tmp_path / "q")
    monkeypatch.setenv(SQ.DEFAULT_ROOT_ENV, root)
    nev = 10
    env = dict(os.environ, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", PSANA_RAY_DRAIN_S="60", PSANA_FAKE_NUM_EVENTS=str(nev),
               PSANA_FAKE_CFG=cfg, PSANA_FAKE_NSEG=str(NSEG), PYTHONPATH=os.pathsep.join([PKG, FAKE, os.path.dirname(HERE), os.environ.get("PYTHONPATH", "")]))
    env.pop("PSANA_FAKE_CM", None)
    cmd = [sys.executable, "-m", "psana_ray.producer", "--exp", EXP, "--run", str(RUN), "--detector_name", det, "--source", "psana",
           "--queue_size", "16", "--batch_size", "4", "--num_consumers", "1", "--queue_name", "shared_queue", "--ray_namespace", "my",
           "--reliable_sentinels"]
    if calib:
        cmd.append("--calib")
    if badmask:
        cmd.append("--uses_bad_pixel_mask")
    if cm:
        cmd += ["--common_mode", cm]
    proc = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    items = []
    try:
        t0 = time.time()
        while True:
            try:
                reader = DataReader(); reader.connect(); break
            except ValueError:
                assert time.time() - t0 < 180 and proc.poll() is None, "producer died: " + (proc.stdout.read() if proc.poll() is not None else "")
                time.sleep(0.2)
        t0 = time.time()
        while not reader.end_of_stream and time.time() - t0 < 240:
            it = reader.read()
            if it is None:
                assert proc.poll() is None or reader.end_of_stream or proc.returncode == 0, "producer died: " + proc.stdout.read()
                time.sleep(0.005); continue
            items.append((it[0], it[1], it[2].cpu().numpy(), it[3], it[2].is_cuda))
        assert reader.end_of_stream
        reader.close()
        assert proc.wait(timeout=120) == 0, proc.stdout.read()
    finally:
        if proc.poll() is None:
            proc.kill()
        try:
            SQ.get_queue("shared_queue", "my").shutdown()
        except Exception:
            pass
    assert len(items) == nev and all(i[4] for i in items)
    ref = _synthetic(det, nev, cfg)
    consts = dict(ref.constants())
    status = consts["mask"].copy()
    status[:, 0, :] = 1; status[:, -1, :] = 1; status[:, :, 0] = 1; status[:, :, -1] = 1
    bad = consts["mask"]
    consts["mask"] = status                                              # det.calib multiplies by the status mask
    if "cfg" in consts:
        consts["cfg"] = consts["cfg"] & 0x1C
    cmpars = None if cm is None else tuple(float(v) if i == 2 else int(v) for i, v in enumerate(cm.split(",")))
    for rank, idx, data, energy, _ in items:
        n = idx                                                           # one rank: local index == event number
        assert energy == ref.photon_energy(n)
        if calib:
            exp = oc.producer_masks(oc.det_calib(ref.spec.kind, ref.raw_event(n), consts, cmpars), bad if badmask else None, None)
        else:
            exp = oc.producer_masks(oc.det_image(ref.spec.kind, ref.raw_event(n), consts, *ref.geometry(), cmpars=cmpars))
        assert data.shape == exp.shape
        kat.assert_same_bits(data, exp, f"event {n}")
