"""End-to-end streaming on a GPU: psana-ray-producer (separate processes, real CLI) -> queue server ->
DataReader in this process, with calibrated batches handed over as CUDA-IPC descriptors.  Every event
is checked bit-for-bit against the oracle."""
import ctypes as C
import os
import subprocess
import sys
import time

import numpy as np
import pytest

import kat
from oracle import psana_calib as oc
from psana_ray import shared_queue as SQ
from psana_ray.data_reader import DataReader
from psana_ray.event_source import SyntheticSource

pytestmark = pytest.mark.gpu
PKG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "psana-ray_b200")


def _launch(env, rank, size, extra):
    e = dict(env)
    e.update(RANK=str(rank), WORLD_SIZE=str(size), LOCAL_RANK=str(rank), PYTHONPATH=PKG + os.pathsep + e.get("PYTHONPATH", ""),
             PSANA_RAY_DRAIN_S="60")
    cmd = [sys.executable, "-m", "psana_ray.producer", "--exp", "mfxl1038923", "--run", "58"] + extra
    return subprocess.Popen(cmd, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)


def _consume(reader, n_expected, timeout=180.0):
    items, t0 = [], time.time()
    while len(items) < n_expected and time.time() - t0 < timeout:
        it = reader.read()
        if it is None:
            if reader.end_of_stream:
                break
            time.sleep(0.005)
            continue
        items.append([it[0], it[1], it[2].cpu().numpy(), it[3], it[2].is_cuda])
    return items


@pytest.mark.parametrize("nranks,calib,cm", [(1, True, None), (2, True, "7,4,50,10"), (1, False, "7,4,50,10")])
def test_producer_cli_to_datareader_over_cuda_ipc(tmp_path, monkeypatch, nranks, calib, cm):
    root = str(tmp_path / "q")
    monkeypatch.setenv(SQ.DEFAULT_ROOT_ENV, root)
    nev, det = 20, "epix10k2M"
    extra = ["--detector_name", det, "--queue_size", "16", "--batch_size", "4", "--num_events", str(nev),
             "--num_consumers", "1", "--queue_name", "shared_queue", "--ray_namespace", "my",
             "--reliable_sentinels"]     # queue_size 16 < 20 events: without it the sentinel is dropped, as in the reference
    if calib:
        extra.append("--uses_bad_pixel_mask")
    if calib:
        extra.append("--calib")
    if cm:
        extra += ["--common_mode", cm]
    procs = [_launch(os.environ, r, nranks, extra) for r in range(nranks)]
    try:
        t0 = time.time()
        while True:                                                 # the consumer starts before or after the producer: wait for the name
            try:
                reader = DataReader(); reader.connect(); break
            except ValueError:
                assert time.time() - t0 < 120 and all(p.poll() is None for p in procs), "producer died: " + "".join(p.stdout.read() for p in procs if p.poll() is not None)
                time.sleep(0.2)
        items = _consume(reader, nev)
        sentinel = None
        t0 = time.time()
        while not reader.end_of_stream and time.time() - t0 < 60:   # the sentinel follows the last event
            sentinel = reader.read(); time.sleep(0.01)
        assert reader.end_of_stream
        reader.close()                                              # releases the last slab -> producers may exit
        for p in procs:
            assert p.wait(timeout=120) == 0, p.stdout.read()
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
        try:
            SQ.get_queue("shared_queue", "my").shutdown()
        except Exception:
            pass
    assert len(items) == nev and all(i[4] for i in items)           # every payload was a CUDA tensor
    cmpars = None if cm is None else (7, 4, 50.0, 10)
    seen = set()
    for rank, idx, data, energy, _ in items:
        src = SyntheticSource("mfxl1038923", 58, det, rank=rank, size=nranks, num_events=nev, common_mode=cmpars)
        n = src.event_numbers()[idx]                                # (rank, idx) is the event key (ref: producer.py:88,101)
        seen.add(n)
        assert energy == src.photon_energy(n)
        c = src.constants()
        if calib:
            exp = oc.producer_masks(oc.det_calib(src.spec.kind, src.raw_event(n), c, cmpars), src.create_bad_pixel_mask(), None)
        else:
            exp = oc.producer_masks(oc.det_image(src.spec.kind, src.raw_event(n), c, *src.geometry(), cmpars=cmpars))
        assert data.shape == exp.shape
        kat.assert_same_bits(data, exp, f"event {n} from rank {rank}")
    assert seen == set(range(nev))
    # per-rank order is preserved through the single FIFO
    for r in range(nranks):
        idxs = [i[1] for i in items if i[0] == r]
        assert idxs == sorted(idxs)


def test_ipc_event_handles():
    import torch
    from psana_ray import _native as N
    lib = N.load()
    ev = C.c_void_p()
    hb = (C.c_ubyte * N.PSR_IPC_HANDLE_BYTES)()
    N.check(lib.psr_ipc_event_create(C.byref(ev), hb, 0))
    assert any(hb)
    st = torch.cuda.current_stream().cuda_stream
    x = torch.ones(1 << 20, device="cuda") * 2
    N.check(lib.psr_event_record(ev, st))
    N.check(lib.psr_event_wait(ev, st))
    N.check(lib.psr_event_synchronize(ev))
    assert float(x.sum()) == 2 * (1 << 20)
    N.check(lib.psr_event_destroy(ev))
    # a mem handle of a plain allocation exports and must not be opened by its own process
    p = C.c_void_p()
    N.check(lib.psr_device_alloc(C.byref(p), 1 << 20, 0))
    mh = (C.c_ubyte * N.PSR_IPC_HANDLE_BYTES)()
    N.check(lib.psr_ipc_export_mem(p, mh))
    assert any(mh)
    N.check(lib.psr_device_free(p, 0))


def test_bridge_over_nccl_between_two_gpus(tmp_path):
    """producer + queue A on GPU 0 -> psana_ray.bridge (NCCL send/recv) -> queue B -> DataReader on GPU 1.
    Two queue roots stand for two boxes; needs two devices (NCCL refuses two ranks on one)."""
    import socket
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root_a, root_b = str(tmp_path / "boxA"), str(tmp_path / "boxB")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    nev, det = 12, "epix10k2M"
    extra = ["--detector_name", det, "--queue_size", "16", "--batch_size", "4", "--num_events", str(nev), "--num_consumers", "1",
             "--queue_name", "shared_queue", "--ray_namespace", "my", "--reliable_sentinels", "--calib", "--common_mode", "7,4,50,10"]
    env_a, env_b = dict(os.environ, PSANA_RAY_DIR=root_a), dict(os.environ, PSANA_RAY_DIR=root_b)
    benv = dict(WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), PYTHONPATH=PKG + os.pathsep + os.environ.get("PYTHONPATH", ""),
                PSANA_RAY_DRAIN_S="60")
    bargs = [sys.executable, "-m", "psana_ray.bridge", "--queue_name", "shared_queue", "--ray_namespace", "my", "--log_level", "WARNING"]
    procs = [_launch(env_a, 0, 1, extra),
             subprocess.Popen(bargs + ["--role", "send", "--device", "0"], env=dict(env_a, RANK="0", **benv), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True),
             subprocess.Popen(bargs + ["--role", "recv", "--device", "1", "--queue_size", "8", "--batch_size", "4", "--nslabs", "3"],
                              env=dict(env_b, RANK="1", **benv), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)]
    try:
        t0 = time.time()
        reader = DataReader(address=root_b, device=1)
        while True:
            try:
                reader.connect(); break
            except ValueError:
                assert time.time() - t0 < 120 and all(p.poll() is None for p in procs), "a process died: " + "".join(p.stdout.read() for p in procs if p.poll() is not None)
                time.sleep(0.2)
        items = []
        for rank, idx, data, energy in reader:                     # stops at the bridged sentinel
            assert data.is_cuda and data.device.index == 1
            items.append((rank, idx, data.cpu().numpy(), energy))
        reader.close()
        for p in procs:
            assert p.wait(timeout=120) == 0, p.stdout.read()
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
        for root in (root_a, root_b):
            try:
                SQ.get_queue("shared_queue", "my", root).shutdown()
            except Exception:
                pass
    assert [i[1] for i in items] == list(range(nev))
    cmpars = (7, 4, 50.0, 10)
    src = SyntheticSource("mfxl1038923", 58, det, rank=0, size=1, num_events=nev, common_mode=cmpars)
    for rank, idx, data, energy in items:
        assert rank == 0 and energy == src.photon_energy(idx)
        kat.assert_same_bits(data, oc.det_calib(src.spec.kind, src.raw_event(idx), src.constants(), cmpars), f"bridged event {idx}")
