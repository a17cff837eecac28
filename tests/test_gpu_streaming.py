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


def _wait_reader(procs, **kw):
    t0 = time.time()
    while True:
        try:
            rd = DataReader(**kw); rd.connect(); return rd
        except ValueError:
            assert time.time() - t0 < 120 and all(p.poll() is None for p in procs), "producer died: " + "".join(p.stdout.read() for p in procs if p.poll() is not None)
            time.sleep(0.2)


def _finish(procs):
    for p in procs:
        if p.poll() is None:
            p.kill()
    try:
        SQ.get_queue("shared_queue", "my").shutdown()
    except Exception:
        pass


def test_items_kept_by_the_consumer_stay_valid_after_their_slabs_are_recycled(tmp_path, monkeypatch):
    """The reference hands the consumer an ndarray it owns (ref: data_reader.py:35).  So does read() by default: a
    consumer that keeps every item on the GPU while the producer's small slab ring is overwritten many times over still
    holds the right frames at the end."""
    monkeypatch.setenv(SQ.DEFAULT_ROOT_ENV, str(tmp_path / "q"))
    nev, det, cm = 40, "epix10k2M", (7, 4, 50.0, 10)
    extra = ["--detector_name", det, "--queue_size", "8", "--batch_size", "4", "--num_events", str(nev), "--num_consumers", "1",
             "--queue_name", "shared_queue", "--ray_namespace", "my", "--reliable_sentinels", "--calib", "--common_mode", "7,4,50,10"]
    procs = [_launch(os.environ, 0, 1, extra)]                     # queue_size 8 / batch 4 -> a ring of 4 slabs for 10 batches
    try:
        reader = _wait_reader(procs)
        kept = [it for it in reader]                               # nothing copied to the host, nothing looked at, until the end
        assert reader.end_of_stream
        reader.close()
        for p in procs:
            assert p.wait(timeout=120) == 0, p.stdout.read()
    finally:
        _finish(procs)
    assert len(kept) == nev and all(it[2].is_cuda for it in kept)
    src = SyntheticSource("mfxl1038923", 58, det, rank=0, size=1, num_events=nev, common_mode=cm)
    for rank, idx, data, energy in kept:
        kat.assert_same_bits(data.cpu().numpy(), oc.det_calib(src.spec.kind, src.raw_event(idx), src.constants(), cm), f"kept event {idx}")


def test_dataset_batches_over_cuda_ipc_and_lease_returns_with_the_next_batch(tmp_path, monkeypatch):
    """SURVEY 8f n2 ("Consumers -> Batches -> PyTorch task", ref: figures/psana-ray-architecture.png,
    examples/psana_consumer.py:28-47): PsanaRayDataset over real CUDA-IPC slabs, zero-copy.  While a step works on
    batch k exactly one slab is out; asking for batch k + 1 hands batch k's slab back to its producer."""
    import torch
    from psana_ray.dataset import PsanaRayDataset
    monkeypatch.setenv(SQ.DEFAULT_ROOT_ENV, str(tmp_path / "q"))
    nev, det = 24, "epix10k2M"
    extra = ["--detector_name", det, "--queue_size", "8", "--batch_size", "4", "--num_events", str(nev), "--num_consumers", "1",
             "--queue_name", "shared_queue", "--ray_namespace", "my", "--reliable_sentinels", "--calib"]
    procs = [_launch(os.environ, 0, 1, extra)]
    got, leases_seen = [], []
    try:
        ds = PsanaRayDataset(zero_copy=True, connect_timeout_s=120.0)
        watch = None
        for data, rank, idx0, energies in torch.utils.data.DataLoader(ds, batch_size=None, num_workers=0):
            if watch is None:
                watch = SQ.get_queue("shared_queue", "my")
            assert data.is_cuda and data.dim() == 4 and data.shape[1:] == (16, 352, 384)
            leases_seen.append(watch.stats()["leases_out"])        # leases held by consumer connections right now
            got.append((rank, idx0, data.cpu().numpy(), list(energies)))   # the step: done with the view before the next batch
        watch.close()
        for p in procs:
            assert p.wait(timeout=120) == 0, p.stdout.read()
    finally:
        _finish(procs)
    assert sum(g[2].shape[0] for g in got) == nev
    assert leases_seen and max(leases_seen) == 1                   # never two slabs out: batch k went back when k + 1 was requested
    src = SyntheticSource("mfxl1038923", 58, det, rank=0, size=1, num_events=nev)
    for rank, idx0, data, energies in got:
        for k in range(data.shape[0]):
            assert energies[k] == src.photon_energy(idx0 + k)
            kat.assert_same_bits(data[k], oc.det_calib(src.spec.kind, src.raw_event(idx0 + k), src.constants(), None), f"event {idx0 + k}")


def test_host_fed_call_begun_twice_before_waiting_gives_both_batches():
    """psr_calib_batch_host_begin: two batches queued back to back (the second begins while the first is in flight and
    shares its staging ring), each with its own completion event; results bit-exact."""
    import torch
    from psana_ray import _native as N, synthetic as syn
    from psana_ray.calib import GpuCalibrator
    lib = N.load()
    spec = syn.EPIX10K2M.with_nseg(4)
    consts = syn.make_constants(spec, seed=5)
    raw = syn.make_raw_frames(spec, consts, 22, seed=6)
    cal = GpuCalibrator(spec, device=0)
    cal.set_constants(consts["peds"], consts["gain"], None, consts["cfg"], consts["mask"])
    cal.set_common_mode((7, 4, 50.0, 10))
    h = torch.from_numpy(raw.view(np.int16)).pin_memory()
    outs = [torch.full((11,) + spec.shape, float("nan"), dtype=torch.float32, device="cuda") for _ in range(2)]
    evs = []
    for k in range(2):
        ev, hb = C.c_void_p(), (C.c_ubyte * N.PSR_IPC_HANDLE_BYTES)()
        N.check(lib.psr_ipc_event_create(C.byref(ev), hb, 0))
        evs.append(ev)
        cal.calib_host_begin(h[11 * k:11 * (k + 1)], d_out=outs[k], done_event=ev.value)
    cal.calib_host_wait(input_only=True)                           # the host buffer may be reused from here
    st = torch.cuda.current_stream().cuda_stream
    for ev in evs:
        N.check(lib.psr_event_wait(ev, st))                        # what a consumer does before touching the batch
    got = [o.clone() for o in outs]
    torch.cuda.synchronize()
    for k in range(2):
        for f in (0, 10):
            kat.assert_same_bits(got[k][f].cpu().numpy(), oc.det_calib(spec.kind, raw[11 * k + f], consts, (7, 4, 50.0, 10)), f"batch {k} frame {f}")
    cal.calib_host_wait()
    for ev in evs:
        N.check(lib.psr_event_destroy(ev))
    cal.close()
