"""Host-side contracts of the reference, checked on CPU: shared_queue (put/get/size, named detached
lookup), DataReader, producer loop.  The reference has no tests; every check
below is derived from the cited reference lines.  The calibrator double is the NumPy oracle, which
is also BASELINE.json configs[0] (the CPU-runnable producer -> queue -> consumer case)."""
import os
import subprocess
import sys
import time

import numpy as np
import pytest

import kat
from oracle import psana_calib as oc
from psana_ray import producer as P
from psana_ray import shared_queue as SQ
from psana_ray.data_reader import DataReader, DataReaderError
from psana_ray.event_source import ImageRetrievalMode, SyntheticSource
from psana_ray.transport import EventBatch

PKG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "psana-ray_b200")


@pytest.fixture()
def qroot(tmp_path, monkeypatch):
    root = str(tmp_path / "q")
    monkeypatch.setenv(SQ.DEFAULT_ROOT_ENV, root)
    made = []
    yield root, made
    for h in made:
        try:
            h.shutdown()
        except Exception:
            pass


def _mk(made, **kw):
    h = SQ.create_queue(**kw)
    assert h is not None
    made.append(h)
    return h


# ---- Queue contract (ref: shared_queue.py:9-31) ------------------------------------------------------

def test_inprocess_queue_contract():
    q = SQ.Queue(maxsize=3)
    assert q.get() is None and q.size() == 0                      # :21, :28
    assert [q.put(i) for i in range(5)] == [True, True, True, False, False]     # :11-14 never evicts
    assert q.size() == 3
    assert [q.get() for _ in range(4)] == [0, 1, 2, None]          # FIFO
    assert q.put(None) and q.size() == 1                          # None is storable (producer.py:125)
    assert q.get_ex() == (True, None) and q.get_ex() == (False, None)


def test_queue_socket_is_private_to_the_user(qroot):
    """Requests are pickles, so connecting means running code as the owner: 0700 directories, 0600 socket."""
    import stat
    root, made = qroot
    _mk(made, queue_name="priv", ray_namespace="nsp", maxsize=2)
    path = SQ.socket_path("priv", "nsp")
    assert stat.S_IMODE(os.stat(path).st_mode) == 0o600
    assert stat.S_IMODE(os.stat(os.path.dirname(path)).st_mode) == 0o700
    assert stat.S_IMODE(os.stat(root).st_mode) == 0o700


def test_weighted_items_keep_maxsize_in_events():
    q = SQ.Queue(maxsize=100)
    assert q.put("a", 64) and not q.put("b", 64) and q.put("c", 36) and not q.put("d", 1)
    assert q.size() == 100
    big = SQ.Queue(maxsize=10)
    assert big.put("batch", 64)                                    # an empty queue always takes one item
    assert not big.put("x", 1)


def test_server_queue_named_namespaced_detached(qroot):
    root, made = qroot
    h = _mk(made, queue_name="my", ray_namespace="default", maxsize=2)
    assert h.put([0, 0, np.arange(4, dtype=np.float32), 9500.0]) is True
    assert h.put("second") is True and h.put("third") is False     # full -> False
    assert h.size() == 2
    # reachable by (name, namespace) from another process; other namespaces do not see it
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from psana_ray.shared_queue import get_queue\n"
            "q = get_queue('my', 'default'); it = q.get(); print(it[0], it[1], it[2].sum(), it[3], q.size())\n"
            "try:\n    get_queue('my', 'other')\nexcept ValueError: print('absent')\n" % PKG)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=dict(os.environ), timeout=60)
    assert out.stdout.split() == ["0", "0", "6.0", "9500.0", "1", "absent"], out.stderr
    assert h.get() == "second" and h.get() is None and h.size() == 0
    with pytest.raises(ValueError):
        SQ.get_queue("nope", "default")
    assert SQ.create_queue(queue_name="my", ray_namespace="default") is None      # name taken -> None (:36-38)


def test_queue_outlives_creator_and_dead_queue_raises(qroot):
    root, made = qroot
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from psana_ray.shared_queue import create_queue\n"
            "q = create_queue('det', 'ns', maxsize=5); q.put('left behind'); print(q.stats()['pid'])\n" % PKG)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=dict(os.environ), timeout=60)
    assert out.returncode == 0, out.stderr
    h = SQ.get_queue("det", "ns")                                  # creator has exited; the queue is still there
    made.append(h)
    assert h.get() == "left behind"
    h2 = SQ.get_queue("det", "ns")
    h2.shutdown()
    time.sleep(0.3)
    with pytest.raises(SQ.QueueDeadError):
        h.size()


def test_leases_release_and_auto_release_on_disconnect(qroot):
    root, made = qroot
    h = _mk(made, queue_name="l", ray_namespace="n", maxsize=10)
    for i in range(3):
        h.put(f"b{i}", weight=2, lease=("own", i, 1))
    c1, c2 = SQ.get_queue("l", "n"), SQ.get_queue("l", "n")
    assert c1.get() == "b0" and c2.get() == "b1" and c2.get() == "b2"
    assert h.poll_released("own") == []
    c1.release(("own", 0, 1))
    assert h.poll_released("own") == [("own", 0, 1)]
    c2.close()                                                    # crashed consumer: its two leases come back
    time.sleep(0.2)
    assert sorted(h.poll_released("own")) == [("own", 1, 1), ("own", 2, 1)]
    c1.close()


# ---- DataReader (ref: data_reader.py:5-48) -----------------------------------------------------------------

def test_datareader_defaults_and_errors(qroot):
    root, made = qroot
    r = DataReader()
    assert (r.address, r.queue_name, r.ray_namespace) == ("auto", "shared_queue", "my")       # :5
    with pytest.raises(RuntimeError):                             # :32-33
        r.read()
    with pytest.raises(ValueError):                               # connect to a queue that is not there re-raises (:19-24)
        r.connect()
    h = _mk(made, queue_name="shared_queue", ray_namespace="my", maxsize=8)
    with DataReader() as rd:                                      # :39-44
        assert rd.read() is None                                  # empty -> None (:21 via :35)
        h.put([3, 7, "payload", None])                            # a plain per-event item passes through untouched
        assert rd.read() == [3, 7, "payload", None]
        h2 = SQ.get_queue("shared_queue", "my"); h2.shutdown(); time.sleep(0.3)
        with pytest.raises(DataReaderError):                      # :36-37
            rd.read()
    assert rd._queue is None


def test_datareader_debatches_in_order_and_releases(qroot):
    root, made = qroot
    h = _mk(made, queue_name="shared_queue", ray_namespace="my", maxsize=100)
    a = np.arange(3 * 2 * 4, dtype=np.float32).reshape(3, 2, 4)
    b = -np.arange(2 * 2 * 4, dtype=np.float32).reshape(2, 2, 4)
    h.put(EventBatch(rank=1, idx0=10, count=3, data=a, photon_energy=[1.0, 2.0, 3.0], lease=("o", 0, 1)), weight=3, lease=("o", 0, 1))
    h.put(EventBatch(rank=0, idx0=0, count=2, data=b, photon_energy=[None, 5.0], lease=("o", 1, 1)), weight=2, lease=("o", 1, 1))
    h.put(None)
    assert h.size() == 6
    with DataReader() as rd:
        got = [rd.read() for _ in range(3)]
        assert [(g[0], g[1], g[3]) for g in got] == [(1, 10, 1.0), (1, 11, 2.0), (1, 12, 3.0)]
        assert all(np.array_equal(g[2], a[i]) for i, g in enumerate(got))
        assert h.poll_released("o") == []                         # still holding batch 0
        g = rd.read()
        assert (g[0], g[1], g[3]) == (0, 0, None) and np.array_equal(g[2], b[0])
        assert h.poll_released("o") == [("o", 0, 1)]              # moved on -> slab 0 released
        rest = rd.read_batch()
        assert rest[1] == 1 and np.array_equal(rest[2], b[1:]) and rest[3] == [5.0]
        assert rd.read() is None and rd.end_of_stream             # sentinel
        assert rd.read() is None and h.size() == 0                # empty: also None (indistinguishable, as in the reference)
    assert h.poll_released("o") == [("o", 1, 1)]


def test_read_batch_walks_batches_and_releases(qroot):
    root, made = qroot
    h = _mk(made, queue_name="shared_queue", ray_namespace="my", maxsize=100)
    for k in range(3):
        h.put(EventBatch(0, 4 * k, 4, np.full((4, 1, 2, 2), k, np.float32), [float(k)] * 4, lease=("o", k, 1)), weight=4, lease=("o", k, 1))
    h.put(None)
    with DataReader() as rd:
        got = []
        while True:
            b = rd.read_batch()
            if b is None:
                assert rd.end_of_stream
                break
            got.append((b[1], b[2].shape[0], float(b[2].mean())))
        assert got == [(0, 4, 0.0), (4, 4, 1.0), (8, 4, 2.0)]
        assert sorted(h.poll_released("o")) == [("o", 0, 1), ("o", 1, 1), ("o", 2, 1)]
        assert rd.read_batch() is None


def test_datareader_iter_stops_at_sentinel(qroot):
    root, made = qroot
    h = _mk(made, queue_name="shared_queue", ray_namespace="my", maxsize=100)
    h.put(EventBatch(0, 0, 2, np.zeros((2, 1, 3, 3), np.float32), [1.0, 2.0]), weight=2)
    h.put(None)
    h.put("after the end")
    with DataReader() as rd:
        items = list(rd)
    assert [i[1] for i in items] == [0, 1] and items[0][2].shape == (1, 3, 3)


# ---- producer (ref: producer.py) ---------------------------------------------------------------------------

def test_flags_and_defaults_are_the_reference_ones():
    a = P.parse_arguments(["--exp", "mfxl1038923", "--run", "58", "--detector_name", "epix10k2M"])
    assert (a.exp, a.run, a.detector_name) == ("mfxl1038923", 58, "epix10k2M")
    assert a.calib is False and a.uses_bad_pixel_mask is False and a.manual_mask_path is None
    assert (a.ray_address, a.ray_namespace, a.queue_name, a.queue_size) == ("auto", "default", "my", 100)    # :24-28
    assert (a.num_consumers, a.max_steps, a.log_level) == (1, None, "INFO")
    with pytest.raises(SystemExit):
        P.parse_arguments(["--run", "1", "--detector_name", "x"])            # --exp required
    with pytest.raises(SystemExit):
        P.parse_arguments(["--exp", "e", "--run", "1", "--detector_name", "x", "--log_level", "TRACE"])


def test_rank_and_size_from_launcher_environment():
    assert (P.get_world({}).rank, P.get_world({}).size) == (0, 1)
    w = P.get_world({"OMPI_COMM_WORLD_RANK": "3", "OMPI_COMM_WORLD_SIZE": "4", "OMPI_COMM_WORLD_LOCAL_RANK": "1"})
    assert (w.rank, w.size, w.local_rank) == (3, 4, 1)
    w = P.get_world({"SLURM_PROCID": "5", "SLURM_NTASKS": "8", "SLURM_LOCALID": "5"})
    assert (w.rank, w.size, w.local_rank) == (5, 8, 5)
    w = P.get_world({"RANK": "1", "WORLD_SIZE": "2"})
    assert (w.rank, w.size, w.local_rank) == (1, 2, 1)


def test_keep_mask_is_bad_then_manual_and_broadcasts_like_np_where():
    shape = (2, 3, 4)
    bad = (np.arange(24).reshape(shape) % 2).astype(np.uint8)
    man = np.arange(24).reshape(shape) % 3 > 0
    keep = P.combine_keep_masks(shape, bad, man)
    d = np.random.default_rng(0).standard_normal(shape).astype(np.float32)
    kat.assert_same_bits(np.where(keep, d, 0), oc.producer_masks(d, bad, man))
    assert P.combine_keep_masks(shape, None, None) is None
    with pytest.raises(ValueError):                               # a calib-shaped mask against a 2-D image
        P.combine_keep_masks((5, 6), bad, None)


class OraclePipeline:
    """Calibrator double: the NumPy oracle on the CPU (tests only)."""

    def __init__(self, source, mode, batch_size, bad=None, manual=None):
        self.src, self.mode, self.batch_size, self.bad, self.manual = source, mode, batch_size, bad, manual
        self.buf = np.zeros((batch_size,) + source.spec.shape, dtype=np.uint16)
        self.calls = []

    def stage(self, k, raw):
        self.buf[k] = raw

    def one(self, raw):
        c = self.src.constants()
        if self.mode == ImageRetrievalMode.image:
            d = oc.det_image(self.src.spec.kind, raw, c, *self.src.geometry(), cmpars=self.src.common_mode())
        else:
            d = oc.det_calib(self.src.spec.kind, raw, c, self.src.common_mode())
        return oc.producer_masks(d, self.bad, self.manual)

    def process(self, n, queue):
        self.calls.append(n)
        return np.stack([self.one(self.buf[k]) for k in range(n)]), None


def _drain(reader):
    out = []
    while True:
        it = reader.read()
        if it is None:
            return out
        out.append(it)


@pytest.mark.parametrize("mode", [ImageRetrievalMode.calib, ImageRetrievalMode.image])
def test_config0_producer_queue_consumer_on_cpu(qroot, mode):
    """BASELINE.json configs[0]: synthetic epix10ka frames, 1 rank, queue_size 400, NumPy calib."""
    root, made = qroot
    h = _mk(made, queue_name="my", ray_namespace="default", maxsize=400)
    src = SyntheticSource("exp", 58, "epix10k2M", rank=0, size=1, num_events=11, nseg=1)
    bad = src.create_bad_pixel_mask() if mode == ImageRetrievalMode.calib else None
    pipe = OraclePipeline(src, mode, batch_size=4, bad=bad)
    P.produce_data(src, pipe, h, P.World(0, 1, 0), num_consumers=2)
    assert pipe.calls == [4, 4, 3] and h.size() == 11 + 2          # ragged tail batch; 2 sentinels
    with DataReader(queue_name="my", ray_namespace="default") as rd:
        items = _drain(rd)
        assert rd.end_of_stream
    assert [(i[0], i[1]) for i in items] == [(0, k) for k in range(11)]              # [rank, idx, data, photon_energy]
    for k, it in enumerate(items):
        assert it[3] == src.photon_energy(k)
        exp = pipe.one(src.raw_event(k))
        assert it[2].shape == exp.shape and (it[2].ndim == 3)                       # (segs,rows,cols) or (1,H,W)
        kat.assert_same_bits(it[2], exp)


def test_event_sharding_index_and_max_steps(qroot):
    root, made = qroot
    h = _mk(made, queue_name="my", ray_namespace="default", maxsize=400)
    for rank in (0, 1, 2):
        src = SyntheticSource("exp", 1, "epix10k2M", rank=rank, size=3, num_events=20, nseg=1)
        assert src.event_numbers() == list(range(rank, 20, 3))                       # nevent % size == rank
        pipe = OraclePipeline(src, ImageRetrievalMode.calib, batch_size=8)
        # size=1 world per call so that no barrier is needed in-process; rank identity still reaches the items
        P.produce_data(src, pipe, h, P.World(rank, 1, rank), num_consumers=0, max_steps=5)
        assert sum(pipe.calls) == 5                                                  # ref :89-91
    with DataReader(queue_name="my", ray_namespace="default") as rd:
        items = _drain(rd)
    assert [(i[0], i[1]) for i in items] == [(r, k) for r in range(3) for k in range(5)]   # idx is rank-local (ref :88)
    src1 = SyntheticSource("exp", 1, "epix10k2M", rank=1, size=3, num_events=20, nseg=1)
    kat.assert_same_bits(items[5 + 2][2], oc.det_calib("epix10ka", src1.raw_event(1 + 3 * 2), src1.constants()))


def test_backoff_schedule_and_dropped_sentinels(qroot, monkeypatch):
    root, made = qroot
    h = _mk(made, queue_name="my", ray_namespace="default", maxsize=2)
    sleeps = []
    monkeypatch.setattr(P.random, "uniform", lambda a, b: 0.25)
    state = {"n": 0}

    def fake_sleep(s):
        sleeps.append(s)
        state["n"] += 1
        if state["n"] == 7:                                        # a consumer finally makes room
            assert h.get() is not None
    monkeypatch.setattr(P.time, "sleep", fake_sleep)
    h.put("x"); h.put("y")                                         # queue full
    assert P._backoff_put(h, "item", 1, None, 0, "t") is True
    assert sleeps == pytest.approx([0.1 + .25, 0.2 + .25, 0.4 + .25, 0.8 + .25, 1.6 + .25, 2.0 + .25, 2.0 + .25])   # ref :85-86,:108-111
    # sentinels: put result is not checked -> silently dropped when the queue is full (ref :125 vs shared_queue :11-14)
    src = SyntheticSource("exp", 1, "epix10k2M", num_events=0, nseg=1)
    P.produce_data(src, OraclePipeline(src, ImageRetrievalMode.calib, 4), h, P.World(0, 1, 0), num_consumers=3)
    assert h.size() == 2 and h.stats()["refused"] >= 3
    # additive --reliable_sentinels: sentinels wait for room like events do
    state["n"] = -100
    drained = []

    def fake_sleep2(s):
        drained.append(h.get())
    monkeypatch.setattr(P.time, "sleep", fake_sleep2)
    P.produce_data(src, OraclePipeline(src, ImageRetrievalMode.calib, 4), h, P.World(0, 1, 0), num_consumers=3, reliable_sentinels=True)
    rest = [h.get() for _ in range(h.size())]
    assert (drained + rest).count(None) == 3


def test_dead_queue_makes_the_producer_return(qroot):
    root, made = qroot
    h = SQ.create_queue(queue_name="my", ray_namespace="default", maxsize=4)
    SQ.get_queue("my", "default").shutdown()
    time.sleep(0.3)
    src = SyntheticSource("exp", 1, "epix10k2M", num_events=6, nseg=1)
    pipe = OraclePipeline(src, ImageRetrievalMode.calib, 2)
    assert P.produce_data(src, pipe, h, P.World(0, 1, 0)) is None                    # ref :112-114
    assert pipe.calls == [2]                                                        # stopped at the first put


def test_initialize_queue_get_or_create_and_failure(qroot):
    root, made = qroot
    q = P.initialize_queue("auto", "default", "my", 7, P.World(0, 1, 0))
    assert q is not None and q.stats()["maxsize"] == 7
    made.append(q)
    q2 = P.initialize_queue("auto", "default", "my", 99, P.World(0, 1, 0))           # existing queue is reused (:43-45)
    assert q2.stats()["maxsize"] == 7 and q2.stats()["pid"] == q.stats()["pid"]
    # a non-zero rank that never finds the queue gives up after max_retries and returns None (:56-71)
    assert P.initialize_queue("auto", "nowhere", "my", 7, P.World(1, 2, 1), max_retries=2, retry_delay=0.01) is None


def test_batch_source_fast_path_same_items(qroot):
    """--batch_source: whole batches served from a pool, no per-event host work; items keep the reference layout."""
    root, made = qroot
    h = _mk(made, queue_name="my", ray_namespace="default", maxsize=400)
    src = SyntheticSource("exp", 3, "epix10k2M", rank=1, size=2, num_events=21, nseg=1)      # 10 local events

    class ViewPipeline(OraclePipeline):
        def process_view(self, raw, queue):
            self.calls.append(raw.shape[0])
            return np.stack([self.one(raw[k]) for k in range(raw.shape[0])]), None

    pipe = ViewPipeline(src, ImageRetrievalMode.calib, batch_size=4)
    P.produce_data(src, pipe, h, P.World(1, 1, 1), num_consumers=1, batch_source=True, max_steps=9)
    assert sum(pipe.calls) == 9 and max(pipe.calls) <= 4
    with DataReader(queue_name="my", ray_namespace="default") as rd:
        items = _drain(rd)
    assert [(i[0], i[1]) for i in items] == [(1, k) for k in range(9)]
    for k, it in enumerate(items):
        assert it[3] == src.photon_energy(1 + 2 * k)
        kat.assert_same_bits(it[2], pipe.one(src.batch_pool_frame(k)))


def test_iterable_dataset_yields_whole_batches(qroot):
    import torch
    from torch.utils.data import DataLoader
    from psana_ray.dataset import PsanaRayDataset
    root, made = qroot
    h = _mk(made, queue_name="shared_queue", ray_namespace="my", maxsize=100)
    for k in range(3):
        h.put(EventBatch(k % 2, 5 * k, 5, np.full((5, 2, 3, 3), k, np.float32), [1.0] * 5, lease=("o", k, 1)), weight=5, lease=("o", k, 1))
    h.put(None)
    seen = []
    for data, rank, idx0, energies in DataLoader(PsanaRayDataset(), batch_size=None):
        assert isinstance(data, torch.Tensor) and tuple(data.shape) == (5, 2, 3, 3) and len(energies) == 5
        seen.append((int(rank), int(idx0), float(data.mean())))
    assert seen == [(0, 0, 0.0), (1, 5, 1.0), (0, 10, 2.0)]
    assert len(h.poll_released("o")) == 3


def test_epix10ka_config_bytes_broadcasts_trbit_per_asic():
    from psana_ray.event_source import epix10ka_config_bytes
    from psana_ray.synthetic import EPIX10KA_MODE_LUT
    pca = np.zeros((352, 384), dtype=np.uint16)
    pca[:, :] = 0xFFF3                      # junk everywhere except bits 3,2 = 00 (auto-ranging)
    pca[10, 20] = 0x000C                    # one fixed-high pixel
    cfg = epix10ka_config_bytes(pca, trbits=[1, 0, 0, 1])
    assert cfg.dtype == np.uint8 and cfg.shape == (352, 384)
    assert cfg[300, 300] == 0x10 and cfg[300, 10] == 0x10          # A0 bottom right, A3 bottom left
    assert cfg[10, 300] == 0x00 and cfg[10, 20] == 0x0C            # A1 top right (trbit 0), pixel config kept
    # decode: trbit 1 + auto -> AHL-H (3) / AHL-L (5); trbit 0 + auto -> AML-M (4) / AML-L (6); fixed high, trbit 0 -> FM (1)
    assert EPIX10KA_MODE_LUT[cfg[300, 300]] == 3 and EPIX10KA_MODE_LUT[cfg[300, 300] | 32] == 5
    assert EPIX10KA_MODE_LUT[cfg[10, 300]] == 4 and EPIX10KA_MODE_LUT[cfg[10, 300] | 32] == 6
    assert EPIX10KA_MODE_LUT[cfg[10, 20]] == 1
