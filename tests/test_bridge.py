"""Cross-box bridge (psana_ray.bridge) on CPU: two queue roots stand for two boxes, a world_size-2 gloo pair
carries the batches.  What arrives in queue B must be what was put into queue A -- same (rank, idx) keys, same
frames bit for bit, same order, loose per-event items passed through as plain items, the end-of-stream sentinel
delivered to every consumer of box B -- and the sender must give every lease of box A back."""
import os
import socket
import subprocess
import sys
import time

import numpy as np

from psana_ray import shared_queue as SQ
from psana_ray.data_reader import DataReader
from psana_ray.transport import EventBatch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "psana-ray_b200")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _bridge(role, rank, root, port, extra=()):
    env = dict(os.environ, PSANA_RAY_DIR=root, RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
               PYTHONPATH=PKG + os.pathsep + os.environ.get("PYTHONPATH", ""), CUDA_VISIBLE_DEVICES="")
    return subprocess.Popen([sys.executable, "-m", "psana_ray.bridge", "--role", role, "--backend", "gloo", "--queue_name", "my",
                             "--ray_namespace", "ns", "--log_level", "WARNING", *extra], env=env, stderr=subprocess.PIPE, text=True)


def test_bridge_forwards_batches_loose_items_and_sentinels(tmp_path):
    root_a, root_b = str(tmp_path / "boxA"), str(tmp_path / "boxB")
    rng = np.random.default_rng(5)
    qa = SQ.create_queue("my", "ns", maxsize=1000, address=root_a)
    assert qa is not None
    batches = []
    for b, n in enumerate((4, 1, 3)):                       # ragged batch sizes
        data = rng.standard_normal((n, 2, 5, 7)).astype(np.float32)
        batches.append(EventBatch(rank=b % 2, idx0=10 * b, count=n, data=data, photon_energy=[9500.0 + i for i in range(n)],
                                  lease=("ownerA", b, 1)))
        assert qa.put(batches[-1], weight=n, lease=batches[-1].lease)
    loose = [1, 77, rng.standard_normal((2, 5, 7)).astype(np.float32), None]
    assert qa.put(loose)
    assert qa.put(None)                                     # end of stream for the sender (one consumer of box A)
    port = _free_port()
    tx = _bridge("send", 0, root_a, port)
    rx = _bridge("recv", 1, root_b, port, ("--queue_size", "5", "--num_consumers", "2"))   # small queue: the back-off path runs
    # consume on box B while the bridge is still running (its queue holds 5 events at a time)
    deadline = time.time() + 120
    rd = DataReader(address=root_b, queue_name="my", ray_namespace="ns")
    while True:
        try:
            rd.connect()
            break
        except ValueError:
            assert time.time() < deadline, "box B's queue never appeared"
            time.sleep(0.05)
    got = list(rd)                                          # stops at the first sentinel
    assert tx.wait(timeout=120) == 0, tx.stderr.read()
    assert rx.wait(timeout=120) == 0, rx.stderr.read()
    want = [(b.rank, b.idx0 + i, b.data[i], b.photon_energy[i]) for b in batches for i in range(b.count)]
    want.append(tuple(loose))
    assert len(got) == len(want)
    for (r, i, d, e), (wr, wi, wd, we) in zip(got, want):
        assert (r, i, e) == (wr, wi, we)
        assert np.asarray(d).dtype == np.float32 and np.array_equal(np.asarray(d), wd)
    qb = SQ.get_queue("my", "ns", root_b)
    assert qb.get_ex() == (True, None) and qb.get_ex() == (False, None)      # the second consumer's sentinel, then empty
    rd.close()
    # every lease of box A came back to its owner once the batch was on the wire
    assert sorted(tuple(l) for l in qa.poll_released("ownerA")) == [("ownerA", 0, 1), ("ownerA", 1, 1), ("ownerA", 2, 1)]
    qa.shutdown(); qb.shutdown()
