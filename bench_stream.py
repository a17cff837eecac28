#!/usr/bin/env python
"""bench_stream.py -- BASELINE.json configs[4]: end-to-end producer -> queue -> multi-consumer streaming
with GPU-resident CUDA-IPC batch hand-off.

    python bench_stream.py --gpus N [--consumers M] [--events-per-rank E] [--batch 64] [--detector epix10k2M]
    python bench_stream.py --impl reference [--events-per-rank E]     # configs[0]: the reference's CPU flow, 1 rank, queue_size 400

Launches N producer ranks (the real CLI, `python -m psana_ray.producer --batch_source`, one per GPU) and M
consumer processes (default one per GPU; consumer c maps the producers' slabs on GPU c %% N, so a consumer
on another GPU reads over NVLink).  Every consumer touches every frame it receives (a sum on its GPU) so that
hand-off without a read does not count.  Prints one JSON line: whole-job calibrated frames/s from the first
event a consumer receives to the last.
"""
import argparse, json, os, subprocess, sys, tempfile, time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "psana-ray_b200")

CONSUMER = r"""
import json, os, sys, time
sys.path.insert(0, %(pkg)r)
import torch
from psana_ray.data_reader import DataReader
dev = int(sys.argv[1]); torch.cuda.set_device(dev)
t_deadline = time.time() + 120
while True:
    try:
        rd = DataReader(queue_name="stream", ray_namespace="bench", device=dev, zero_copy=True, expected_producers=int(sys.argv[2]) if len(sys.argv) > 2 else 0); rd.connect(); break
    except ValueError:
        if time.time() > t_deadline: raise
        time.sleep(0.05)
n = 0; fb = 0; marks = []; acc = torch.zeros((), dtype=torch.float64, device="cuda"); t_first = None
while True:
    b = rd.read_batch()
    if b is None:
        if rd.end_of_stream: break
        time.sleep(0.0005); continue
    if t_first is None: t_first = time.time()
    acc += b[2].sum(dtype=torch.float64)          # the consumer really reads the frames
    n += b[2].shape[0]; fb = b[2][0].numel() * 4
    marks.append((time.time(), n))
torch.cuda.synchronize(); t_last = time.time()
rd.close()
print(json.dumps({"n": n, "t_first": t_first, "t_last": t_last, "sum": float(acc), "device": dev, "frame_bytes": fb, "marks": marks[::4]}))
"""


# ---- reference arm: BASELINE.json configs[0], the reference's own per-event flow on the host cores -------------------------
# One producer rank calibrates each event with the NumPy oracle (what `det.calib` does inside psana_wrapper's iter_events),
# applies the two np.where masks (ref: producer.py:92-95), and puts the pickled [rank, idx, ndarray, photon_energy] item with the
# reference's back-off (ref: producer.py:99-117); a consumer polls DataReader.read() (ref: examples/psana_consumer.py:28-47).
# The queue server of this repo stands in for the detached Ray actor (Ray is not installable here).
CPU_PRODUCER = r"""
import sys, time
sys.path.insert(0, %(pkg)r); sys.path.insert(0, %(root)r)
import numpy as np
from oracle import psana_calib as O                      # bench reference arm: the one place outside tests that runs the oracle
from psana_ray import synthetic as S
from psana_ray.shared_queue import create_queue, get_queue
from psana_ray.producer import _backoff_put
det, nev, qsize, ncons, pool = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), 16
spec = S.detector_spec(det)
consts = S.make_constants(spec, seed=0)
consts["gfac"] = O.divide_protected(np.ones_like(consts["gain"]), consts["gain"])
raw = S.make_raw_frames(spec, consts, pool, seed=1)
bad = consts["mask"]; consts = dict(consts, mask=None)   # det.calib without mask; the producer applies it (--uses_bad_pixel_mask)
create_queue("stream", "bench", maxsize=qsize)
q = get_queue("stream", "bench")
for idx in range(nev):
    data = O.det_calib(spec.kind, raw[idx %% pool], consts)
    data = O.producer_masks(data, bad_pixel_mask=bad)
    if not _backoff_put(q, [0, idx, data, 9500.0], 1, None, 0, f"idx={idx}"):
        sys.exit(2)
for _ in range(ncons):
    while not q.put(None):
        time.sleep(0.01)
"""

CPU_CONSUMER = r"""
import json, sys, time
sys.path.insert(0, %(pkg)r)
from psana_ray.data_reader import DataReader
t_deadline = time.time() + 120
while True:
    try:
        rd = DataReader(queue_name="stream", ray_namespace="bench"); rd.connect(); break
    except ValueError:
        if time.time() > t_deadline: raise
        time.sleep(0.05)
n = 0; acc = 0.0; t_first = None
for rank, idx, data, photon_energy in rd:
    if t_first is None: t_first = time.time()
    acc += float(data.sum(dtype="float64")); n += 1
t_last = time.time()
rd.close()
print(json.dumps({"n": n, "t_first": t_first, "t_last": t_last, "sum": acc}))
"""


def reference_arm(a):
    root = tempfile.mkdtemp(prefix="psr_stream_ref_")
    env = dict(os.environ, PSANA_RAY_DIR=root, CUDA_VISIBLE_DEVICES="")
    ncons = a.consumers or 1
    srcs = {}
    for name, text in (("producer", CPU_PRODUCER), ("consumer", CPU_CONSUMER)):
        srcs[name] = os.path.join(root, name + ".py")
        with open(srcs[name], "w") as fh:
            fh.write(text % {"pkg": PKG, "root": ROOT})
    nev = a.events_per_rank
    t_launch = time.time()
    prod = subprocess.Popen([sys.executable, srcs["producer"], a.detector, str(nev), str(a.queue_size), str(ncons)], env=env)
    cons = [subprocess.Popen([sys.executable, srcs["consumer"]], env=env, stdout=subprocess.PIPE, text=True) for _ in range(ncons)]
    outs = [json.loads(c.communicate(timeout=900)[0].strip().splitlines()[-1]) for c in cons]
    rc = prod.wait(timeout=60)
    try:
        sys.path.insert(0, PKG)
        from psana_ray.shared_queue import get_queue
        q = get_queue("stream", "bench", root); stats = q.stats(); q.shutdown()
    except Exception:
        stats = None
    total = sum(o["n"] for o in outs)
    t0 = min(o["t_first"] for o in outs if o["t_first"]); t1 = max(o["t_last"] for o in outs)
    line = {"impl": "reference", "metric": "streamed calibrated frames/sec, producer -> queue -> consumer on the host", "value": total / (t1 - t0),
            "unit": "frames/s", "n_gpus": 0, "producers": 1, "consumers": ncons, "events": total, "expected_events": nev, "seconds": t1 - t0,
            "seconds_with_startup": t1 - t_launch, "producer_exit_codes": [rc], "queue_stats": stats,
            "config": {"workload": "configs[0]: synthetic epix10k2M through producer -> shared_queue -> DataReader on CPU, 1 rank, NumPy calib",
                       "detector": a.detector, "queue_size": a.queue_size, "queue": "this repo's unix-socket queue server standing in for the Ray actor",
                       "cores": 1 + ncons}}
    print(json.dumps(line), flush=True)
    return 0 if total == nev and rc == 0 else 1


def bridge_arm(a):
    """the cross-box leg, with two queue roots on one machine standing for two boxes."""
    import socket
    if a.gpus < 2:
        print("--bridge needs --gpus 2 (NCCL refuses two ranks on one device)", file=sys.stderr)
        return 2
    ncons = a.consumers or 1
    root_a, root_b = tempfile.mkdtemp(prefix="psr_boxA_"), tempfile.mkdtemp(prefix="psr_boxB_")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    base = dict(os.environ, PYTHONPATH=PKG + os.pathsep + os.environ.get("PYTHONPATH", ""), PSANA_RAY_DRAIN_S="120")
    env_a, env_b = dict(base, PSANA_RAY_DIR=root_a), dict(base, PSANA_RAY_DIR=root_b)
    cons_src = os.path.join(root_b, "consumer.py")
    with open(cons_src, "w") as fh:
        fh.write(CONSUMER % {"pkg": PKG})
    nev = a.events_per_rank
    pargs = ["--exp", "bench", "--run", "1", "--detector_name", a.detector, "--queue_name", "stream", "--ray_namespace", "bench",
             "--queue_size", str(a.queue_size), "--batch_size", str(a.batch), "--num_events", str(nev), "--num_consumers", "1",
             "--reliable_sentinels", "--batch_source", "--log_level", "WARNING"] + ([] if a.image else ["--calib"])
    if a.common_mode:
        pargs += ["--common_mode", a.common_mode]
    bargs = ["--queue_name", "stream", "--ray_namespace", "bench", "--log_level", "WARNING"]
    denv = dict(WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    prod = subprocess.Popen([sys.executable, "-m", "psana_ray.producer"] + pargs, env=dict(env_a, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0"))
    tx = subprocess.Popen([sys.executable, "-m", "psana_ray.bridge", "--role", "send", "--device", "0"] + bargs, env=dict(env_a, RANK="0", **denv))
    rx = subprocess.Popen([sys.executable, "-m", "psana_ray.bridge", "--role", "recv", "--device", "1", "--queue_size", str(a.queue_size),
                           "--batch_size", str(a.batch), "--num_consumers", str(ncons)] + bargs, env=dict(env_b, RANK="1", **denv))
    cons = [subprocess.Popen([sys.executable, cons_src, "1"], env=env_b, stdout=subprocess.PIPE, text=True) for _ in range(ncons)]
    outs = [json.loads(c.communicate(timeout=240)[0].strip().splitlines()[-1]) for c in cons]
    rcs = [p.wait(timeout=300) for p in (prod, tx, rx)]
    sys.path.insert(0, PKG)
    from psana_ray.shared_queue import get_queue
    stats = {}
    for name, root in (("A", root_a), ("B", root_b)):
        try:
            q = get_queue("stream", "bench", root); stats[name] = q.stats(); q.shutdown()
        except Exception:
            stats[name] = None
    total = sum(o["n"] for o in outs)
    t0 = min(o["t_first"] for o in outs if o["t_first"]); t1 = max(o["t_last"] for o in outs)
    line = {"metric": "streamed calibrated frames/sec, producer -> queue A -> NCCL bridge -> queue B -> consumer", "value": total / (t1 - t0),
            "unit": "frames/s", "n_gpus": 2, "producers": 1, "consumers": ncons, "events": total, "expected_events": nev, "seconds": t1 - t0,
            "exit_codes": {"producer": rcs[0], "bridge_send": rcs[1], "bridge_recv": rcs[2]}, "queue_stats": stats,
            "payload_GBps": total / (t1 - t0) * outs[0].get("frame_bytes", 0) / 1e9,
            "config": {"workload": "SURVEY 8f n3: cross-box leg emulated on one machine (two queue roots), GPU 0 -> GPU 1 over NCCL send/recv",
                       "detector": a.detector, "batch": a.batch, "queue_size": a.queue_size, "image": a.image, "common_mode": a.common_mode}}
    print(json.dumps(line), flush=True)
    return 0 if total == nev and all(r == 0 for r in rcs) else 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"], help="reference = configs[0]: the reference's per-event CPU flow (1 rank, NumPy calib, pickled frames)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--consumers", type=int, default=0)
    ap.add_argument("--events-per-rank", type=int, default=2048)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--queue-size", type=int, default=0, help="events; default 1024 per producer rank.  400 (the reference README) makes refused puts and their 0.1-2 s back-off dominate")
    ap.add_argument("--detector", default="epix10k2M")
    ap.add_argument("--image", action="store_true")
    ap.add_argument("--common-mode", default=None)
    ap.add_argument("--bridge", action="store_true", help="two emulated boxes on one machine: producer + queue A on GPU 0, psana_ray.bridge "
                    "(NCCL send/recv) to queue B, consumers on GPU 1; needs --gpus 2")
    a = ap.parse_args()
    if a.impl == "reference":
        if a.events_per_rank == 2048:
            a.events_per_rank = 300                # ~15-30 s of NumPy calibration
        if a.queue_size == 0:
            a.queue_size = 400                     # configs[0] / the reference README
        return reference_arm(a)
    if a.queue_size == 0:
        a.queue_size = 1024 * max(1, a.gpus)       # room for 16 batches per producer: a refused put costs a 0.1-2 s back-off
    if a.bridge:
        return bridge_arm(a)
    ncons = a.consumers or a.gpus
    root = tempfile.mkdtemp(prefix="psr_stream_")
    env = dict(os.environ, PSANA_RAY_DIR=root, PYTHONPATH=PKG + os.pathsep + os.environ.get("PYTHONPATH", ""), PSANA_RAY_DRAIN_S="120")
    cons_src = os.path.join(root, "consumer.py")
    with open(cons_src, "w") as fh:
        fh.write(CONSUMER % {"pkg": PKG})
    nev = a.events_per_rank * a.gpus
    pargs = ["--exp", "bench", "--run", "1", "--detector_name", a.detector, "--queue_name", "stream", "--ray_namespace", "bench",
             "--queue_size", str(a.queue_size), "--batch_size", str(a.batch), "--num_events", str(nev), "--num_consumers", str(ncons),
             "--reliable_sentinels", "--batch_source", "--log_level", "WARNING"]
    if not a.image:
        pargs.append("--calib")
    if a.common_mode:
        pargs += ["--common_mode", a.common_mode]
    prods = [subprocess.Popen([sys.executable, "-m", "psana_ray.producer"] + pargs,
                              env=dict(env, RANK=str(r), WORLD_SIZE=str(a.gpus), LOCAL_RANK=str(r))) for r in range(a.gpus)]
    cons = [subprocess.Popen([sys.executable, cons_src, str(c % a.gpus), str(a.gpus)], env=env, stdout=subprocess.PIPE, text=True) for c in range(ncons)]
    outs = [json.loads(c.communicate(timeout=240)[0].strip().splitlines()[-1]) for c in cons]
    rcs = [p.wait(timeout=300) for p in prods]
    try:
        sys.path.insert(0, PKG)
        from psana_ray.shared_queue import get_queue
        q = get_queue("stream", "bench", root); stats = q.stats(); q.shutdown()
    except Exception:
        stats = None
    total = sum(o["n"] for o in outs)
    t0 = min(o["t_first"] for o in outs if o["t_first"]); t1 = max(o["t_last"] for o in outs)
    # steady state: the second half of the run (the first batches pay for CUDA-IPC mapping of every producer's slabs and for
    # enabling peer access between every pair of GPUs)
    tm = 0.5 * (t0 + t1)
    def upto(o, t):
        k = 0
        for tt, nn in o.get("marks", []):
            if tt <= t: k = nn
        return k
    steady = sum(o["n"] - upto(o, tm) for o in outs) / (t1 - tm)
    line = {"metric": "streamed calibrated frames/sec, producer -> queue -> consumer over CUDA IPC", "value": total / (t1 - t0), "unit": "frames/s",
            "steady_state_value": steady,
            "n_gpus": a.gpus, "producers": a.gpus, "consumers": ncons, "events": total, "expected_events": nev, "seconds": t1 - t0,
            "per_consumer": [o["n"] for o in outs], "producer_exit_codes": rcs, "queue_stats": stats,
            "config": {"workload": "configs[4]: producer -> queue -> multi-consumer streaming, GPU-resident CUDA-IPC hand-off",
                       "detector": a.detector, "batch": a.batch, "queue_size": a.queue_size, "image": a.image, "common_mode": a.common_mode,
                       "source": "synthetic, batches served from a pinned pool (no per-event host work)"}}
    print(json.dumps(line), flush=True)
    return 0 if total == nev and all(r == 0 for r in rcs) else 1


if __name__ == "__main__":
    sys.exit(main())
