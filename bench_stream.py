#!/usr/bin/env python
"""bench_stream.py -- BASELINE.json configs[4]: end-to-end producer -> queue -> multi-consumer streaming
with GPU-resident CUDA-IPC batch hand-off.

    python bench_stream.py --gpus N [--consumers M] [--events-per-rank E] [--batch 64] [--detector epix10k2M]

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
        rd = DataReader(queue_name="stream", ray_namespace="bench", device=dev); rd.connect(); break
    except ValueError:
        if time.time() > t_deadline: raise
        time.sleep(0.05)
n = 0; acc = torch.zeros((), dtype=torch.float64, device="cuda"); t_first = None
while True:
    b = rd.read_batch()
    if b is None:
        if rd.end_of_stream: break
        time.sleep(0.0005); continue
    if t_first is None: t_first = time.time()
    acc += b[2].sum(dtype=torch.float64)          # the consumer really reads the frames
    n += b[2].shape[0]
torch.cuda.synchronize(); t_last = time.time()
rd.close()
print(json.dumps({"n": n, "t_first": t_first, "t_last": t_last, "sum": float(acc), "device": dev}))
"""


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--consumers", type=int, default=0)
    ap.add_argument("--events-per-rank", type=int, default=2048)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--queue-size", type=int, default=1024, help="events; 400 (the reference README) makes refused puts and their 0.1-2 s back-off dominate")
    ap.add_argument("--detector", default="epix10k2M")
    ap.add_argument("--image", action="store_true")
    ap.add_argument("--common-mode", default=None)
    a = ap.parse_args()
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
    cons = [subprocess.Popen([sys.executable, cons_src, str(c % a.gpus)], env=env, stdout=subprocess.PIPE, text=True) for c in range(ncons)]
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
    line = {"metric": "streamed calibrated frames/sec, producer -> queue -> consumer over CUDA IPC", "value": total / (t1 - t0), "unit": "frames/s",
            "n_gpus": a.gpus, "producers": a.gpus, "consumers": ncons, "events": total, "expected_events": nev, "seconds": t1 - t0,
            "per_consumer": [o["n"] for o in outs], "producer_exit_codes": rcs, "queue_stats": stats,
            "config": {"workload": "configs[4]: producer -> queue -> multi-consumer streaming, GPU-resident CUDA-IPC hand-off",
                       "detector": a.detector, "batch": a.batch, "queue_size": a.queue_size, "image": a.image, "common_mode": a.common_mode,
                       "source": "synthetic, batches served from a pinned pool (no per-event host work)"}}
    print(json.dumps(line), flush=True)
    return 0 if total == nev and all(r == 0 for r in rcs) else 1


if __name__ == "__main__":
    sys.exit(main())
