"""The N > 1 path on CPU: world_size-2 gloo.  Events shard by rank (nevent % size == rank), every rank
calibrates its own shard (here with the oracle as the calibrator double), there is no data-path
collective; torch.distributed only brackets the timing (barrier, max over ranks)."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import json, os, sys, time
sys.path.insert(0, os.path.join(%(root)r, "psana-ray_b200")); sys.path.insert(0, %(root)r)
import numpy as np, torch.distributed as dist
from psana_ray import dist_util as du
from psana_ray.event_source import SyntheticSource
from oracle import psana_calib as oc
rank, world = du.init("gloo")
src = SyntheticSource("exp", 7, "epix10k2M", rank=rank, size=world, num_events=10, nseg=1)
du.barrier()
t0 = time.perf_counter()
sums = {}
for n, (raw, e) in zip(src.event_numbers(), src.iter_raw_events()):
    sums[n] = float(np.float64(oc.det_calib("epix10ka", raw, src.constants()).sum(dtype=np.float64)))
du.barrier()
dt = time.perf_counter() - t0
tmax = du.max_over_ranks(dt)
allsums = [None] * world
dist.all_gather_object(allsums, sums)
alltimes = [None] * world
dist.all_gather_object(alltimes, dt)
if rank == 0:
    print(json.dumps({"tmax": tmax, "times": alltimes, "sums": allsums, "value": du.whole_job_throughput(5, world, tmax)}))
du.finalize()
"""


def test_two_ranks_gloo_shard_without_collectives(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29631", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)),
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    res = json.loads(outs[0][0].strip().splitlines()[-1])
    s0, s1 = ({int(k): v for k, v in d.items()} for d in res["sums"])
    assert sorted(s0) == [0, 2, 4, 6, 8] and sorted(s1) == [1, 3, 5, 7, 9]          # nevent % size == rank, disjoint, complete
    assert res["tmax"] == max(res["times"])                                        # max over ranks, not rank 0's clock
    assert abs(res["value"] - 10 / res["tmax"]) < 1e-9                              # whole-job throughput
    # every rank computed what a single rank would have computed for the same event
    sys.path.insert(0, os.path.join(ROOT, "psana-ray_b200"))
    from oracle import psana_calib as oc
    from psana_ray.event_source import SyntheticSource
    one = SyntheticSource("exp", 7, "epix10k2M", rank=0, size=1, num_events=10, nseg=1)
    for n in range(10):
        exp = float(np.float64(oc.det_calib("epix10ka", one.raw_event(n), one.constants()).sum(dtype=np.float64)))
        assert (s0 if n % 2 == 0 else s1)[n] == exp


def test_numa_binding_is_a_no_op_without_nvml_or_when_disabled(monkeypatch):
    """bind_to_gpu_numa must never fail a rank: no GPU / no NVML -> None and the affinity is untouched."""
    import os
    from psana_ray import dist_util as du
    before = os.sched_getaffinity(0)
    monkeypatch.setenv("PSANA_RAY_NO_NUMA_BIND", "1")
    assert du.bind_to_gpu_numa(0) is None and os.sched_getaffinity(0) == before
    monkeypatch.delenv("PSANA_RAY_NO_NUMA_BIND")
    got = du.bind_to_gpu_numa(0)                      # here: no GPU, NVML init fails -> None
    assert got is None or set(got) <= before
    os.sched_setaffinity(0, before)
