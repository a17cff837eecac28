#!/usr/bin/env python
"""bench.py -- calibrated frames/s of the psana-ray hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload calib|cm|image|image_nocm|jungfrau] [--impl ours|reference]

A step is one pass of the hot path over one batch of synthetic raw frames (default workload =
BASELINE.json configs[1]: epix10k2M pedestal + gain-switch calibration, batch 64, one B200).  Prints
ONE JSON line: the device-resident throughput (`value`), the end-to-end number through the C ABI
from pinned host memory with the WHOLE calibrated batch copied back (`e2e`; the GPU-resident hand-off
beside it as `e2e.device_resident_out`), the roofline of the step and the CPU baseline (NumPy oracle
on this box's host cores).  The other BASELINE.json workloads (configs[2] with and without the
image, configs[3] Jungfrau) are measured briefly in the same run into `workloads`, each with its own
`value`, `roofline` and `e2e`.  For N > 1 launch with torch.distributed.run, one rank per GPU;
events shard by rank, there is no data-path collective (weak scaling).
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "psana-ray_b200"), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np

BATCH = 64
METRIC = "calibrated epix10k2M frames/sec"
WORKLOADS = {
    # name -> (detector, common-mode cmpars, image?, BASELINE.json config it is)
    "calib":    ("epix10k2M", None, False, "configs[1]: epix10k2M pedestal + gain-switch calibration, batch 64, constants on-chip across the batch"),
    "cm":       ("epix10k2M", (7, 4, 100.0, 10), False, "configs[2] (calib part): epix10k2M calib + per-bank common-mode median + mask, batch 64"),
    "image":    ("epix10k2M", (7, 4, 100.0, 10), True, "configs[2]: epix10k2M calib + per-bank common-mode median + mask + assembled image, batch 64"),
    "image_nocm": ("epix10k2M", None, True, "epix10k2M calib fused with image assembly, batch 64"),
    "jungfrau": ("jungfrau4M", None, False, "configs[3]: Jungfrau 4M 3-gain-stage calibration, batch 64 per GPU"),
}


# =====================================================================================================
# CPU reference arm: the NumPy oracle (the reference's own CPU path, psana det.calib, is not installable
# offline -- see oracle/psana_calib.py), one process per host core like `mpirun -n N`.
# =====================================================================================================

def _cpu_worker(rank, nprocs, detector, cmpars, image, frames, warm, rounds, barrier, q):
    from oracle import psana_calib as oc
    from psana_ray import synthetic as syn
    spec = syn.detector_spec(detector)
    consts = syn.make_constants(spec, seed=0)
    consts["gfac"] = oc.divide_protected(np.ones_like(consts["gain"]), consts["gain"])   # psana caches gfac once per run
    raw = syn.make_raw_frames(spec, consts, max(1, min(frames, 4)), seed=100 + rank)
    geo = syn.make_pixel_index_arrays(spec) if image else None

    def one(f):
        r = raw[f % raw.shape[0]]
        if image:
            return oc.det_image(spec.kind, r, consts, geo[0], geo[1], cmpars)
        return oc.det_calib(spec.kind, r, consts, cmpars)

    for f in range(warm):
        one(f)
    spans = []
    for _ in range(rounds):
        barrier.wait()
        t0 = time.perf_counter()
        for f in range(frames):
            one(f)
        t1 = time.perf_counter()
        spans.append((t0, t1))
    q.put((rank, spans))


def run_cpu_reference(detector, cmpars, image, nprocs, frames_per_proc, rounds, warm=1):
    """Returns list of per-round wall seconds for nprocs*frames_per_proc frames."""
    ctx = mp.get_context("fork")
    barrier = ctx.Barrier(nprocs)
    q = ctx.Queue()
    procs = [ctx.Process(target=_cpu_worker, args=(r, nprocs, detector, cmpars, image, frames_per_proc, warm, rounds, barrier, q))
             for r in range(nprocs)]
    for p in procs:
        p.start()
    res = [q.get() for _ in procs]
    for p in procs:
        p.join()
    walls = []
    for i in range(rounds):
        t0 = min(s[i][0] for _, s in res)
        t1 = max(s[i][1] for _, s in res)
        walls.append(t1 - t0)
    return walls


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# =====================================================================================================
# clocks
# =====================================================================================================

class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples = []
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                try:
                    rs = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.perf_counter(), mhz, rs))
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self, t0, t1):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable"}
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        note = "timed region"
        if len(inside) < 3:
            inside = [s for s in self.samples if s[0] >= t0]
            note = "timed region + probe loop under the same load"
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        reasons = sorted({name for b, name in self.REASONS.items() if bits & b})
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(inside), "sampled": note}


# =====================================================================================================

def algorithmic_bytes_per_frame(spec, image_shape):
    """2 B raw in + 4 B float out per pixel (image mode: 2 B x npix in + 4 B x H x W out)."""
    if image_shape is not None:
        return 2 * spec.npix + 4 * image_shape[0] * image_shape[1]
    return 6 * spec.npix


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as fh:
            d = json.load(fh)
        return float(d["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def ncu_traffic(workload):
    """dram bytes per launch of the dominant kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as fh:
            return json.load(fh).get(workload)
    except Exception:
        return None


def workload_config(name, B, world):
    """The `config` object of the JSON line -- one definition for both arms, so the driver sees the same workload."""
    from psana_ray import synthetic as syn
    detector, cmpars, image, cfgname = WORKLOADS[name]
    spec = syn.detector_spec(detector)
    image_shape = None
    if image:
        iX, iY = syn.make_pixel_index_arrays(spec)
        image_shape = [int(iX.max()) + 1, int(iY.max()) + 1]
    out_px = image_shape[0] * image_shape[1] if image else spec.npix
    return {"workload": cfgname, "detector": spec.name, "frame_shape": list(spec.shape), "batch_per_gpu": B,
            "common_mode": list(cmpars) if cmpars else None, "image": image_shape,
            "sharding": "events by rank, no collective" if world > 1 else "single GPU",
            "l2": f"inputs larger than L2: {2 * spec.npix * B / 1e6:.0f} MB raw in + {4 * out_px * B / 1e6:.0f} MB out per step vs 126 MB L2",
            "seed": 1000}


def reference_arm(args, rank, world):
    if rank != 0:
        return 0
    detector, cmpars, image, cfgname = WORKLOADS[args.workload]
    nprocs = args.cpu_procs or host_cores()
    fpp = args.cpu_frames_per_proc
    walls = run_cpu_reference(detector, cmpars, image, nprocs, fpp, rounds=args.warmup + args.steps, warm=1)
    timed = walls[args.warmup:]
    total = sum(timed)
    frames_per_step = nprocs * fpp
    value = frames_per_step * len(timed) / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(timed),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.workload, args.batch, world),
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": nprocs, "kind": "port",
                         "sample": f"bounded sample of the workload: {frames_per_step} frames per step ({fpp} per process, one process per host core "
                                   f"like mpirun -n N), {len(timed)} timed steps; reference CPU path = NumPy restatement of psana det.calib "
                                   "(psana / ray / mpi4py are not installable offline)"},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)
    return 0


def tiled_raw(syn, spec, consts, B, seed):
    """B raw frames: 16 distinct synthetic frames, tiled; the repeats get their lowest two ADC bits flipped so that every
    frame differs while the noise statistics stay those of SURVEY.md 8d (the common-mode medians depend on them)."""
    raw = syn.make_raw_frames(spec, consts, min(B, 16), seed=seed)
    if raw.shape[0] < B:
        reps = (B + raw.shape[0] - 1) // raw.shape[0]
        raw = np.concatenate([raw] * reps)[:B].copy()
        raw[16:] ^= (np.arange(B - 16, dtype=np.uint16) % 3 + 1)[:, None, None, None]
    return raw


def pcie_probe(torch, du, world, rank, nbytes=256 << 20, reps=8):
    """This box's pinned-copy rates with every rank copying at once, H2D and D2H as SEPARATE barrier-bracketed phases
    (a probe that mixes directions across drifting ranks is not a denominator).  Returns per-rank min / median and the
    node aggregate, GB/s."""
    h_buf = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda")

    def phase(dst, src, rounds=3):
        dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
        best = 0.0
        for _ in range(rounds):                       # the best of a few rounds: one slow round (page placement, a neighbour's burst) is not the link
            du.barrier(); torch.cuda.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            for _ in range(reps):
                dst.copy_(src, non_blocking=True)
            a1.record(); torch.cuda.synchronize()
            best = max(best, reps * nbytes / (a0.elapsed_time(a1) / 1e3) / 1e9)
        du.barrier()
        return best

    out = {}
    for name, dst, src in (("h2d", d_buf, h_buf), ("d2h", h_buf, d_buf)):
        mine = phase(dst, src)
        allr = sorted(du.gather_floats(mine, device="cuda"))
        out[name] = {"this_rank": mine, "per_rank_min": allr[0], "per_rank_median": allr[len(allr) // 2], "node_aggregate": sum(allr)}
    out["how"] = f"{nbytes >> 20} MiB pinned cudaMemcpyAsync x {reps} per rank, best of 3 rounds, all {world} ranks at once, one direction per barrier-bracketed phase, before the e2e loops"
    del h_buf, d_buf
    return out


def measure(torch, du, syn, GpuCalibrator, name, B, steps, warmup, e2e_steps, dev, rank, world, frames_per_block=0,
            sampler=None, pcie=None, wc_staging=False):
    """One workload on this rank's GPU: device-resident throughput (CUDA events per step, max over ranks), the roofline
    of its step, and the end-to-end numbers through the C ABI from pinned host memory."""
    detector, cmpars, image, cfgname = WORKLOADS[name]
    spec = syn.detector_spec(detector)
    consts = syn.make_constants(spec, seed=0)
    raw = tiled_raw(syn, spec, consts, B, seed=1000 + rank)
    cal = GpuCalibrator(spec, device=dev)
    cal.set_constants(consts["peds"], consts.get("gain"), consts.get("offs"), consts.get("cfg"), consts.get("mask"))
    cal.set_common_mode(cmpars)
    image_shape = None
    if image:
        image_shape = cal.set_geometry(*syn.make_pixel_index_arrays(spec))
        cal.reserve(B)
    if frames_per_block:
        cal.set_option("frames_per_block", frames_per_block)

    def barrier():
        du.barrier()
        torch.cuda.synchronize()

    wc_holder = None
    if wc_staging:                                    # write-combined pinned staging: the CPU only ever writes the raw frames
        import ctypes as C
        from psana_ray import _native as N
        lib = N.load(); p = C.c_void_p()
        N.check(lib.psr_host_alloc_wc(C.byref(p), raw.nbytes))
        wc_holder = (lib, p)
        h_np = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_int16)), shape=(raw.size,)).reshape(raw.shape)
        np.copyto(h_np, raw.view(np.int16))
        h_raw = h_np
        d_raw = torch.from_numpy(raw.view(np.int16)).cuda()
    else:
        h_raw = torch.from_numpy(raw.view(np.int16)).pin_memory()
        d_raw = h_raw.cuda(non_blocking=True)
    d_out = torch.empty(cal.out_shape(B, image), dtype=torch.float32, device="cuda")
    h_out = torch.empty(cal.out_shape(B, image), dtype=torch.float32).pin_memory()
    stream = torch.cuda.current_stream()
    launches_per_step = cal.launches_per_batch(B, image)

    # ---- device-resident throughput ---------------------------------------------------------------
    for _ in range(warmup):
        cal.calib(d_raw, out=d_out, image=image)
    barrier()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    if sampler is not None:
        sampler.start()
    t_wall0 = time.perf_counter()
    evs[0].record(stream)
    for i in range(steps):
        cal.calib(d_raw, out=d_out, image=image)
        evs[i + 1].record(stream)
    barrier()
    t_wall1 = time.perf_counter()
    total_ms = evs[0].elapsed_time(evs[-1])
    step_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
    clocks = None
    if sampler is not None:
        if len([s for s in sampler.samples if t_wall0 <= s[0] <= t_wall1]) < 3:     # short region: probe the clocks under the same load
            t_end = time.perf_counter() + 0.25
            while time.perf_counter() < t_end:
                for _ in range(20):
                    cal.calib(d_raw, out=d_out, image=image)
                torch.cuda.synchronize()
        sampler.stop()
        clocks = sampler.summary(t_wall0, t_wall1)
    total_ms_max = du.max_over_ranks(total_ms, device="cuda")
    value = du.whole_job_throughput(B * steps, world, total_ms_max / 1e3)

    # ---- end to end through the C ABI from HOST buffers ---------------------------------------------------
    def timed(fn, n):
        barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        barrier()
        # the calls block on the library's own streams, so the host clock brackets the device work
        return du.whole_job_throughput(B * n, world, du.max_over_ranks(time.perf_counter() - t0, device="cuda"))

    # (1) the headline: pinned host raw in, the WHOLE calibrated batch back in pinned host memory -- what the reference's
    #     consumer gets from DataReader.read() (ref: data_reader.py:35 returns host arrays)
    for _ in range(3):
        cal.calib_host(h_raw, h_out=h_out, image=image)
    e2e_host = timed(lambda: cal.calib_host(h_raw, h_out=h_out, image=image), e2e_steps)
    # (2) the product's hand-off: the calibrated batch STAYS on the GPU (CUDA-IPC handle to the consumers); one calibrated
    #     frame is read back as the step's result
    h_frame = torch.empty(tuple(d_out.shape[1:]), dtype=torch.float32).pin_memory()

    def resident_step():
        cal.calib_host(h_raw, d_out=d_out, image=image)      # synchronous: returns when d_out is final
        h_frame.copy_(d_out[B - 1], non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(3):
        resident_step()
    e2e_resident = timed(resident_step, e2e_steps)

    peak, peak_src = measured_peaks()
    alg_bytes = algorithmic_bytes_per_frame(spec, image_shape) * B
    avg_launch_s = (sum(step_ms) / len(step_ms)) / 1e3
    achieved = alg_bytes / avg_launch_s / 1e9
    h2d_bytes = raw.nbytes
    d2h_bytes = h_out.numel() * 4
    res = {
        "metric": METRIC.replace("epix10k2M", detector), "value": value, "unit": "frames/s", "steps": steps,
        "ms_per_step": total_ms_max / steps,
        "config": workload_config(name, B, world),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic(name), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "avg_launch_ms": avg_launch_s * 1e3,
                     "min_launch_ms": min(step_ms), "kernel_launches_per_step": launches_per_step},
        "e2e": {"value": e2e_host, "unit": "frames/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                "steps": e2e_steps,
                "path": "psr_calib_batch_host: pinned host raw -> H2D chunks on a side stream -> kernels -> D2H of every calibrated "
                        "chunk on a third stream -> the whole calibrated batch in pinned host memory",
                "pcie_frac_of_63GBs_d2h": e2e_host / world * (d2h_bytes / B) / 1e9 / 63.0,
                "device_resident_out": {
                    "value": e2e_resident, "unit": "frames/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": h_frame.numel() * 4,
                    "path": "same H2D + kernels; the batch stays on the GPU for the CUDA-IPC hand-off, one calibrated frame read back",
                    "pcie_frac_of_63GBs_h2d": e2e_resident / world * (h2d_bytes / B) / 1e9 / 63.0}},
        "gpu_launches": steps * launches_per_step,
    }
    if pcie is not None:
        res["e2e"]["pcie_measured_gbs"] = pcie
        # against what THIS box gives with every rank copying at once (slowest rank = the job's pace)
        res["e2e"]["frac_of_measured_d2h"] = e2e_host / world * (d2h_bytes / B) / 1e9 / pcie["d2h"]["per_rank_min"]
        res["e2e"]["device_resident_out"]["frac_of_measured_h2d"] = e2e_resident / world * (h2d_bytes / B) / 1e9 / pcie["h2d"]["per_rank_min"]
    if clocks is not None:
        res["clocks"] = clocks
    cal.close()
    if wc_holder is not None:
        res["e2e"]["staging"] = "write-combined pinned (psr_host_alloc_wc)"
        wc_holder[0].psr_host_free(wc_holder[1])
    del d_raw, d_out, h_out, h_raw
    torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="calib")
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--frames-per-block", type=int, default=0)
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = min(steps, 20)")
    ap.add_argument("--cpu-procs", type=int, default=0, help="CPU reference processes (0 = all host cores)")
    ap.add_argument("--cpu-frames-per-proc", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--also", default="jungfrau,cm,image,image_nocm",
                    help="other BASELINE.json workloads measured into the line's `workloads` object (comma list, '' = none)")
    ap.add_argument("--also-steps", type=int, default=10)
    ap.add_argument("--wc-staging", action="store_true", help="raw frames staged in write-combined pinned memory (psr_host_alloc_wc)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        return reference_arm(args, rank, world)

    detector, cmpars, image, cfgname = WORKLOADS[args.workload]

    # ---- CPU baseline first (fork before CUDA is initialised); rank 0 at N = 1 only ---------------
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        nprocs = min(args.cpu_procs or host_cores(), 32)
        fpp = 8
        walls = run_cpu_reference(detector, cmpars, image, nprocs, fpp, rounds=2, warm=1)
        w1 = run_cpu_reference(detector, cmpars, image, 1, 4, rounds=2, warm=1)
        cpu_baseline = {"value": nprocs * fpp / walls[-1], "unit": "frames/s", "cores": nprocs, "kind": "port",
                        "single_process_frames_per_sec": 4 / w1[-1],
                        "sample": f"{nprocs} processes x {fpp} frames of the same workload (NumPy oracle), second of two rounds; "
                                  f"host has {host_cores()} usable cores",
                        "frames_per_sec_per_core": fpp / walls[-1]}

    import torch
    from psana_ray import synthetic as syn
    from psana_ray.calib import GpuCalibrator

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    dev = local_rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    from psana_ray import dist_util as du
    numa_cpus = du.bind_to_gpu_numa(dev)              # before any pinned allocation: staging memory next to the GPU
    du.init("nccl", device=torch.device("cuda", dev))

    pcie = pcie_probe(torch, du, world, rank)         # the box's real pinned-copy rates, before anything else loads the links
    e2e_steps = args.e2e_steps or min(args.steps, 20)
    main_res = measure(torch, du, syn, GpuCalibrator, args.workload, args.batch, args.steps, args.warmup, e2e_steps, dev, rank, world,
                       frames_per_block=args.frames_per_block, sampler=ClockSampler(dev), pcie=pcie, wc_staging=args.wc_staging)
    others = {}
    for name in [w for w in args.also.split(",") if w and w != args.workload]:
        r = measure(torch, du, syn, GpuCalibrator, name, args.batch, args.also_steps, 3, min(args.also_steps, 5), dev, rank, world, pcie=pcie)
        others[name] = {k: r[k] for k in ("metric", "value", "unit", "steps", "ms_per_step", "config", "roofline", "e2e", "gpu_launches")}

    if rank == 0:
        line = {
            "metric": main_res["metric"], "value": main_res["value"], "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": main_res["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": main_res["config"], "roofline": main_res["roofline"], "cpu_baseline": cpu_baseline, "e2e": main_res["e2e"],
            "workloads": others,
            "numa_bind_cpus": (f"{numa_cpus[0]}-{numa_cpus[-1]} ({len(numa_cpus)})" if numa_cpus else None),
            "gpu_launches": main_res["gpu_launches"],
            "clocks": main_res.get("clocks"),
        }
        _emit(line)
    du.finalize()
    return 0


def _emit(line):
    """The JSON line goes to the process's ORIGINAL stdout; everything else written to fd 1 meanwhile (NCCL's version banner,
    library chatter) was diverted to stderr, so stdout carries exactly one line."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT, data)


_REAL_STDOUT = 1

if __name__ == "__main__":
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    sys.exit(main())
