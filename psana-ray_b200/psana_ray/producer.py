"""psana-ray-producer -- same CLI and loop as the reference, batched GPU calibration underneath.

Reference: ``psana_ray/producer.py``.  Kept: every flag name and default of ``parse_arguments``
(:17-33; ``--calib`` absent means image mode, :156-159), get-or-create of the queue on rank 0,
barrier, lookup with 10 retries x 1 s (:35-71), the per-rank ``enumerate`` index and the item layout
``[rank, idx, data, photon_energy]`` (:88,:101), mask order bad-pixel then manual (:92-95), 2-D ->
``(1, H, W)`` (:96-97), queue-full back-off ``min(2.0, 0.1 * 2**n) + U(0, 0.5)`` s (:85-86,:105-111),
dead queue -> return (:112-114), ``--max_steps`` cut-off (:89-91), barrier then ``num_consumers``
``None`` sentinels from rank 0 whose put result is not checked (:119-130), SIGINT on rank 0 (:73-76).

Changed (the north star): the source yields RAW frames; a rank stages them into a pinned host
buffer, and ``psr_calib_batch_host`` copies a batch H2D on a side stream and calibrates it on the
rank's own GPU (``device = local_rank % ngpu``); the batch is queued as ONE ``EventBatch`` whose
payload is a CUDA-IPC descriptor.  Ray -> unix-socket queue server, MPI -> mpi4py if importable,
else rank/size from the launcher's environment and the barrier through the queue server.
"""
import argparse
import logging
import os
import random
import signal
import sys
import time
import traceback

import numpy as np

from .event_source import ImageRetrievalMode, SyntheticSource
from .shared_queue import QueueDeadError, QueueNotFoundError, create_queue, get_queue
from .transport import EventBatch


def parse_arguments(argv=None):
    parser = argparse.ArgumentParser(description="PsanaWrapper Data Producer")
    # -- the reference's 13 flags, names and defaults verbatim (ref: producer.py:17-33) --
    parser.add_argument("--exp", type=str, required=True, help="Experiment name")
    parser.add_argument("--run", type=int, required=True, help="Run number")
    parser.add_argument("--detector_name", type=str, required=True, help="Detector name")
    parser.add_argument("--calib", action="store_true", help="Use calib mode")
    parser.add_argument("--uses_bad_pixel_mask", action="store_true", help="Use bad pixel mask")
    parser.add_argument("--manual_mask_path", type=str, default=None, help="Path to a manual mask in npy")
    parser.add_argument("--ray_address", type=str, default="auto", help="Address of the Ray cluster")
    parser.add_argument("--ray_namespace", type=str, default="default", help="Ray namespace to use for both queues")
    parser.add_argument("--queue_name", type=str, default='my', help="Queue name")
    parser.add_argument("--queue_size", type=int, default=100, help="Maximum queue size")
    parser.add_argument("--num_consumers", type=int, default=1, help="Number of consumer processes expected.")
    parser.add_argument("--max_steps", type=int, default=None, help="Maximum number of steps before terminating")
    parser.add_argument("--log_level", type=str, default="INFO", choices=["DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"],
                        help="Logging level")
    # -- additive: how the GPU path is driven --
    parser.add_argument("--batch_size", type=int, default=64, help="Events calibrated per GPU launch / queue item")
    parser.add_argument("--source", type=str, default="synthetic", choices=["synthetic", "psana"],
                        help="Event source: seeded synthetic frames (default) or a real psana run")
    parser.add_argument("--num_events", type=int, default=1024, help="Synthetic source: events in the run (all ranks)")
    parser.add_argument("--common_mode", type=str, default=None, help="cmpars as 'alg,mode,cormax[,npix_min]'")
    parser.add_argument("--calib_mask", type=str, default="status", choices=["status", "total", "none"],
                        help="psana source: the mask det.calib itself multiplies by (psana default: pixel status)")
    parser.add_argument("--epix10ka_cfg_path", type=str, default=None,
                        help="psana source: npy of per-pixel epix10ka config bytes (bits 3,2 pixel gain config | bit 4 trbit), "
                             "instead of reading them from the detector's configuration")
    parser.add_argument("--reliable_sentinels", action="store_true",
                        help="Retry the end-of-stream sentinels with the put back-off instead of dropping them when the queue is full "
                             "(the reference drops them: producer.py:125 ignores put's result)")
    parser.add_argument("--batch_source", action="store_true",
                        help="Synthetic source only: serve whole batches from a pinned pool (no per-event host work); for throughput runs")
    parser.add_argument("--device", type=int, default=None, help="CUDA device (default: local rank %% device count)")
    return parser.parse_args(argv)


# ---------------------------------------------------------------------------------------------------
# rank / size / barrier: MPI if there is one, else the launcher's environment (ref: producer.py:138-140)
# ---------------------------------------------------------------------------------------------------
_RANK_ENVS = (("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK"),
              ("PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID"),
              ("SLURM_PROCID", "SLURM_NTASKS", "SLURM_LOCALID"),
              ("RANK", "WORLD_SIZE", "LOCAL_RANK"))


class World:
    def __init__(self, rank, size, local_rank, comm=None):
        self.rank, self.size, self.local_rank, self.comm = rank, size, local_rank, comm
        self._nbar = 0

    def barrier(self, queue, tag):
        if self.comm is not None:
            self.comm.Barrier()
        elif self.size > 1:
            self._nbar += 1
            queue.barrier(f"{tag}:{self._nbar}", self.size)

    def finalize(self):
        if self.comm is not None:
            from mpi4py import MPI
            MPI.Finalize()


def get_world(env=None):
    env = os.environ if env is None else env
    try:
        from mpi4py import MPI                     # absent in this image; used when present
        comm = MPI.COMM_WORLD
        if comm.Get_size() > 1 or not any(r in env for r, _, _ in _RANK_ENVS):
            lr = int(env.get("OMPI_COMM_WORLD_LOCAL_RANK", env.get("MPI_LOCALRANKID", comm.Get_rank())))
            return World(comm.Get_rank(), comm.Get_size(), lr, comm)
    except ImportError:
        pass
    for r, s, l in _RANK_ENVS:
        if r in env and s in env:
            return World(int(env[r]), int(env[s]), int(env.get(l, env[r])))
    return World(0, 1, 0)


def initialize_queue(address, namespace, queue_name, queue_size, world, max_retries=10, retry_delay=1):
    """The reference's ``initialize_ray`` (ref: producer.py:35-71) over the socket queue server."""
    rank = world.rank
    try:
        queue = None
        if rank == 0:
            try:
                queue = get_queue(queue_name, namespace, address)
                logging.info(f"Connected to existing input queue: {queue_name} in namespace: {namespace}")
            except ValueError:      # QueueNotFoundError is a ValueError, as Ray's get_actor raises
                queue = create_queue(queue_name=queue_name, ray_namespace=namespace, maxsize=queue_size, address=address)
                logging.info(f"Rank {rank}: Shared input queue created successfully.")
        else:
            logging.info(f"Rank {rank}: Waiting for shared input queue to be created...")

        # Wait for rank 0 to create the queue.  Without MPI the barrier itself goes through the queue
        # server, so the other ranks reach it by the retry loop below.
        if world.comm is not None:
            world.comm.Barrier()

        retries = 0
        while retries < max_retries:
            try:
                queue = get_queue(queue_name, namespace, address)
                logging.info(f"Rank {rank}: Successfully connected to shared queue.")
                world.barrier(queue, "init")
                return queue
            except ValueError:
                retries += 1
                logging.info(f"Rank {rank}: Attempt {retries}/{max_retries} to connect to shared queue failed. Retrying...")
                time.sleep(retry_delay)
        raise TimeoutError(f"Rank {rank}: Timeout waiting for shared_queue actor")
    except Exception as e:
        logging.error(f"Rank {rank}: Error in initialize_ray: {e}")
        return None


def signal_handler(sig, frame):
    logging.error("Ctrl+C pressed. Shutting down...")
    sys.exit(0)


# ---------------------------------------------------------------------------------------------------
# batch pipeline: pinned staging -> H2D on a side stream -> sm_100a kernels -> slab -> IPC descriptor
# ---------------------------------------------------------------------------------------------------
def combine_keep_masks(out_shape, bad_pixel_mask, manual_pixel_mask):
    """The two ``np.where(mask, data, 0)`` passes of ref: producer.py:92-95 as one keep mask.

    ``out_shape`` is the per-event data shape BEFORE the 2-D -> (1,H,W) step (masks act on that array),
    and each mask must broadcast against it exactly as ``np.where`` would require."""
    keep = None
    for m in (bad_pixel_mask, manual_pixel_mask):
        if m is None:
            continue
        m = np.broadcast_to(np.asarray(m) != 0, out_shape)      # raises like np.where on a shape mismatch
        keep = m.copy() if keep is None else (keep & m)
    return keep


class GpuBatchPipeline:
    """Owns the rank's GpuCalibrator, one pinned raw staging buffer and the ring of output slabs."""

    def __init__(self, source, psana_mode, device, batch_size, nslabs, bad_pixel_mask=None, manual_pixel_mask=None):
        import torch
        from .calib import GpuCalibrator
        from .transport import SlabRing
        self.torch = torch
        self.spec = source.spec
        self.image = psana_mode == ImageRetrievalMode.image
        self.batch_size = int(batch_size)
        self.device = device
        consts = source.constants()
        self.spec = source.spec                       # a psana source learns nseg from the constants
        self.cal = GpuCalibrator(self.spec, device=device)
        self.cal.set_constants(consts["peds"], consts.get("gain"), consts.get("offs"), consts.get("cfg"), consts.get("mask"))
        self.cal.set_common_mode(source.common_mode())
        if self.image:
            hw = self.cal.set_geometry(*source.geometry())
            self.event_shape = (1,) + tuple(hw)       # ref: producer.py:96-97
            keep = combine_keep_masks(tuple(hw), bad_pixel_mask, manual_pixel_mask)
            self.cal.reserve(self.batch_size)
        else:
            self.event_shape = self.spec.shape
            keep = combine_keep_masks(self.spec.shape, bad_pixel_mask, manual_pixel_mask)
        if keep is not None:
            self.cal.set_keep_mask(keep, image=self.image)
        self.h_raw = torch.empty((self.batch_size,) + self.spec.shape, dtype=torch.int16).pin_memory()
        self._h_raw_np = self.h_raw.numpy().view(np.uint16)
        self.ring = SlabRing(nslabs, (self.batch_size,) + self.event_shape, device)
        self._registered = False
        # publishing is asynchronous, but the rank stays at most this many batches ahead of its GPU: without a bound it
        # would fill the queue in microseconds and spend its time in the reference's put back-off (0.1-2 s per refusal)
        self.max_in_flight = int(os.environ.get("PSANA_RAY_IN_FLIGHT", "2"))
        self._in_flight = []

    def stage(self, k, raw_frame):
        self._h_raw_np[k] = raw_frame

    def register(self, queue):
        """Once, before the first batch: publish the ring's handle table, so that consumers map every slab and event when
        they connect instead of inside the stream."""
        if not self._registered:
            try:
                queue.register_slabs(self.ring.table())
            except Exception:
                pass
            self._registered = True

    def poll(self, queue):
        self.register(queue)
        self.ring.reclaim(queue.poll_released(self.ring.owner))

    def _throttle(self, i):
        from . import _native as N
        import ctypes as C
        while len(self._in_flight) >= self.max_in_flight:
            N.check(self.cal._lib.psr_event_synchronize(C.c_void_p(self.ring.events[self._in_flight.pop(0)])))
        self._in_flight.append(i)

    def process(self, n, queue):
        """Calibrate the n staged frames into a free slab.  Returns (payload, lease) or None if no slab is free."""
        self.poll(queue)
        got = self.ring.acquire()
        if got is None:
            return None
        i, slab, lease = got
        self._throttle(i)
        d_out = slab[:n].view((n,) + (self.cal.image_shape if self.image else self.spec.shape))
        # queued, not waited for: the slab's inter-process event is recorded behind the kernels and the descriptor is
        # published at once; only the staging buffer is waited for, because the next events are staged into it
        self.cal.calib_host_begin(self.h_raw[:n], d_out=d_out, image=self.image, done_event=self.ring.events[i])
        self.cal.calib_host_wait(input_only=True)
        return self.ring.ref(i, n), lease

    def process_view(self, h_raw_view, queue):
        """As process(), for n frames already contiguous in (pinned) host memory: no staging copy."""
        n = h_raw_view.shape[0]
        self.poll(queue)
        got = self.ring.acquire()
        if got is None:
            return None
        i, slab, lease = got
        self._throttle(i)
        d_out = slab[:n].view((n,) + (self.cal.image_shape if self.image else self.spec.shape))
        # the pool the view comes from is read-only, so nothing is waited for here: the next batch's copies queue up
        # right behind this one's
        self.cal.calib_host_begin(h_raw_view, d_out=d_out, image=self.image, done_event=self.ring.events[i])
        return self.ring.ref(i, n), lease

    def unprocess(self, lease):
        self.ring.give_back(lease)

    def close(self):
        try:
            self.cal.calib_host_wait()                # nothing of ours may still be writing a slab that is about to be freed
        except Exception:
            pass
        self.cal.close()
        self.ring.close()


def _wait_for_room(queue, weight, maxsize, poll=0.0005, limit=5.0, poll_max=0.008):
    """Additive: look before putting.  A refused put costs the reference's back-off below -- at least 0.1 s, thousands of
    events at GPU rates -- so a producer that knows the queue's size waits in half-millisecond steps until its batch
    fits (an empty queue always takes one item).  The put itself and its back-off are unchanged; with several producers
    a put can still be refused and then backs off exactly as the reference does."""
    if not maxsize:
        return
    t0 = time.time()
    try:
        while time.time() - t0 < limit:
            sz = queue.size()
            if sz == 0 or sz + weight <= maxsize:
                return
            time.sleep(poll)
            poll = min(poll * 2, poll_max)        # several waiting producers must not keep the single-threaded queue busy
    except Exception:
        return


def _backoff_put(queue, item, weight, lease, rank, what, maxsize=None):
    """The reference's blocking put (ref: producer.py:99-117).  Returns False if the queue is dead."""
    _wait_for_room(queue, weight, maxsize)
    base_delay_in_sec = 0.1
    max_delay_in_sec = 2.0
    retries = 0
    while True:
        try:
            success = queue.put(item, weight=weight, lease=lease)
            if success:
                logging.debug(f"Rank {rank} produced: {what}")
                return True
            logging.info(f"Rank {rank}: Queue is full, waiting...")
            delay_in_sec = min(max_delay_in_sec, base_delay_in_sec * (2 ** retries))
            jitter_in_sec = random.uniform(0, 0.5)
            time.sleep(delay_in_sec + jitter_in_sec)
            if delay_in_sec < max_delay_in_sec:
                retries += 1
        except QueueDeadError:
            logging.error(f"Rank {rank}: Queue actor is dead. Exiting...")
            return False
        except Exception as e:
            logging.error(f"Rank {rank}: Error in produce_data: {e}")
            time.sleep(1)


def produce_batches(source, pipeline, queue, world, max_steps=None, sleep=time.sleep, queue_size=None):
    """Batch fast path of the event loop: the source hands out whole batches in pinned memory.  Same items,
    same per-rank index, same back-off as produce_data; returns False if the queue died."""
    rank = world.rank
    for raw, energies, idx0 in source.iter_raw_batches(pipeline.batch_size):
        n = raw.shape[0]
        if max_steps is not None:
            if idx0 >= max_steps:
                break
            if idx0 + n > max_steps:
                n = max_steps - idx0
                raw, energies = raw[:n], energies[:n]
        waits = 0
        while True:
            res = pipeline.process_view(raw, queue)
            if res is not None:
                break
            sleep(min(0.05, 0.0005 * (2 ** waits)))
            waits = min(waits + 1, 8)
        payload, lease = res
        item = EventBatch(rank=rank, idx0=idx0, count=n, data=payload, photon_energy=list(energies), lease=lease)
        if not _backoff_put(queue, item, n, lease, rank, f"idx={idx0}..{idx0 + n - 1}", queue_size):
            return False
        logging.debug(f"Rank {rank} produced: idx={idx0}..{idx0 + n - 1}")
    return True


def produce_data(source, pipeline, queue, world, num_consumers=1, max_steps=None, sleep=time.sleep, reliable_sentinels=False,
                 batch_source=False, queue_size=None):
    """The reference's ``produce_data`` (ref: producer.py:78-130) with events grouped into batches.

    ``pipeline`` needs ``batch_size``, ``stage(k, raw)``, ``process(n, queue) -> (payload, lease) | None``."""
    rank = world.rank
    B = pipeline.batch_size
    staged, energies, idx0 = 0, [], 0

    def flush():
        nonlocal staged, energies, idx0
        if staged == 0:
            return True
        waits = 0
        while True:
            res = pipeline.process(staged, queue)
            if res is not None:
                break
            logging.info(f"Rank {rank}: all output slabs are still with consumers, waiting...")
            sleep(min(2.0, 0.01 * (2 ** waits)))
            waits = min(waits + 1, 8)
        payload, lease = res
        item = EventBatch(rank=rank, idx0=idx0, count=staged, data=payload, photon_energy=list(energies), lease=lease)
        ok = _backoff_put(queue, item, staged, lease, rank, f"idx={idx0}..{idx0 + staged - 1} | shape={getattr(payload, 'shape', None)}", queue_size)
        logging.info(f"Rank {rank} produced: idx={idx0}..{idx0 + staged - 1} | shape={tuple(getattr(payload, 'shape', ()))}")
        idx0 += staged
        staged, energies = 0, []
        return ok

    alive = True
    if batch_source:
        if not produce_batches(source, pipeline, queue, world, max_steps, sleep, queue_size):
            return
    for idx, (raw, photon_energy) in enumerate(source.iter_raw_events() if not batch_source else ()):
        if max_steps is not None and idx >= max_steps:
            logging.info(f"Rank {rank}: Reached max_steps {max_steps}, terminating")
            break
        pipeline.stage(staged, raw)
        energies.append(photon_energy)
        staged += 1
        if staged == B:
            alive = flush()
            if not alive:
                return  # Exit the function if the queue actor is dead
    if alive and not flush():
        return

    # Signal completion
    world.barrier(queue, "end")
    if rank == 0:
        # Put sentinel values to signal end of data (one per consumer); as in the reference the
        # result of put is not checked, so a full queue drops sentinels (ref: producer.py:121-125)
        try:
            for _ in range(num_consumers):
                if reliable_sentinels:
                    if not _backoff_put(queue, None, 1, None, rank, "sentinel"):
                        break
                else:
                    queue.put(None)
            logging.info(f"Rank 0: {num_consumers} sentinel values sent successfully")
        except QueueDeadError:
            logging.info("Rank 0: Queue actor is dead. Unable to send sentinels.")
        except Exception as e:
            logging.info(f"Rank 0: Error putting sentinel values: {e}")


def make_source(args, world):
    cm = None
    if args.common_mode:
        p = args.common_mode.split(",")
        cm = (int(p[0]), int(p[1]), float(p[2])) + ((int(p[3]),) if len(p) > 3 else ())
    if args.source == "psana":
        from .event_source import PsanaSource
        cfg = np.load(args.epix10ka_cfg_path) if args.epix10ka_cfg_path else None
        return PsanaSource(exp=args.exp, run=args.run, detector_name=args.detector_name, rank=world.rank, size=world.size,
                           epix10ka_cfg=cfg, common_mode=cm, calib_mask=args.calib_mask)
    return SyntheticSource(exp=args.exp, run=args.run, detector_name=args.detector_name, rank=world.rank, size=world.size,
                           num_events=args.num_events, common_mode=cm)


def main(argv=None):
    args = parse_arguments(argv)

    logging.basicConfig(level=getattr(logging, args.log_level),
                        format='%(asctime)s - %(levelname)s - %(message)s')

    world = get_world()
    rank = world.rank

    if rank == 0:
        signal.signal(signal.SIGINT, signal_handler)

    queue = initialize_queue(args.ray_address, args.ray_namespace, args.queue_name, args.queue_size, world)
    if queue is None:
        world.finalize()
        return

    source = make_source(args, world)
    psana_mode = dict(
        calib=ImageRetrievalMode.calib,
        image=ImageRetrievalMode.image,
    )['calib' if args.calib else 'image']

    pipeline = None
    try:
        import torch
        from . import _native
        _native.load()                                # fail loudly: there is no CPU fallback
        device = args.device if args.device is not None else world.local_rank % max(1, torch.cuda.device_count())
        from .dist_util import bind_to_gpu_numa
        cpus = bind_to_gpu_numa(device)              # pinned staging buffers are allocated below: first touch next to the GPU
        if cpus:
            logging.info(f"Rank {rank}: bound to CPUs {cpus[0]}-{cpus[-1]} next to cuda:{device}")
        bad = source.create_bad_pixel_mask() if args.uses_bad_pixel_mask else None
        manual = np.load(args.manual_mask_path) if args.manual_mask_path is not None else None
        nslabs = max(2, -(-args.queue_size // args.batch_size)) + 2
        pipeline = GpuBatchPipeline(source, psana_mode, device, args.batch_size, nslabs, bad, manual)
        pipeline.register(queue)
        produce_data(source, pipeline, queue, world, args.num_consumers, args.max_steps,
                     reliable_sentinels=args.reliable_sentinels, batch_source=args.batch_source, queue_size=args.queue_size)
    except Exception as e:
        logging.error(f"Rank {rank}: Unhandled exception in main: {e}")
        logging.error("Traceback:")
        logging.error(traceback.format_exc())
        raise
    finally:
        # Keep the slabs alive until every lease has come back: a consumer that still holds a batch (or has not yet read
        # one that sits in the queue) must never find freed memory behind its mapping.  Leases of consumers that died are
        # returned by the queue server when their connection drops, so this ends; it is cut short only if the queue is
        # dead, or after PSANA_RAY_DRAIN_S seconds when that is set (unset = wait).
        if pipeline is not None:
            t0, limit, said = time.time(), os.environ.get("PSANA_RAY_DRAIN_S"), 0.0
            try:
                while pipeline.ring.out and (limit is None or time.time() - t0 < float(limit)):
                    pipeline.poll(queue)
                    if time.time() - t0 - said >= 10.0:
                        said = time.time() - t0
                        logging.info(f"Rank {rank}: waiting for {len(pipeline.ring.out)} batch(es) still with consumers or in the queue")
                    time.sleep(0.05)
            except Exception:
                pass
            pipeline.close()
        world.finalize()


if __name__ == "__main__":
    main()
