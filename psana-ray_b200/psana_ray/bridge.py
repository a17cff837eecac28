"""Cross-box leg of the stream: forwards queued batches from one box's queue to another box's queue.

The reference reaches consumers on other nodes through the Ray cluster (ref: README.md:23-30 -- every
``put``/``get`` is an RPC to the detached actor wherever it runs, frames pickled through the object
store, ref: producer.py:101, data_reader.py:35).  The queue server and the CUDA-IPC hand-off of this
package are same-box by construction, so a second box gets its own queue and a pair of bridge
processes in between::

    box A:  producers -> queue A -> bridge --role send  ==NCCL send/recv==>  bridge --role recv -> queue B -> consumers  :box B

Both ends are ordinary clients of their local queue: the sender is a consumer of queue A (it maps each
slab through CUDA IPC and gives the lease back when the batch is on the wire), the receiver is a producer
of queue B (its own ring of slabs, the reference's back-off on a full queue, ``num_consumers`` reliable
sentinels at the end).  The payload moves GPU memory to GPU memory with ``torch.distributed`` point-to-point
(backend ``nccl``: GPUDirect RDMA between boxes, NVLink inside one); nothing is pickled except a few hundred
bytes of per-batch metadata.  With backend ``gloo`` the same protocol carries host arrays (CPU tests, or a
box without a GPU on the consuming side).  This is a point-to-point exchange, not a collective.

Wire format per message: ``int64[2] = (kind, meta_len)``, then ``uint8[meta_len]`` = pickled metadata, then
for kinds BATCH/LOOSE the float32 payload of the shape the metadata names.
"""
from __future__ import annotations

import argparse
import logging
import os
import pickle
import sys
import time

import numpy as np

from .data_reader import DataReader
from .shared_queue import create_queue, get_queue
from .transport import EventBatch

KIND_BATCH, KIND_END, KIND_LOOSE = 0, 1, 2


# ---- wire ------------------------------------------------------------------------------------------------
def send_message(dist, torch, peer, kind, meta, payload, dev):
    """One message to ``peer``.  ``payload`` is a contiguous float32 tensor on ``dev`` or None."""
    blob = pickle.dumps(meta, protocol=pickle.HIGHEST_PROTOCOL)
    dist.send(torch.tensor([kind, len(blob)], dtype=torch.int64, device=dev), peer)
    dist.send(torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev), peer)
    if payload is not None:
        dist.send(payload, peer)


def recv_header(dist, torch, peer, dev):
    """(kind, meta) of the next message from ``peer``; the payload, if the kind has one, is still to be received."""
    head = torch.empty(2, dtype=torch.int64, device=dev)
    dist.recv(head, peer)
    kind, n = (int(v) for v in head.tolist())
    buf = torch.empty(n, dtype=torch.uint8, device=dev)
    dist.recv(buf, peer)
    return kind, pickle.loads(buf.cpu().numpy().tobytes())


# ---- sender: a consumer of the local queue ---------------------------------------------------------------------
def run_sender(reader, dist, torch, peer, dev, poll=0.0005, log=logging):
    """Forwards everything ``reader`` yields until the end-of-stream sentinel; returns the number of events sent."""
    nsent = 0
    while True:
        item = reader.read_batch()
        if item is None:
            if reader.end_of_stream:
                send_message(dist, torch, peer, KIND_END, {"events": nsent}, None, dev)
                return nsent
            time.sleep(poll)
            continue
        rank, idx0, data, energy = item
        loose = not isinstance(energy, list)                   # a plain per-event item of a reference-style producer
        t = torch.as_tensor(data)
        if loose:
            t = t[None]
        t = t.to(device=dev, dtype=torch.float32).contiguous()
        meta = {"rank": rank, "idx0": idx0, "count": int(t.shape[0]), "shape": tuple(t.shape), "photon_energy": energy}
        send_message(dist, torch, peer, KIND_LOOSE if loose else KIND_BATCH, meta, t, dev)
        if t.is_cuda:
            torch.cuda.current_stream().synchronize()          # the slab goes back to its producer on the next read_batch
        nsent += meta["count"]
        log.debug(f"bridge: sent rank={rank} idx={idx0}..{idx0 + meta['count'] - 1}")


# ---- receiver: a producer of the local queue ---------------------------------------------------------------------
class HostSlabs:
    """Payload sink of the gloo/CPU mode: a fresh host array per batch, handed to the queue by value."""

    def acquire(self, queue, shape, torch):
        t = torch.empty(shape, dtype=torch.float32)
        return t, (lambda: t.numpy()), None

    def close(self):
        pass


class GpuSlabs:
    """Payload sink of the nccl mode: the receiver's own ring of CUDA-IPC slabs (as a producer's)."""

    def __init__(self, device, batch_size, nslabs):
        self.device, self.batch_size, self.nslabs, self.ring = device, int(batch_size), int(nslabs), None

    def acquire(self, queue, shape, torch, sleep=time.sleep):
        from .transport import SlabRing
        if self.ring is None:
            self.ring = SlabRing(self.nslabs, (max(self.batch_size, shape[0]),) + tuple(shape[1:]), self.device)
        if shape[0] > self.ring.batch_shape[0] or tuple(shape[1:]) != self.ring.batch_shape[1:]:
            raise ValueError(f"batch of shape {shape} does not fit the receiver's slabs {self.ring.batch_shape}")
        waits = 0
        while True:
            self.ring.reclaim(queue.poll_released(self.ring.owner))
            got = self.ring.acquire()
            if got is not None:
                break
            sleep(min(2.0, 0.001 * (2 ** waits)))
            waits = min(waits + 1, 11)
        i, slab, lease = got
        return slab[:shape[0]], (lambda: self.ring.ref(i, shape[0], with_event=False)), lease      # published only once final (see run_receiver)

    def close(self):
        if self.ring is not None:
            self.ring.close()


def run_receiver(queue, slabs, dist, torch, peer, dev, num_consumers=1, put=None, log=logging):
    """Receives until END, queues every batch with the reference's back-off; returns the number of events queued."""
    from .producer import _backoff_put
    put = put or _backoff_put
    nrecv = 0
    while True:
        kind, meta = recv_header(dist, torch, peer, dev)
        if kind == KIND_END:
            for _ in range(num_consumers):                     # reliable sentinels: a full queue must not drop them
                while not queue.put(None):
                    time.sleep(0.01)
            if meta.get("events") != nrecv:
                log.error(f"bridge: peer sent {meta.get('events')} events, {nrecv} arrived")
            return nrecv
        buf, payload_of, lease = slabs.acquire(queue, tuple(meta["shape"]), torch)
        dist.recv(buf, peer)
        if buf.is_cuda:
            torch.cuda.current_stream().synchronize()          # the batch is final before its descriptor is visible
        n = meta["count"]
        if kind == KIND_LOOSE:
            arr = buf[0].cpu().numpy() if buf.is_cuda else payload_of()[0]
            item, weight = [meta["rank"], meta["idx0"], arr, meta["photon_energy"]], 1
            if lease is not None:
                slabs.ring.give_back(lease)
                lease = None
        else:
            item = EventBatch(rank=meta["rank"], idx0=meta["idx0"], count=n, data=payload_of(),
                              photon_energy=list(meta["photon_energy"]), lease=lease)
            weight = n
        if not put(queue, item, weight, lease, meta["rank"], f"bridged idx={meta['idx0']}..{meta['idx0'] + n - 1}"):
            raise RuntimeError("bridge: the local queue died")
        nrecv += n


# ---- CLI ----------------------------------------------------------------------------------------------------------
def parse_arguments(argv=None):
    p = argparse.ArgumentParser(description="psana-ray bridge: forwards a queue's batches to a queue on another box")
    p.add_argument("--role", required=True, choices=["send", "recv"])
    p.add_argument("--backend", default="nccl", choices=["nccl", "gloo"])
    p.add_argument("--peer", type=int, default=None, help="rank of the other end (default: 1 - RANK)")
    p.add_argument("--ray_address", type=str, default="auto")
    p.add_argument("--ray_namespace", type=str, default="default")
    p.add_argument("--queue_name", type=str, default="my")
    p.add_argument("--queue_size", type=int, default=100, help="recv: size of the queue it creates if none exists")
    p.add_argument("--num_consumers", type=int, default=1, help="recv: sentinels to queue at the end of the stream")
    p.add_argument("--batch_size", type=int, default=64, help="recv: capacity of one slab, in events")
    p.add_argument("--nslabs", type=int, default=8)
    p.add_argument("--device", type=int, default=None, help="CUDA device (default: LOCAL_RANK)")
    p.add_argument("--connect_timeout", type=float, default=120.0, help="send: seconds to wait for the local queue to appear")
    p.add_argument("--log_level", type=str, default="INFO", choices=["DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"])
    return p.parse_args(argv)


def main(argv=None):
    a = parse_arguments(argv)
    logging.basicConfig(level=getattr(logging, a.log_level), format="%(asctime)s %(levelname)s %(message)s")
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "2"))
    if world != 2:
        raise SystemExit("the bridge is a pair of processes: WORLD_SIZE must be 2")
    peer = a.peer if a.peer is not None else 1 - rank
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if a.backend == "nccl":
        if not torch.cuda.is_available():
            raise SystemExit("bridge --backend nccl needs a CUDA device (no CPU fallback; use --backend gloo for host arrays)")
        device = a.device if a.device is not None else int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count()
        torch.cuda.set_device(device)
        dev = torch.device("cuda", device)
        dist.init_process_group("nccl", rank=rank, world_size=2, device_id=dev)
    else:
        device, dev = None, torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=2)
    try:
        if a.role == "send":
            reader = DataReader(address=a.ray_address, queue_name=a.queue_name, ray_namespace=a.ray_namespace, device=device, zero_copy=True)   # the batch goes straight onto the wire
            deadline = time.time() + a.connect_timeout
            while True:
                try:
                    reader.connect()
                    break
                except ValueError:
                    if time.time() > deadline:
                        raise
                    time.sleep(0.1)
            n = run_sender(reader, dist, torch, peer, dev)
            reader.close()
            logging.info(f"bridge send: {n} events forwarded")
        else:
            try:
                queue = get_queue(a.queue_name, a.ray_namespace, a.ray_address)
            except ValueError:
                create_queue(a.queue_name, a.ray_namespace, a.queue_size, a.ray_address)
                queue = get_queue(a.queue_name, a.ray_namespace, a.ray_address)
            slabs = GpuSlabs(device, a.batch_size, a.nslabs) if a.backend == "nccl" else HostSlabs()
            n = run_receiver(queue, slabs, dist, torch, peer, dev, num_consumers=a.num_consumers)
            logging.info(f"bridge recv: {n} events queued")
            if a.backend == "nccl":                             # keep the slabs alive until consumers gave every lease back
                t_end = time.time() + float(os.environ.get("PSANA_RAY_DRAIN_S", "30"))
                while slabs.ring is not None and slabs.ring.out and time.time() < t_end:
                    slabs.ring.reclaim(queue.poll_released(slabs.ring.owner))
                    time.sleep(0.01)
            slabs.close()
    finally:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
