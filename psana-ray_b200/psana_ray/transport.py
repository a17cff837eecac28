"""GPU-resident hand-off of calibrated batches between same-box processes.

The reference pickles every calibrated frame into the Ray object store on ``put`` and out of it on
``get`` (ref: psana_ray/producer.py:101, data_reader.py:35): at least two host copies of 8.65 MB
per event plus a synchronous RPC.  Here a producer calibrates a batch into a slab of its own GPU
memory and queues an ``EventBatch`` whose ``data`` is a ``GpuBatchRef``: the slab's 64-byte CUDA-IPC
handle, an offset and a shape.  A consumer maps the slab once (``psr_ipc_open_mem``) and reads the
frames in place -- over NVLink/NVSwitch if it sits on another GPU of the box.

Slab lifetime: every queued batch carries a lease ``(owner, slab, generation)``; the consumer
releases it when it moves on (or the queue server does when the consumer's connection drops) and
the producer reuses a slab only after its lease came back.

Publishing is asynchronous: every slab has an inter-process CUDA event.  The producer queues the
copies and kernels of a batch, records the slab's event behind them and puts the descriptor at once
(the reference's put returns only after the data exists, ref: producer.py:99-103); the consumer makes
its stream wait on that event before it touches the slab, so the queue round trip overlaps the GPU
work and the next batch's H2D copy.  A producer registers its slab table (memory and event handles)
with the queue once; consumers map all of it when they connect, not inside the stream.
"""
from __future__ import annotations

import ctypes as C
import os
import uuid
from dataclasses import dataclass, field

import numpy as np


@dataclass
class GpuBatchRef:
    """Picklable descriptor of a float32 batch living in a producer's GPU slab."""
    handle: bytes              # cudaIpcMemHandle_t (64 bytes) of the slab's base allocation
    offset: int                # byte offset of the batch inside the slab
    shape: tuple               # (n, segs, rows, cols) or (n, 1, H, W)
    device: int                # producer's CUDA device index
    pid: int = field(default_factory=os.getpid)
    dtype: str = "float32"
    event: bytes | None = None # cudaIpcEventHandle_t (64 bytes) of the slab's "batch is final" event; None = already final

    @property
    def nbytes(self):
        return int(np.prod(self.shape)) * 4


@dataclass
class EventBatch:
    """One queue item: ``count`` consecutive events of one rank.  ``DataReader.read()`` hands them out
    one at a time as the reference's ``[rank, idx, data, photon_energy]`` (ref: producer.py:101)."""
    rank: int
    idx0: int                  # per-rank enumerate index of the first event (ref: producer.py:88)
    count: int
    data: object               # GpuBatchRef | np.ndarray (count, ...) float32
    photon_energy: list
    lease: tuple | None = None


class _CudaArray:
    """Minimal __cuda_array_interface__ carrier so torch.as_tensor can wrap a mapped pointer."""

    def __init__(self, ptr, shape, owner=None):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 3, "strides": None}
        self._owner = owner


class SlabMapper:
    """Consumer side: opens each producer slab (and its event) once and hands out tensor views into it."""

    def __init__(self, device: int = 0):
        from . import _native as N
        self._N = N
        self._lib = N.load()
        self.device = device
        self._open = {}            # memory handle bytes -> base pointer
        self._events = {}          # event handle bytes -> cudaEvent_t

    def _open_mem(self, handle):
        base = self._open.get(handle)
        if base is None:
            p = C.c_void_p()
            hb = (C.c_ubyte * self._N.PSR_IPC_HANDLE_BYTES).from_buffer_copy(handle)
            self._N.check(self._lib.psr_ipc_open_mem(hb, self.device, C.byref(p)))
            base = p.value
            self._open[handle] = base
        return base

    def _open_event(self, handle):
        ev = self._events.get(handle)
        if ev is None:
            p = C.c_void_p()
            hb = (C.c_ubyte * self._N.PSR_IPC_HANDLE_BYTES).from_buffer_copy(handle)
            self._N.check(self._lib.psr_ipc_event_open(hb, self.device, C.byref(p)))
            ev = p.value
            self._events[handle] = ev
        return ev

    def premap(self, tables):
        """Open every slab and event of the registered producers now (at connect), so that the first batches of a stream
        do not pay for cudaIpcOpenMemHandle and peer-access set-up.  Tables of producers that are gone are skipped."""
        n = 0
        for t in tables:
            if t.get("pid") == os.getpid():
                continue
            for h, e in zip(t.get("handles", ()), t.get("events", ())):
                try:
                    self._open_mem(h)
                    if e is not None:
                        self._open_event(e)
                    n += 1
                except Exception:
                    break                      # a dead producer's handles cannot be opened
        return n

    def map(self, ref: GpuBatchRef, stream=None):
        """Tensor view of the batch; the current (or given) stream is made to wait for the batch to be final first."""
        import torch
        if ref.pid == os.getpid() and ref.handle not in self._open:
            raise RuntimeError("a CUDA-IPC handle cannot be opened in the process that exported it")
        base = self._open_mem(ref.handle)
        if ref.event is not None:
            ev = self._open_event(ref.event)
            st = stream if stream is not None else torch.cuda.current_stream(torch.device("cuda", self.device)).cuda_stream
            self._N.check(self._lib.psr_event_wait(C.c_void_p(ev), C.c_void_p(st)))
        arr = _CudaArray(base + ref.offset, ref.shape, owner=self)
        return torch.as_tensor(arr, device=torch.device("cuda", self.device))

    def close(self):
        for ev in self._events.values():
            try:
                self._lib.psr_event_destroy(C.c_void_p(ev))
            except Exception:
                pass
        self._events.clear()
        for base in self._open.values():
            try:
                self._lib.psr_ipc_close_mem(C.c_void_p(base), self.device)
            except Exception:
                pass
        self._open.clear()


class SlabRing:
    """Producer side: a ring of output slabs in GPU memory, exported once for CUDA IPC."""

    def __init__(self, nslabs: int, batch_shape: tuple, device: int):
        import torch
        from . import _native as N
        self._N = N
        self._lib = N.load()
        self.device = device
        self.owner = f"{os.getpid()}-{uuid.uuid4().hex[:8]}"
        self.batch_shape = tuple(batch_shape)
        # plain cudaMalloc blocks (not torch's caching allocator): an IPC handle names a whole allocation,
        # so each slab is its own allocation and a batch sits at offset 0 of it
        nbytes = int(np.prod(self.batch_shape)) * 4
        self._ptrs, self.slabs, self.handles, self.offsets, self.events, self.event_handles = [], [], [], [], [], []
        for _ in range(nslabs):
            p = C.c_void_p()
            N.check(self._lib.psr_device_alloc(C.byref(p), nbytes, device))
            self._ptrs.append(p.value)
            self.slabs.append(torch.as_tensor(_CudaArray(p.value, self.batch_shape, owner=self),
                                              device=torch.device("cuda", device)))
            hb = (C.c_ubyte * N.PSR_IPC_HANDLE_BYTES)()
            N.check(self._lib.psr_ipc_export_mem(p, hb))
            self.handles.append(bytes(hb))
            self.offsets.append(0)
            ev, eh = C.c_void_p(), (C.c_ubyte * N.PSR_IPC_HANDLE_BYTES)()
            N.check(self._lib.psr_ipc_event_create(C.byref(ev), eh, device))   # "the batch in this slab is final"
            self.events.append(ev.value)
            self.event_handles.append(bytes(eh))
        self.generation = [0] * nslabs
        self.free = list(range(nslabs))
        self.out = {}              # lease -> slab index

    def table(self):
        """What a consumer needs to map this ring ahead of time (registered with the queue once)."""
        return {"owner": self.owner, "pid": os.getpid(), "device": self.device, "handles": list(self.handles),
                "events": list(self.event_handles), "batch_shape": self.batch_shape}

    def close(self):
        self.slabs = []
        for e in self.events:
            try:
                self._lib.psr_event_destroy(C.c_void_p(e))
            except Exception:
                pass
        self.events = []
        for p in self._ptrs:
            try:
                self._lib.psr_device_free(C.c_void_p(p), self.device)
            except Exception:
                pass
        self._ptrs = []

    def reclaim(self, leases):
        for lease in leases:
            i = self.out.pop(tuple(lease), None)
            if i is not None:
                self.free.append(i)

    def acquire(self):
        """(slab index, tensor, lease) or None when every slab is still out with consumers."""
        if not self.free:
            return None
        i = self.free.pop(0)
        self.generation[i] += 1
        lease = (self.owner, i, self.generation[i])
        self.out[lease] = i
        return i, self.slabs[i], lease

    def give_back(self, lease):
        """Undo acquire() for a batch that was never published."""
        i = self.out.pop(tuple(lease), None)
        if i is not None:
            self.free.insert(0, i)

    def ref(self, i, count, with_event=True):
        shape = (count,) + self.batch_shape[1:]
        return GpuBatchRef(self.handles[i], self.offsets[i], shape, self.device, event=self.event_handles[i] if with_event else None)
