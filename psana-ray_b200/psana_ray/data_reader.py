"""DataReader -- the consumer handle of the reference, same methods, GPU-resident payload.

Reference: ``psana_ray/data_reader.py`` -- ctor defaults ``address='auto', queue_name="shared_queue",
ray_namespace='my'`` (:5); ``connect()`` (:11-24); ``read()`` returns one item or ``None`` (:31-35),
raises ``RuntimeError`` when unconnected (:32-33) and ``DataReaderError`` when the queue is dead
(:36-37); ``close()`` (:26-29); context manager (:39-44); ``DataReaderError`` (:46-48).

An item is still the reference's ``[rank, idx, data, photon_energy]`` (ref: producer.py:101);
``data`` is now a CUDA ``torch.Tensor`` -- ``(segs, rows, cols)`` or ``(1, H, W)`` float32 -- taken
from the producer's slab, mapped once per slab through CUDA IPC, instead of an ndarray unpickled
from the object store.  Producers queue whole batches; ``read()`` de-batches locally, so most calls
cost no round trip at all.  Additive: ``read_batch()``, ``__iter__`` (stops at the end-of-stream
sentinel, which ``read()`` cannot tell from an empty queue -- both are ``None`` in the reference).

LIFETIME.  The reference hands every consumer an ndarray it owns.  By default so does this reader:
``read()`` / ``read_batch()`` / iteration return a device-to-device COPY of the frame(s) (a few
microseconds on the GPU's own bandwidth), which the caller may keep as long as it likes.  With
``zero_copy=True`` they return VIEWS into the producer's slab instead: those are valid only until the
reader moves past their batch -- the next ``read_batch()``, the ``read()`` that crosses into the next
batch, or ``close()`` -- at which point the slab goes back to its producer and is overwritten.  In
both modes the reader waits for the work queued on the consumer's stream (the copies, or whatever the
caller launched on the views) before it returns a slab, so nothing in flight reads a recycled slab.
"""
import time

from .shared_queue import QueueDeadError, get_queue
from .transport import EventBatch, GpuBatchRef


class DataReader:
    def __init__(self, address='auto', queue_name="shared_queue", ray_namespace='my', device=None, zero_copy=False, premap=True,
                 expected_producers=0, premap_timeout_s=30.0):
        self.address = address
        self.queue_name = queue_name
        self.ray_namespace = ray_namespace
        self.device = device          # CUDA device GPU batches are mapped on; None = current device at first use
        self.zero_copy = zero_copy    # True: hand out views into the producer's slab (see LIFETIME above)
        self.premap = premap          # map every registered producer slab at connect(), not on first touch
        self.expected_producers = int(expected_producers)   # connect() waits (up to premap_timeout_s) until this many producers
        self.premap_timeout_s = premap_timeout_s            # have registered their slabs, so all mapping happens before the stream
        self._queue = None
        self._mapper = None
        self._batch = None            # EventBatch being handed out
        self._tensor = None           # its data, mapped
        self._cursor = 0
        self.end_of_stream = False

    # -- reference surface --------------------------------------------------------------------------------
    def connect(self):
        try:
            self._queue = get_queue(self.queue_name, self.ray_namespace, self.address)
        except Exception as e:
            print(f"Error getting queue actor: {e}")
            self.close()
            raise
        if self.premap:
            try:
                t0 = time.time()
                tables = self._queue.list_slabs()
                while len(tables) < self.expected_producers and time.time() - t0 < self.premap_timeout_s:
                    time.sleep(0.02)
                    tables = self._queue.list_slabs()
                if tables:
                    self._ensure_mapper().premap(tables)
            except Exception:
                pass                                  # mapping falls back to first touch

    def close(self):
        self._finish_batch()
        if self._queue is not None:
            self._queue.close()
        self._queue = None
        if self._mapper is not None:
            self._mapper.close()
            self._mapper = None

    def read(self):
        """One ``[rank, idx, data, photon_energy]``, or ``None`` when nothing is available."""
        found, item = self._read_ex()
        return item if found else None

    def __enter__(self):
        self.connect()
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.close()

    # -- additive -------------------------------------------------------------------------------------------
    def read_batch(self):
        """The rest of the current batch (or the next one) as ``[rank, idx0, data(n, ...), photon_energies]``; None if empty."""
        if self._queue is None:
            raise RuntimeError("DataReader is not connected. Call connect() first.")
        if self._batch is not None and self._cursor >= self._batch.count:
            self._finish_batch()                      # hands the slab back to its producer
        if self._batch is None:
            if not self._fetch():
                return None
            if self._pending_sentinel:
                self._pending_sentinel = False
                return None
            if self._loose is not None:               # a plain per-event item: a batch of one
                item, self._loose = self._loose, None
                return item
        b, c = self._batch, self._cursor
        out = [b.rank, b.idx0 + c, self._own(self._tensor[c:]), list(b.photon_energy[c:])]
        self._cursor = b.count
        return out

    def __iter__(self):
        """Yields items until the producer's end-of-stream sentinel; polls while the queue is merely empty."""
        while True:
            found, item = self._read_ex()
            if found and item is None:
                return
            if not found:
                time.sleep(0.001)
                continue
            yield item

    # -- internals --------------------------------------------------------------------------------------------
    def _own(self, t):
        """What the caller gets: a copy it owns (default), or the view itself (zero_copy)."""
        if self.zero_copy or not hasattr(t, "clone"):
            return t
        return t.clone()

    def _finish_batch(self):
        b, t = self._batch, self._tensor
        self._batch = self._tensor = None
        self._cursor = 0
        if t is not None and getattr(t, "is_cuda", False):
            # the copies handed out (or the caller's kernels on the views) are queued on the consumer's stream: they must
            # have read the slab before its producer may overwrite it
            import torch
            torch.cuda.current_stream(t.device).synchronize()
        if b is not None and b.lease is not None and self._queue is not None:
            try:
                self._queue.release(b.lease)
            except Exception:
                pass

    def _ensure_mapper(self):
        if self._mapper is None:
            import torch
            from .transport import SlabMapper
            dev = self.device if self.device is not None else torch.cuda.current_device()
            self._mapper = SlabMapper(dev)
        return self._mapper

    def _map(self, data):
        if isinstance(data, GpuBatchRef):
            return self._ensure_mapper().map(data)    # the current stream now waits for the producer's "batch is final" event
        return data

    def _fetch(self):
        """Pull the next queue item.  Returns False if the queue was empty; sets end_of_stream on a sentinel."""
        try:
            found, item = self._queue.get_ex()
        except QueueDeadError as e:
            raise DataReaderError("Queue actor is dead.") from e
        if not found:
            return False
        if item is None:
            self.end_of_stream = True
            self._pending_sentinel = True
            return True
        if isinstance(item, EventBatch):
            self._batch, self._cursor = item, 0
            self._tensor = self._map(item.data)
        else:
            self._loose = item            # a plain item put by some other client: hand it through untouched
        return True

    _pending_sentinel = False
    _loose = None

    def _read_ex(self):
        """(found, item): found=False means the queue was empty; (True, None) is the end-of-stream sentinel."""
        if self._queue is None:
            raise RuntimeError("DataReader is not connected. Call connect() first.")
        if self._batch is not None and self._cursor >= self._batch.count:
            self._finish_batch()
        if self._batch is None:
            if not self._fetch():
                return False, None
            if self._pending_sentinel:
                self._pending_sentinel = False
                return True, None
            if self._loose is not None:
                item, self._loose = self._loose, None
                return True, item
        b, c = self._batch, self._cursor
        self._cursor += 1
        return True, [b.rank, b.idx0 + c, self._own(self._tensor[c]), b.photon_energy[c]]


class DataReaderError(Exception):
    """Custom exception for DataReader errors."""
    pass
