"""Consumer-side batching into PyTorch: the "Consumers -> Batches -> PyTorch task" box of the reference's
architecture figure (ref: figures/psana-ray-architecture.png; README.md:31-35), which the reference leaves to
user code (ref: examples/psana_consumer.py:30-40 polls event by event).

``PsanaRayDataset`` is a ``torch.utils.data.IterableDataset`` over a ``DataReader`` that yields whole
GPU-resident batches -- ``(data (n, ...), rank, idx0, photon_energies)``.  By default ``data`` is a device-side
copy the step owns.  With ``zero_copy=True`` it is a view into the producer's slab: no copy at all, but valid
only until the NEXT batch is requested -- at that point the reader waits for the consumer stream's queued work
and hands the slab back to its producer, so a step must be done with (or have cloned what it keeps of) a batch
before it asks for another.  Use it with ``DataLoader(ds, batch_size=None)`` and ``num_workers=0``: CUDA-IPC
mappings belong to the consuming process.
"""
from __future__ import annotations

import time

import torch
from torch.utils.data import IterableDataset

from .data_reader import DataReader


class PsanaRayDataset(IterableDataset):
    def __init__(self, address="auto", queue_name="shared_queue", ray_namespace="my", device=None,
                 poll_s=0.001, zero_copy=False, connect_timeout_s=0.0):
        self.kw = dict(address=address, queue_name=queue_name, ray_namespace=ray_namespace, device=device, zero_copy=zero_copy)
        self.poll_s, self.connect_timeout_s = poll_s, connect_timeout_s

    def __iter__(self):
        reader = DataReader(**self.kw)
        t0 = time.time()
        while True:
            try:
                reader.connect()
                break
            except ValueError:
                if time.time() - t0 >= self.connect_timeout_s:
                    raise
                time.sleep(0.05)
        try:
            while True:
                b = reader.read_batch()
                if b is None:
                    if reader.end_of_stream:
                        return
                    time.sleep(self.poll_s)
                    continue
                rank, idx0, data, energies = b
                if not isinstance(data, torch.Tensor):
                    data = torch.as_tensor(data)
                yield data, rank, idx0, energies
        finally:
            reader.close()
