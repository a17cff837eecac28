"""Consumer-side batching into PyTorch: the "Consumers -> Batches -> PyTorch task" box of the reference's
architecture figure (ref: figures/psana-ray-architecture.png; README.md:31-35), which the reference leaves to
user code (ref: examples/psana_consumer.py:30-40 polls event by event).

``PsanaRayDataset`` is a ``torch.utils.data.IterableDataset`` over a ``DataReader`` that yields whole
GPU-resident batches -- ``(data (n, ...), rank, idx0, photon_energies)`` -- with zero copies: ``data`` is a view
into the producer's slab.  The slab is handed back to the producer when the NEXT batch is requested, so a
training/inference step must finish with (or clone) a batch before it asks for another.  Use it with
``DataLoader(ds, batch_size=None)`` and ``num_workers=0``: CUDA-IPC mappings belong to the consuming process.
"""
from __future__ import annotations

import time

import torch
from torch.utils.data import IterableDataset

from .data_reader import DataReader


class PsanaRayDataset(IterableDataset):
    def __init__(self, address="auto", queue_name="shared_queue", ray_namespace="my", device=None,
                 poll_s=0.001, clone=False, connect_timeout_s=0.0):
        self.kw = dict(address=address, queue_name=queue_name, ray_namespace=ray_namespace, device=device)
        self.poll_s, self.clone, self.connect_timeout_s = poll_s, clone, connect_timeout_s

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
                yield (data.clone() if self.clone else data), rank, idx0, energies
        finally:
            reader.close()
