"""Event sources with the shape of ``psana_wrapper.PsanaWrapperSmd`` as the reference uses it.

Reference call sites (the class itself is third-party, [EXT]): constructed with ``exp=, run=,
detector_name=`` (ref: psana_ray/producer.py:150-154), ``create_bad_pixel_mask()`` (:81),
``iter_events(mode=)`` yielding ``(data, photon_energy)`` (:88), modes ``ImageRetrievalMode.calib /
.image`` (:156-159).  Events are sharded round-robin over MPI ranks by psana's ``MPIDataSource``:
rank r sees the events with ``nevent % size == r``.

The GPU path moves calibration out of the source: a source here yields RAW uint16 frames
(``iter_raw_events``) and exposes the per-run constants once (``constants()``, ``geometry()``), and
the producer calibrates whole batches on its GPU.

  * ``SyntheticSource`` -- seeded frames of the named detector shape (psana and the LCLS file
    system are not available offline); used by the CLI by default, by the tests and by bench.py.
  * ``PsanaSource`` -- the same interface over a real psana ``MPIDataSource`` + ``Detector``; only
    importable where psana is installed; untested here.
"""
from __future__ import annotations

import enum
import zlib

import numpy as np

from . import synthetic as syn


class ImageRetrievalMode(enum.Enum):
    raw = 0
    calib = 1
    image = 2


class SyntheticSource:
    def __init__(self, exp, run, detector_name, rank=0, size=1, num_events=256, pool=8, cfg_variant="aml",
                 common_mode=None, nseg=None):
        self.exp, self.run, self.detector_name = exp, int(run), detector_name
        self.rank, self.size = int(rank), int(size)
        self.num_events = int(num_events)
        self.spec = syn.detector_spec(detector_name)
        if nseg is not None:
            self.spec = self.spec.with_nseg(nseg)
        self.seed = zlib.crc32(f"{exp}:{self.run}:{detector_name}".encode()) & 0x7FFFFFFF
        self._consts = syn.make_constants(self.spec, seed=self.seed, cfg_variant=cfg_variant)
        self._cmpars = common_mode
        self._pool = syn.make_raw_frames(self.spec, self._consts, pool, seed=self.seed + 1)
        self._geo = None

    # -- per-run constants (psana: det.pedestals / gain / offset / mask / common_mode / indexes_x,y) ------
    def constants(self):
        return self._consts

    def common_mode(self):
        return self._cmpars

    def geometry(self):
        if self._geo is None:
            self._geo = syn.make_pixel_index_arrays(self.spec)
        return self._geo

    def create_bad_pixel_mask(self):
        """det.mask(run, calib=True, status=True, edges=True, central=True, unbond=True, unbondnbrs=True): 1 = good."""
        return self._consts["mask"]

    # -- events -----------------------------------------------------------------------------------------------
    def raw_event(self, nevent):
        """Deterministic raw frame of global event number ``nevent`` (any rank computes the same bytes)."""
        base = self._pool[nevent % self._pool.shape[0]]
        twist = np.uint16((nevent * 2654435761 >> 7) & 0x01FF)       # perturbs ADC bits only, keeps the gain bits
        return base ^ twist

    def photon_energy(self, nevent):
        return 9500.0 + 0.25 * (nevent % 400)

    def iter_raw_events(self):
        for n in range(self.rank, self.num_events, self.size):          # nevent % size == rank
            yield self.raw_event(n), self.photon_energy(n)

    def event_numbers(self):
        return list(range(self.rank, self.num_events, self.size))

    # -- batch fast path (throughput runs) --------------------------------------------------------------------
    def iter_raw_batches(self, batch_size, pinned=True):
        """Yields ``(raw (n, nseg, rows, cols) uint16, photon_energies, first_local_index)`` with zero per-event
        host work: the rank's local event i is frame ``i % P`` of a pool of P = 2 * batch_size frames held in
        pinned memory, so a batch is a contiguous view that the H2D copy reads directly.  (The per-event
        ``iter_raw_events`` stays the reference-shaped path; this one exists so that a throughput measurement
        is limited by PCIe and the GPU, not by NumPy in the source.)"""
        P = 2 * batch_size
        reps = -(-P // self._pool.shape[0])
        pool = np.concatenate([self._pool] * reps)[:P]
        p0 = self._pool.shape[0]
        if P > p0:                                   # make the repeated frames differ in their ADC bits
            pool[p0:] ^= (np.arange(P - p0, dtype=np.uint16) % 257 + 1)[:, None, None, None]
        holder = None
        if pinned:
            try:
                import torch
                holder = torch.from_numpy(pool.view(np.int16)).pin_memory()
                pool = holder.numpy().view(np.uint16)
            except Exception:
                holder = None
        self._batch_pool = (pool, holder)
        nlocal = len(range(self.rank, self.num_events, self.size))
        i = 0
        while i < nlocal:
            n = min(batch_size, nlocal - i)
            off = i % P
            if off + n > P:
                n = P - off
            energies = [self.photon_energy(self.rank + (i + k) * self.size) for k in range(n)]
            yield pool[off:off + n], energies, i
            i += n

    def batch_pool_frame(self, local_index):
        """Frame served for local event ``local_index`` by iter_raw_batches (for checks)."""
        pool, _ = self._batch_pool
        return pool[local_index % pool.shape[0]]


def epix10ka_config_bytes(pixel_config, trbits, asic_quadrants=((1, 1), (0, 1), (0, 0), (1, 0))):
    """Per-pixel config byte the library wants (``psr_set_constants`` ``h_cfg``) from what psana's epix10ka
    configuration object holds for ONE segment: ``pixel_config`` = asicPixelConfigArray (rows, cols) whose bits 3,2
    are the pixel gain configuration, ``trbits`` = the four ASICs' trbit values.  The trbit goes to bit 4 of every
    pixel of its ASIC's quadrant, as psana's ``cbits_config_epix10ka`` does [EXT].  ``asic_quadrants[i]`` =
    (row half, column half) of ASIC i; the default is the layout restated from psana (A0 bottom right, A1 top right,
    A2 top left, A3 bottom left) and is a parameter because it could not be checked offline."""
    pca = np.asarray(pixel_config)
    rows, cols = pca.shape[-2:]
    cfg = (pca & 0xC).astype(np.uint8)
    rh, ch = rows // 2, cols // 2
    for asic, (r, c) in enumerate(asic_quadrants):
        if trbits[asic]:
            cfg[..., r * rh:(r + 1) * rh, c * ch:(c + 1) * ch] |= 0x10
    return cfg


class PsanaSource:
    """psana-backed source (LCLS-I psana1).  Needs ``import psana``; raises ImportError otherwise."""

    def __init__(self, exp, run, detector_name, rank=0, size=1, epix10ka_cfg=None):
        import psana                                                      # noqa: F401  (not available offline)
        self._psana = psana
        self.exp, self.run, self.detector_name = exp, int(run), detector_name
        self.rank, self.size = rank, size
        self.ds = psana.MPIDataSource(f"exp={exp}:run={self.run}:smd")
        self.det = psana.Detector(detector_name)
        self.ebeam = psana.Detector("EBeam")
        self.spec = syn.detector_spec(detector_name)
        self._cfg = epix10ka_cfg

    def constants(self):
        r = self.run
        out = {"peds": np.asarray(self.det.pedestals(r), dtype=np.float32),
               "gain": np.asarray(self.det.gain(r), dtype=np.float32),
               "mask": np.asarray(self.create_bad_pixel_mask(), dtype=np.uint8)}
        if self.spec.kind == "jungfrau":
            offs = self.det.offset(r)
            if offs is not None:
                out["offs"] = np.asarray(offs, dtype=np.float32)
        else:
            if self._cfg is None:
                raise RuntimeError("epix10ka needs the per-pixel config bytes (pixel gain bits | trbit << 4); "
                                   "pass epix10ka_cfg built from the detector's configuration object")
            out["cfg"] = np.asarray(self._cfg, dtype=np.uint8)
        self.spec = self.spec.with_nseg(out["peds"].shape[1])
        return out

    def common_mode(self):
        cm = self.det.common_mode(self.run)
        return None if cm is None else tuple(cm)

    def geometry(self):
        return (np.asarray(self.det.indexes_x(self.run), dtype=np.int32),
                np.asarray(self.det.indexes_y(self.run), dtype=np.int32))

    def create_bad_pixel_mask(self):
        return self.det.mask(self.run, calib=True, status=True, edges=True, central=True, unbond=True, unbondnbrs=True)

    def iter_raw_events(self):
        for evt in self.ds.events():                                      # MPIDataSource shards nevent % size == rank
            raw = self.det.raw(evt)
            if raw is None:
                continue
            eb = self.ebeam.get(evt)
            yield np.asarray(raw, dtype=np.uint16), (eb.ebeamPhotonEnergy() if eb is not None else None)
