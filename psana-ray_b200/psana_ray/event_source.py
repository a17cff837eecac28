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
    importable where psana is installed.  Here it runs against ``tests/fake_psana``, a
    stub of the psana calls it makes.
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


def _epix10ka_cfg_from_psana(det, env, source, shape):
    """The (nseg, rows, cols) config bytes of an epix10ka-family detector from psana's own objects [EXT].

    psana's helper is preferred -- ``Detector.UtilsEpix10ka.cbits_config_epix10ka_any`` is what
    ``gain_maps_epix10ka_any`` itself decodes, so its ASIC -> quadrant placement of the trbits is the
    authoritative one.  Its call convention differs between psana releases (config object vs detector), so
    both are tried; if the module has neither, the bytes are built here from the configuration object
    (``epix10ka_config_bytes``, whose quadrant layout is restated, not verified).  Raises RuntimeError with
    what was tried when nothing yields an array of the detector's shape."""
    tried = []
    want = tuple(shape)

    def fit(a):
        a = np.asarray(a)
        if a.shape == want[1:]:
            a = a[None]
        if a.shape != want:
            raise ValueError(f"shape {a.shape}, expected {want}")
        return (a & 0x1C).astype(np.uint8)

    try:
        from Detector import UtilsEpix10ka as ue
    except ImportError as e:
        ue = None
        tried.append(f"import Detector.UtilsEpix10ka: {e}")
    cob = None
    if ue is not None:
        getter = getattr(ue, "get_epix10ka_any_config_object", None)
        if getter is not None:
            try:
                cob = getter(env, source)
            except Exception as e:                                        # noqa: BLE001
                tried.append(f"get_epix10ka_any_config_object: {e!r}")
        fn = getattr(ue, "cbits_config_epix10ka_any", None)
        if fn is not None:
            for arg, label in ((cob, "config object"), (det, "detector")):
                if arg is None:
                    continue
                try:
                    return fit(fn(arg))
                except Exception as e:                                    # noqa: BLE001
                    tried.append(f"cbits_config_epix10ka_any({label}): {e!r}")
    if cob is not None:
        try:
            elems = [cob.elemCfg(i) for i in range(cob.numberOfElements())] if hasattr(cob, "numberOfElements") else [cob]
            segs = [epix10ka_config_bytes(np.asarray(e.asicPixelConfigArray()), [e.asics(i).trbit() for i in range(4)]) for e in elems]
            return fit(np.stack(segs))
        except Exception as e:                                            # noqa: BLE001
            tried.append(f"config object -> epix10ka_config_bytes: {e!r}")
    raise RuntimeError("could not build the epix10ka per-pixel config bytes from psana (pass epix10ka_cfg=...); tried: " + "; ".join(tried))


class PsanaSource:
    """psana-backed source (LCLS-I psana1) with the shape of ``PsanaWrapperSmd`` as the reference uses it
    (ref: producer.py:81,88,150-154), but yielding RAW frames and the run's constants.  Needs ``import psana``.

    ``calib_mask`` is the mask ``det.calib`` itself multiplies by (psana: the pixel-status mask, ``mbits=1``) --
    NOT the producer's bad-pixel mask, which the reference applies afterwards and only under
    ``--uses_bad_pixel_mask`` (ref: producer.py:81,92-93); that one is ``create_bad_pixel_mask()`` and goes to
    ``set_keep_mask``.  "status" | "total" | "none", or an array.
    ``common_mode``: cmpars override (``--common_mode``); None = the run's ``common_mode`` calibration constant.
    ``epix10ka_cfg``: per-pixel config bytes override; None = read from the detector's configuration (above).
    """

    def __init__(self, exp, run, detector_name, rank=0, size=1, epix10ka_cfg=None, common_mode=None, calib_mask="status"):
        import psana                                                      # noqa: F401  (not available offline)
        self._psana = psana
        self.exp, self.run, self.detector_name = exp, int(run), detector_name
        self.rank, self.size = rank, size
        self.ds = psana.MPIDataSource(f"exp={exp}:run={self.run}:smd")
        self.det = psana.Detector(detector_name)
        try:
            self.ebeam = psana.Detector("EBeam")
        except Exception:                                                 # noqa: BLE001  (runs without an EBeam record)
            self.ebeam = None
        self.spec = syn.detector_spec(detector_name)
        self._cfg = epix10ka_cfg
        self._cmpars = common_mode
        self._calib_mask = calib_mask

    def _mask_for_calib(self, shape):
        m = self._calib_mask
        if m is None or (isinstance(m, str) and m == "none"):
            return None
        if isinstance(m, str):
            if m == "status":
                arr = self.det.mask(self.run, calib=False, status=True, edges=False, central=False, unbond=False, unbondnbrs=False)
            elif m == "total":
                arr = self.det.mask(self.run, calib=True, status=True, edges=True, central=True, unbond=True, unbondnbrs=True)
            else:
                raise ValueError(f"calib_mask = {m!r}")
        else:
            arr = m
        arr = np.asarray(arr)
        return (arr.reshape(shape) != 0).astype(np.uint8)

    def constants(self):
        r = self.run
        peds = np.asarray(self.det.pedestals(r), dtype=np.float32)
        if peds.ndim == 3:                                                # single-segment detectors come without the segment axis
            peds = peds[:, None]
        self.spec = self.spec.with_nseg(peds.shape[1])
        full = (self.spec.nmodes,) + self.spec.shape
        out = {"peds": peds.reshape(full)}
        gain = self.det.gain(r)
        out["gain"] = np.ones(full, dtype=np.float32) if gain is None else np.asarray(gain, dtype=np.float32).reshape(full)
        mask = self._mask_for_calib(self.spec.shape)
        if mask is not None:
            out["mask"] = mask
        if self.spec.kind == "jungfrau":
            offs = self.det.offset(r)
            if offs is not None:
                out["offs"] = np.asarray(offs, dtype=np.float32).reshape(full)
        else:
            cfg = self._cfg
            if cfg is None:
                cfg = _epix10ka_cfg_from_psana(self.det, getattr(self.ds, "env", lambda: None)(), getattr(self.det, "source", None), self.spec.shape)
            out["cfg"] = np.asarray(cfg, dtype=np.uint8).reshape(self.spec.shape)
        return out

    def common_mode(self):
        if self._cmpars is not None:
            return tuple(self._cmpars)
        cm = self.det.common_mode(self.run)
        if cm is None:
            return None
        cm = [float(v) for v in np.asarray(cm).ravel()]
        if len(cm) < 3 or int(cm[0]) == 0:
            return None
        return (int(cm[0]), int(cm[1]), float(cm[2])) + ((int(cm[3]),) if len(cm) > 3 else ())

    def geometry(self):
        return (np.asarray(self.det.indexes_x(self.run), dtype=np.int32).reshape(self.spec.shape),
                np.asarray(self.det.indexes_y(self.run), dtype=np.int32).reshape(self.spec.shape))

    def create_bad_pixel_mask(self):
        """``PsanaWrapperSmd.create_bad_pixel_mask`` (ref: producer.py:81): 1 = good; the producer's keep mask."""
        m = self.det.mask(self.run, calib=True, status=True, edges=True, central=True, unbond=True, unbondnbrs=True)
        return np.asarray(m).reshape((-1,) + self.spec.shape[1:])

    def iter_raw_events(self):
        for evt in self.ds.events():                                      # MPIDataSource shards nevent % size == rank
            raw = self.det.raw(evt)
            if raw is None:                                               # detector missing from this event
                continue
            energy = None
            if self.ebeam is not None:
                eb = self.ebeam.get(evt)
                if eb is not None:
                    energy = eb.ebeamPhotonEnergy()
            yield np.asarray(raw).astype(np.uint16, copy=False).reshape((-1,) + self.spec.shape[1:]), energy
