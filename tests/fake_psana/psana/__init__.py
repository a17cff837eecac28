"""Stub of the part of psana (LCLS-I "psana1") that ``psana_ray.event_source.PsanaSource`` touches.

TEST INFRASTRUCTURE: psana cannot be installed offline, so the real-data adapter is exercised
against this stand-in, which serves the seeded synthetic run of ``psana_ray.synthetic`` through
psana-shaped calls -- ``MPIDataSource(...).events()``, ``Detector(name).raw / pedestals / gain /
offset / mask / common_mode / indexes_x / indexes_y``, ``Detector("EBeam").get(evt)`` -- with
psana's conventions (mask 1 = good; events sharded ``nevent % size == rank``).  It also answers
``det.calib(evt, cmpars)`` / ``det.image(evt)`` from the oracle, so ``tools/dump_psana_golden.py``
can be rehearsed here.  Put ``tests/fake_psana`` on PYTHONPATH to use it.

Environment: PSANA_FAKE_NUM_EVENTS (default 64), PSANA_FAKE_CFG (aml | mixed), PSANA_FAKE_CM
("alg,mode,cormax[,npix_min]" = the run's common_mode constant; default none), PSANA_FAKE_NSEG (fewer
segments than the real detector, for quick tests), rank/size from the
same launcher variables the producer reads.
"""
import os

import numpy as np

from psana_ray import synthetic as syn
from psana_ray.event_source import SyntheticSource

_CURRENT = {}


def _rank_size():
    for r, s in (("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE"), ("PMI_RANK", "PMI_SIZE"),
                 ("SLURM_PROCID", "SLURM_NTASKS"), ("RANK", "WORLD_SIZE")):
        if r in os.environ and s in os.environ:
            return int(os.environ[r]), int(os.environ[s])
    return 0, 1


class _Event:
    def __init__(self, nevent):
        self.nevent = nevent


class _Env:
    pass


class MPIDataSource:
    def __init__(self, ds_string):
        kv = dict(p.split("=") for p in ds_string.split(":") if "=" in p)
        self.exp, self.run = kv["exp"], int(kv["run"])
        self.rank, self.size = _rank_size()
        self.num_events = int(os.environ.get("PSANA_FAKE_NUM_EVENTS", "64"))
        _CURRENT["ds"] = self

    def env(self):
        return _Env()

    def events(self):
        for n in range(self.rank, self.num_events, self.size):
            yield _Event(n)


class _EBeamData:
    def __init__(self, e):
        self._e = e

    def ebeamPhotonEnergy(self):
        return self._e


class Detector:
    def __init__(self, name, env=None):
        ds = _CURRENT["ds"]
        self.name, self.source, self.env = name, name, env
        if name == "EBeam":
            self._src = None
            self._any = None
            return
        _CURRENT["area"] = name
        nseg = os.environ.get("PSANA_FAKE_NSEG")
        self._src = SyntheticSource(ds.exp, ds.run, name, num_events=ds.num_events,
                                    cfg_variant=os.environ.get("PSANA_FAKE_CFG", "aml"), nseg=int(nseg) if nseg else None)
        self._c = self._src.constants()

    # ---- EBeam ----
    def get(self, evt):
        return _EBeamData(9500.0 + 0.25 * (evt.nevent % 400))          # SyntheticSource.photon_energy

    # ---- area detector ----
    def raw(self, evt):
        return self._src.raw_event(evt.nevent)

    def pedestals(self, run):
        return self._c["peds"]

    def gain(self, run):
        return self._c["gain"]

    def offset(self, run):
        return self._c.get("offs")

    def status_mask(self):
        """Pixel-status mask: the synthetic mask with the edge rows / columns good again (1 = good)."""
        m = self._c["mask"].copy()
        m[:, 0, :] = 1; m[:, -1, :] = 1; m[:, :, 0] = 1; m[:, :, -1] = 1
        return m

    def mask(self, run, calib=False, status=False, edges=False, central=False, unbond=False, unbondnbrs=False, **kw):
        m = np.ones(self._src.spec.shape, dtype=np.uint8)
        if status:
            m &= self.status_mask()
        if edges:                                                        # what the synthetic mask adds to the status mask
            m[:, 0, :] = 0; m[:, -1, :] = 0; m[:, :, 0] = 0; m[:, :, -1] = 0
        return m

    def common_mode(self, run):
        cm = os.environ.get("PSANA_FAKE_CM")
        return None if not cm else [float(v) for v in cm.split(",")]

    def indexes_x(self, run):
        return self._src.geometry()[0]

    def indexes_y(self, run):
        return self._src.geometry()[1]

    # ---- what the GPU path replaces; answered by the oracle so the golden-dump tool can be rehearsed ----
    def _consts_for_calib(self):
        c = dict(self._c)
        c["mask"] = self.status_mask()
        return c

    def calib(self, evt, cmpars=None, **kw):
        from oracle import psana_calib as oc
        if cmpars is None:
            cm = self.common_mode(0)
            cmpars = None if cm is None else (int(cm[0]), int(cm[1]), float(cm[2])) + ((int(cm[3]),) if len(cm) > 3 else ())
        return oc.det_calib(self._src.spec.kind, self.raw(evt), self._consts_for_calib(), cmpars)

    def image(self, evt, nda_in=None, **kw):
        from oracle import psana_calib as oc
        nda = self.calib(evt) if nda_in is None else nda_in
        return oc.img_from_pixel_arrays(self.indexes_x(0), self.indexes_y(0), nda)
