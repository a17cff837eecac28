#!/usr/bin/env python
"""Dump golden vectors from a REAL psana run, on a box that has psana (LCLS-I "psana1"): the kit that pins the oracle.

    python tools/dump_psana_golden.py --exp mfxl1038923 --run 58 --detector_name epix10k2M \
        --events 4 [--cmpars 7,2,100,10] [--out tests/golden/psana_epix10k2M_r58.npz]

For N events of the run it stores what psana ITSELF computes on the CPU -- ``det.calib(evt, cmpars=...)`` and
``det.image(evt)`` (the calls behind ``PsanaWrapperSmd.iter_events``, ref: psana_ray/producer.py:88,156-159) -- next to
everything needed to recompute them: ``det.raw(evt)``, pedestals, gain, offset, the per-pixel epix10ka config bytes, the
status mask ``det.calib`` multiplies by, the producer's bad-pixel mask (ref: producer.py:81), the run's common-mode
constant and the geometry index arrays.  ``tests/test_psana_golden.py`` then checks BOTH the NumPy oracle and the GPU path
against the file: gain decode and masks bit-exact, calibrated pixels within 1e-5 relative, medians exact (any difference
in a common-mode median shows as a whole bank / row / column off by more than the tolerance).

The file is a few tens of MB per event for a 2M-pixel detector: keep --events small (2-4) and pick events with
both gain ranges populated.  Nothing here runs on the GPU; it needs only psana + numpy (and this repo on PYTHONPATH
for the epix10ka config-byte helper).  Rehearse it offline with PYTHONPATH=tests/fake_psana.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "psana-ray_b200"))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--exp", required=True)
    ap.add_argument("--run", type=int, required=True)
    ap.add_argument("--detector_name", required=True)
    ap.add_argument("--events", type=int, default=4)
    ap.add_argument("--skip", type=int, default=0, help="events to skip first")
    ap.add_argument("--cmpars", default=None, help="'alg,mode,cormax[,npix_min]' passed to det.calib; default: psana's own (the run's constant)")
    ap.add_argument("--no-image", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)

    from psana_ray.event_source import PsanaSource
    src = PsanaSource(a.exp, a.run, a.detector_name)          # same adapter the producer uses: what it reads is what is dumped
    det = src.det
    consts = src.constants()
    cmpars = None
    if a.cmpars:
        p = a.cmpars.split(",")
        cmpars = (int(p[0]), int(p[1]), float(p[2])) + ((int(p[3]),) if len(p) > 3 else ())
    run_cm = src.common_mode()
    raws, calibs, images, energies = [], [], [], []
    for i, evt in enumerate(src.ds.events()):
        if i < a.skip:
            continue
        raw = det.raw(evt)
        if raw is None:
            continue
        raws.append(np.asarray(raw, dtype=np.uint16).reshape(src.spec.shape))
        cal = det.calib(evt, cmpars=cmpars) if cmpars is not None else det.calib(evt)
        calibs.append(np.asarray(cal, dtype=np.float32).reshape(src.spec.shape))
        if not a.no_image:
            images.append(np.asarray(det.image(evt, nda_in=cal), dtype=np.float32))
        eb = src.ebeam.get(evt) if src.ebeam is not None else None
        energies.append(np.nan if eb is None else eb.ebeamPhotonEnergy())
        if len(raws) >= a.events:
            break
    if not raws:
        raise SystemExit("no event of this run has the detector")
    out = a.out or os.path.join(ROOT, "tests", "golden", f"psana_{src.spec.name}_r{a.run}.npz")
    payload = dict(
        format=np.array("psana-ray golden v1"), exp=np.array(a.exp), run=np.array(a.run), detector_name=np.array(a.detector_name),
        kind=np.array(src.spec.kind), raw=np.stack(raws), calib=np.stack(calibs), photon_energy=np.array(energies, dtype=np.float64),
        peds=consts["peds"], gain=consts["gain"], bad_pixel_mask=np.asarray(src.create_bad_pixel_mask(), dtype=np.uint8),
        cmpars_used=np.array([] if cmpars is None else list(cmpars), dtype=np.float64),
        cmpars_run=np.array([] if run_cm is None else list(run_cm), dtype=np.float64),
    )
    if "mask" in consts:
        payload["calib_mask"] = consts["mask"]
    if "offs" in consts:
        payload["offs"] = consts["offs"]
    if "cfg" in consts:
        payload["cfg"] = consts["cfg"]
    if images:
        gx, gy = src.geometry()
        payload.update(image=np.stack(images), iX=gx, iY=gy)
    np.savez_compressed(out, **payload)
    print(f"wrote {out}: {len(raws)} events of {src.spec.name} {src.spec.shape}, cmpars used {cmpars if cmpars else run_cm}")


if __name__ == "__main__":
    main()
