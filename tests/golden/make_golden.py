#!/usr/bin/env python
"""Generate tests/golden/kat_v1.json -- hand-derived known-answer vectors for the calibration path.

The reference holds no golden vectors for this path and psana cannot be imported offline
(parity unpinned), so these vectors are authored from the *definitions*
in SURVEY.md section 8a-EXT / BASELINE.md section 2, independently of ``oracle/psana_calib.py``:

  * decode tables are written out literally from the known-answer table;
  * numeric expectations come from a scalar, per-pixel, pure-Python restatement below
    (np.float32 scalars, Python ``sorted``), which shares no code with the NumPy oracle.

Run from the repo root:  python tests/golden/make_golden.py
Floats are stored as uint32 bit patterns (``f32bits``) so the fixtures are exact.
"""
import json
import os

import numpy as np

F = np.float32

# --- literal decode table (SURVEY.md 8a-EXT "Gain-decode known-answer table") -------------------
# (cfg bits (3,2), trbit, data bit 14) -> (row index into peds/gain or None, common-mode eligible)
EPIX_TABLE = {}
for tr in (0, 1):
    for d14 in (0, 1):
        EPIX_TABLE[(0b10, tr, d14)] = (2, False)       # FL
        EPIX_TABLE[(0b01, tr, d14)] = (None, False)    # no map -> ped 0, factor 1
for d14 in (0, 1):
    EPIX_TABLE[(0b11, 1, d14)] = (0, True)             # FH
    EPIX_TABLE[(0b11, 0, d14)] = (1, True)             # FM
EPIX_TABLE[(0b00, 1, 0)] = (3, True)                   # AHL-H
EPIX_TABLE[(0b00, 0, 0)] = (4, True)                   # AML-M
EPIX_TABLE[(0b00, 1, 1)] = (5, False)                  # AHL-L
EPIX_TABLE[(0b00, 0, 1)] = (6, False)                  # AML-L

# raw bits (15,14) -> (index or None, eligible)
JF_TABLE = {0b00: (0, True), 0b01: (1, True), 0b11: (2, False), 0b10: (None, False)}


def bits(x):
    return int(np.array(x, dtype=np.float32).view(np.uint32))


def gfac_of(g):
    g = F(g)
    return F(1) / g if g != 0 else F(0)


def median_f32(vals):
    v = sorted(F(x) for x in vals)
    n = len(v)
    return (v[(n - 1) // 2] + v[n // 2]) / F(2)


def cm_2d(a, elig, idxs, cormax, npix_min):
    sel = [i for i in idxs if elig[i]]
    if len(sel) < npix_min:
        return
    cm = median_f32([a[i] for i in sel])
    if abs(cm) > cormax:
        return
    for i in sel:
        a[i] = F(a[i] - cm)


def cm_line(a, elig, idxs, cormax, npix_min):
    """rows / cols guard style: npix > npix_min and |cm| < cormax, both strict."""
    sel = [i for i in idxs if elig[i]]
    if not (len(sel) > npix_min):
        return
    cm = median_f32([a[i] for i in sel])
    if not (abs(cm) < cormax):
        return
    for i in sel:
        a[i] = F(a[i] - cm)


def common_mode_segment(a, elig, rows, cols, nbanks, mode, cormax, npix_min):
    """a, elig: flat row-major lists for one segment. Order: 4 (banks per half), 1 (rows per bank), 2 (cols per half)."""
    h = rows // 2
    bw = cols // nbanks
    if mode & 4:
        for r0 in (0, h):
            for b in range(nbanks):
                cm_2d(a, elig, [r * cols + c for r in range(r0, r0 + h) for c in range(b * bw, (b + 1) * bw)], cormax, npix_min)
    if mode & 1:
        for b in range(nbanks):
            for r in range(rows):
                cm_line(a, elig, [r * cols + c for c in range(b * bw, (b + 1) * bw)], cormax, npix_min)
    if mode & 2:
        for r0 in (0, h):
            for c in range(cols):
                cm_line(a, elig, [r * cols + c for r in range(r0, r0 + h)], cormax, npix_min)


def epix_calib_scalar(raw, cfg, peds, gain, mask, shape, cmpars):
    nseg, rows, cols = shape
    n = rows * cols
    out = []
    for s in range(nseg):
        a, elig, fac = [], [], []
        for i in range(n):
            r = int(raw[s * n + i]); cf = int(cfg[s * n + i])
            idx, el = EPIX_TABLE[((cf >> 2) & 3, (cf >> 4) & 1, (r >> 14) & 1)]
            ped = F(0) if idx is None else F(peds[idx][s * n + i])
            g = F(1) if idx is None else gfac_of(gain[idx][s * n + i])
            a.append(F(F(r & 0x3FFF) - ped))
            fac.append(g)
            elig.append(bool(el) and (mask is None or mask[s * n + i] > 0))
        if cmpars is not None and cmpars[1] > 0:
            common_mode_segment(a, elig, rows, cols, 8, cmpars[1], cmpars[2], cmpars[3])
        for i in range(n):
            v = F(a[i] * fac[i])
            if mask is not None:
                v = F(v * F(mask[s * n + i]))
            out.append(v)
    return out


def jf_calib_scalar(raw, peds, gain, offs, mask, shape, cmpars):
    nseg, rows, cols = shape
    n = rows * cols
    out = []
    for s in range(nseg):
        a, elig, fac = [], [], []
        for i in range(n):
            r = int(raw[s * n + i])
            idx, el = JF_TABLE[r >> 14]
            v = F(r & 0x3FFF)
            if idx is not None:
                v = F(v - F(peds[idx][s * n + i]))
                v = F(v - (F(offs[idx][s * n + i]) if offs is not None else F(0)))
            else:
                v = F(v - F(0))
            a.append(v)
            fac.append(F(1) if idx is None else (gfac_of(gain[idx][s * n + i]) if gain is not None else F(1)))
            elig.append(bool(el) and (mask is None or mask[s * n + i] > 0))
        if cmpars is not None and cmpars[1] > 0:
            common_mode_segment(a, elig, rows, cols, 16, cmpars[1], cmpars[2], cmpars[3])
        for i in range(n):
            v = F(a[i] * fac[i])
            if mask is not None:
                v = F(v * F(mask[s * n + i]))
            out.append(v)
    return out


def main():
    rng = np.random.default_rng(20241013)
    G = {"version": 1, "note": "see make_golden.py; f32 values as uint32 bit patterns"}

    # 1. epix10ka decode: all 64 cbits values (bits 5..0; bits 1,0 are don't-care)
    dec = []
    for cb in range(64):
        idx, el = EPIX_TABLE[((cb >> 2) & 3, (cb >> 4) & 1, (cb >> 5) & 1)]
        dec.append({"cbits": cb, "cfg": cb & 31, "d14": (cb >> 5) & 1, "mode": 255 if idx is None else idx, "cm_eligible": bool(el)})
    G["epix10ka_decode"] = dec

    # 2. jungfrau decode
    G["jungfrau_decode"] = [{"gain_bits": gb, "mode": 255 if JF_TABLE[gb][0] is None else JF_TABLE[gb][0],
                             "cm_eligible": JF_TABLE[gb][1]} for gb in range(4)]

    # 3. single-pixel calibration cases, written out by hand
    #    raw, cfg, (ped, gain) of the selected mode, mask -> expected
    G["epix10ka_pixels"] = [
        # raw=0x4123: d14=1, value 0x123=291; cfg=0 (AML, trbit 0) -> AML-L idx 6; (291-100.5)*(1/2) = 95.25
        {"raw": 0x4123, "cfg": 0, "mode": 6, "ped": bits(100.5), "gain": bits(2.0), "mask": 1, "expect": bits(95.25)},
        # same pixel, d14=0 -> AML-M idx 4; (291-41)*(1/4) = 62.5
        {"raw": 0x0123, "cfg": 0, "mode": 4, "ped": bits(41.0), "gain": bits(4.0), "mask": 1, "expect": bits(62.5)},
        # zero gain -> gfac 0 -> 0
        {"raw": 0x0123, "cfg": 0, "mode": 4, "ped": bits(41.0), "gain": bits(0.0), "mask": 1, "expect": bits(0.0)},
        # masked pixel: (291-41)*0.25*0 = 0
        {"raw": 0x0123, "cfg": 0, "mode": 4, "ped": bits(41.0), "gain": bits(4.0), "mask": 0, "expect": bits(0.0)},
        # cfg(3,2)=01 matches nothing: ped 0, factor 1 -> 291
        {"raw": 0x0123, "cfg": 0b00100, "mode": 255, "ped": bits(999.0), "gain": bits(7.0), "mask": 1, "expect": bits(291.0)},
        # FH (cfg 11, trbit 1): 16383 - 0.5 = 16382.5, gain 8 -> 2047.8125; bit 14 ignored by FH
        {"raw": 0x7FFF, "cfg": 0b11100, "mode": 0, "ped": bits(0.5), "gain": bits(8.0), "mask": 1, "expect": bits(2047.8125)},
        # FL with junk in cfg bits 1,0: (10 - 12) * (1/0.5) = -4
        {"raw": 10, "cfg": 0b01011, "mode": 2, "ped": bits(12.0), "gain": bits(0.5), "mask": 1, "expect": bits(-4.0)},
        # three separate roundings: (5000 - 0.1f) * (1/3f): computed by the scalar restatement
        {"raw": 5000, "cfg": 0b10000, "mode": 3, "ped": bits(0.1), "gain": bits(3.0), "mask": 1,
         "expect": bits(F(F(5000) - F(0.1)) * (F(1) / F(3)))},
    ]
    G["jungfrau_pixels"] = [
        # g0: (1000 - 100.5 - 2.25) * (1/40) = 22.43125 -> float32
        {"raw": 1000, "mode": 0, "ped": bits(100.5), "offs": bits(2.25), "gain": bits(40.0), "mask": 1,
         "expect": bits(F(F(F(1000) - F(100.5)) - F(2.25)) * (F(1) / F(40)))},
        # g1: raw = 01|value 8000
        {"raw": 0x4000 | 8000, "mode": 1, "ped": bits(7000.0), "offs": bits(-1.5), "gain": bits(-2.0), "mask": 1,
         "expect": bits(F(F(F(8000) - F(7000)) - F(-1.5)) * (F(1) / F(-2)))},
        # g2: raw = 11|value
        {"raw": 0xC000 | 9000, "mode": 2, "ped": bits(9100.0), "offs": bits(0.0), "gain": bits(-0.125), "mask": 1,
         "expect": bits(800.0)},
        # invalid 10: ped 0, offset 0, factor 1 -> the 14-bit value
        {"raw": 0x8000 | 1234, "mode": 255, "ped": bits(55.0), "offs": bits(66.0), "gain": bits(77.0), "mask": 1,
         "expect": bits(1234.0)},
        # masked
        {"raw": 1000, "mode": 0, "ped": bits(100.5), "offs": bits(2.25), "gain": bits(40.0), "mask": 0, "expect": bits(0.0)},
        # zero gain
        {"raw": 1000, "mode": 0, "ped": bits(100.5), "offs": bits(2.25), "gain": bits(0.0), "mask": 1, "expect": bits(0.0)},
    ]

    # 4. divide_protected
    gains = [2.0, 0.0, -4.0, 0.164, 16.4, 1e-30, 0.0]
    G["divide_protected"] = {"gain": [bits(g) for g in gains], "gfac": [bits(gfac_of(g)) for g in gains]}

    # 5. median / guard cases on one group of values (all eligible unless listed); expectations by hand
    G["median_cases"] = [
        {"name": "odd", "values": [5, 1, 9, 3, 7], "eligible": [1, 1, 1, 1, 1], "median": bits(5.0)},
        {"name": "even", "values": [5, 1, 9, 3], "eligible": [1, 1, 1, 1], "median": bits(4.0)},
        {"name": "even_fractional", "values": [0.5, 0.25, 8, -3], "eligible": [1, 1, 1, 1], "median": bits(0.375)},
        {"name": "masked_out", "values": [100, 1, 2, 3, 200], "eligible": [0, 1, 1, 1, 0], "median": bits(2.0)},
        {"name": "ties", "values": [2, 2, 2, 7, 7, 1], "eligible": [1, 1, 1, 1, 1, 1], "median": bits(2.0)},
    ]
    # guard semantics: npix_min=4, cormax=3.0
    G["guard_cases"] = {
        "npix_min": 4, "cormax": bits(3.0),
        "cases": [
            # values, eligible -> after 2d-style, after line-style (rows/cols)
            {"name": "npix_eq_npix_min", "values": [1, 2, 3, 4], "eligible": [1, 1, 1, 1],
             "after_2d": [-1.5, -0.5, 0.5, 1.5], "after_line": [1, 2, 3, 4]},          # 2d: n<min false -> applied; line: n>min false
            {"name": "npix_below", "values": [1, 2, 3, 50], "eligible": [1, 1, 1, 0],
             "after_2d": [1, 2, 3, 50], "after_line": [1, 2, 3, 50]},
            {"name": "cm_eq_cormax", "values": [1, 3, 3, 3, 9], "eligible": [1, 1, 1, 1, 1],
             "after_2d": [-2, 0, 0, 0, 6], "after_line": [1, 3, 3, 3, 9]},              # 2d: |cm|>cormax false -> applied; line: |cm|<cormax false
            {"name": "cm_above_cormax", "values": [1, 4, 4, 4, 9], "eligible": [1, 1, 1, 1, 1],
             "after_2d": [1, 4, 4, 4, 9], "after_line": [1, 4, 4, 4, 9]},
            {"name": "neg_cm_eq_cormax", "values": [-1, -3, -3, -3, -9], "eligible": [1, 1, 1, 1, 1],
             "after_2d": [2, 0, 0, 0, -6], "after_line": [-1, -3, -3, -3, -9]},
            {"name": "applied_eligible_only", "values": [1, 2, 2, 2, 5, 40], "eligible": [1, 1, 1, 1, 1, 0],
             "after_2d": [-1, 0, 0, 0, 3, 40], "after_line": [-1, 0, 0, 0, 3, 40]},
            {"name": "fully_masked", "values": [1, 2, 3, 4, 5], "eligible": [0, 0, 0, 0, 0],
             "after_2d": [1, 2, 3, 4, 5], "after_line": [1, 2, 3, 4, 5]},
        ]}

    # 6. scatter with duplicate targets (last in flat C order wins) and gaps (= 0)
    G["scatter"] = {
        "iX": [[0, 0, 2], [2, 1, 0]], "iY": [[0, 1, 1], [1, 3, 0]], "W": [[10, 11, 12], [13, 14, 15]],
        "img": [[15, 11, 0, 0], [0, 0, 0, 14], [0, 13, 0, 0]],
    }

    # 7. tiny full-pipeline cases through the scalar restatement: epix (2,4,16) -> banks of 2x2,
    #    jungfrau (1,4,32) -> banks of 2x2; every common-mode bit word 0..7; npix_min=2
    tiny = []
    shape = (2, 4, 16)
    n = int(np.prod(shape))
    cfg = rng.choice(np.array([0b00000, 0b10000, 0b11100, 0b01100, 0b01000, 0b00100, 0b00011], dtype=np.uint8), size=n)
    peds = rng.uniform(90.0, 110.0, size=(7, n)).astype(np.float32)
    gain = rng.choice(np.array([2.0, 4.0, 0.5, 0.0, 3.0], dtype=np.float32), size=(7, n))
    mask = (rng.random(n) > 0.15).astype(np.uint8)
    raw = (rng.integers(80, 130, size=n).astype(np.uint16) | (rng.integers(0, 2, size=n).astype(np.uint16) << 14))
    tiny.append({"kind": "epix10ka", "shape": list(shape), "alg": 7, "cormax": bits(6.0), "npix_min": 2,
                 "raw": raw.tolist(), "cfg": cfg.tolist(), "mask": mask.tolist(),
                 "peds": peds.view(np.uint32).tolist(), "gain": gain.view(np.uint32).tolist(),
                 "expect_by_mode": {str(m): [bits(v) for v in epix_calib_scalar(raw, cfg, peds, gain, mask, shape, (7, m, 6.0, 2))]
                                    for m in range(8)},
                 "expect_nomask_mode7": [bits(v) for v in epix_calib_scalar(raw, cfg, peds, gain, None, shape, (7, 7, 6.0, 2))]})
    shape = (1, 4, 32)
    n = int(np.prod(shape))
    peds = rng.uniform(90.0, 110.0, size=(3, n)).astype(np.float32)
    offs = rng.uniform(-2.0, 2.0, size=(3, n)).astype(np.float32)
    gain = rng.choice(np.array([40.0, -1.5, -0.125, 0.0], dtype=np.float32), size=(3, n))
    mask = (rng.random(n) > 0.15).astype(np.uint8)
    gb = rng.choice(np.array([0, 0, 0, 0, 1, 1, 3, 2], dtype=np.uint16), size=n)
    raw = (rng.integers(80, 130, size=n).astype(np.uint16) | (gb << 14)).astype(np.uint16)
    tiny.append({"kind": "jungfrau", "shape": list(shape), "alg": 7, "cormax": bits(6.0), "npix_min": 2,
                 "raw": raw.tolist(), "mask": mask.tolist(),
                 "peds": peds.view(np.uint32).tolist(), "gain": gain.view(np.uint32).tolist(),
                 "offs": offs.view(np.uint32).tolist(),
                 "expect_by_mode": {str(m): [bits(v) for v in jf_calib_scalar(raw, peds, gain, offs, mask, shape, (7, m, 6.0, 2))]
                                    for m in range(8)},
                 "expect_nogain_nooffs_mode5": [bits(v) for v in jf_calib_scalar(raw, peds, None, None, mask, shape, (7, 5, 6.0, 2))]})
    G["tiny_pipelines"] = tiny

    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kat_v1.json")
    with open(path, "w") as fh:
        json.dump(G, fh, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
