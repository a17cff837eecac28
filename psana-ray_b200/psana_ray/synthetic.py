"""Seeded synthetic detector data: raw frames, per-run calibration constants, geometry.

The reference reads real LCLS runs through psana (ref: psana_ray/producer.py:150-154); psana is
not installable offline, so synthetic frames of the named detector shapes stand in (SURVEY.md
section 8d gives the recipe followed here).  Everything is generated from
``numpy.random.default_rng(seed)`` so that tests, ``bench.py`` and the CPU baseline see the same
bytes.  This module is host-side NumPy only and holds no calibration arithmetic.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class DetectorSpec:
    name: str
    kind: str          # "epix10ka" | "jungfrau"
    nseg: int
    rows: int
    cols: int
    nbanks: int        # column banks per segment row (common mode)
    nmodes: int        # gain modes in the constants (7 epix10ka, 3 jungfrau)

    @property
    def shape(self):
        return (self.nseg, self.rows, self.cols)

    @property
    def npix(self):
        return self.nseg * self.rows * self.cols

    def with_nseg(self, nseg):
        return DetectorSpec(self.name, self.kind, int(nseg), self.rows, self.cols, self.nbanks, self.nmodes)


EPIX10K2M = DetectorSpec("epix10k2M", "epix10ka", 16, 352, 384, 8, 7)
JUNGFRAU4M = DetectorSpec("jungfrau4M", "jungfrau", 8, 512, 1024, 16, 3)

_BY_NAME = {
    "epix10k2m": EPIX10K2M, "epix10ka2m": EPIX10K2M, "epix10ka": EPIX10K2M, "epix10k": EPIX10K2M,
    "jungfrau4m": JUNGFRAU4M, "jungfrau": JUNGFRAU4M,
}


def detector_spec(detector_name: str) -> DetectorSpec:
    """Map a ``--detector_name`` (ref: producer.py:21) to a shape; aliases like MfxEndstation.0:Epix10ka2M.0."""
    key = detector_name.lower()
    if key in _BY_NAME:
        return _BY_NAME[key]
    for k, v in _BY_NAME.items():
        if k in key:
            return v
    raise ValueError(f"unknown detector_name {detector_name!r}; known: {sorted(set(_BY_NAME))}")


# --- epix10ka config-bit decode, as a 64-entry table over cbits = cfg bits 4..2 | d14 << 5 --------
# Rows of the known-answer table in SURVEY.md section 8a-EXT.  255 = no gain map matches.
def epix10ka_mode_of_cbits(cbits: int) -> int:
    if (cbits & 28) == 28: return 0   # FH
    if (cbits & 28) == 12: return 1   # FM
    if (cbits & 12) == 8:  return 2   # FL
    if (cbits & 60) == 16: return 3   # AHL-H
    if (cbits & 60) == 0:  return 4   # AML-M
    if (cbits & 60) == 48: return 5   # AHL-L
    if (cbits & 60) == 32: return 6   # AML-L
    return 255


EPIX10KA_MODE_LUT = np.array([epix10ka_mode_of_cbits(c) for c in range(64)], dtype=np.uint8)

_EPIX_GAIN_NOMINAL = np.array([16.4, 5.5, 0.164, 16.4, 5.5, 0.164, 0.164], dtype=np.float32)
_JF_GAIN_NOMINAL = np.array([41.5, -1.39, -0.11], dtype=np.float32)
_JF_PED_RANGE = ((2500.0, 3500.0), (13000.0, 14500.0), (14000.0, 15500.0))


def make_constants(spec: DetectorSpec, seed: int = 0, cfg_variant: str = "aml", with_offsets: bool = True,
                   bad_fraction: float = 0.01, zero_gain_fraction: float = 0.001) -> dict:
    """Per-run constants: ``peds``, ``gain`` (+ ``offs``, ``cfg``), ``mask`` (uint8, 1 = good).

    ``cfg_variant``: "aml" = every pixel auto-ranging medium/low (cfg(3,2)=00, trbit 0 -> modes
    4/6, the usual SFX setting); "mixed" = blocks of every config-bit pattern incl. the
    non-matching 01 and both trbit values per ASIC quadrant.
    """
    rng = np.random.default_rng(seed)
    shp = spec.shape
    out = {}
    if spec.kind == "epix10ka":
        peds = rng.uniform(2500.0, 3500.0, size=(7,) + shp).astype(np.float32)
        gain = (_EPIX_GAIN_NOMINAL.reshape(7, 1, 1, 1) *
                (1.0 + 0.02 * rng.standard_normal(size=(7,) + shp, dtype=np.float32))).astype(np.float32)
        cfg = np.zeros(shp, dtype=np.uint8)
        if cfg_variant == "mixed":
            # pixel gain-config bits (3,2): four patterns in 16x16 blocks, plus junk in bits 1,0
            pat = rng.integers(0, 4, size=(spec.nseg, (spec.rows + 15) // 16, (spec.cols + 15) // 16), dtype=np.uint8)
            pat = np.repeat(np.repeat(pat, 16, axis=1), 16, axis=2)[:, :spec.rows, :spec.cols]
            cfg |= (pat << 2).astype(np.uint8)
            cfg |= rng.integers(0, 4, size=shp, dtype=np.uint8)          # bits 1,0: ignored by decode
            # trbit per ASIC: 4 ASICs per segment = quadrants of (rows/2, cols/2)
            tr = rng.integers(0, 2, size=(spec.nseg, 2, 2), dtype=np.uint8)
            tr = np.repeat(np.repeat(tr, spec.rows // 2, axis=1), spec.cols // 2, axis=2)
            cfg |= (tr << 4).astype(np.uint8)
        elif cfg_variant != "aml":
            raise ValueError(cfg_variant)
        out["cfg"] = cfg
    else:
        peds = np.stack([rng.uniform(lo, hi, size=shp).astype(np.float32) for lo, hi in _JF_PED_RANGE])
        gain = (_JF_GAIN_NOMINAL.reshape(3, 1, 1, 1) *
                (1.0 + 0.02 * rng.standard_normal(size=(3,) + shp, dtype=np.float32))).astype(np.float32)
        if with_offsets:
            out["offs"] = (5.0 * rng.standard_normal(size=(3,) + shp, dtype=np.float32)).astype(np.float32)
    if zero_gain_fraction > 0:
        gain[rng.random(size=gain.shape, dtype=np.float32) < zero_gain_fraction] = 0.0   # exercises divide_protected
    mask = (rng.random(size=shp, dtype=np.float32) >= bad_fraction).astype(np.uint8)
    mask[:, 0, :] = 0; mask[:, -1, :] = 0; mask[:, :, 0] = 0; mask[:, :, -1] = 0          # edges
    out.update(peds=peds, gain=gain, mask=mask)
    return out


def make_raw_frames(spec: DetectorSpec, consts: dict, nframes: int, seed: int = 1,
                    switched_fraction: float = 0.10, photon_fraction: float = 0.02,
                    out: np.ndarray | None = None) -> np.ndarray:
    """(nframes, nseg, rows, cols) uint16 raw frames with gain bits set.

    epix10ka: 14-bit ADC = clip(rint(N(ped_of_actual_mode, 3) + photons)), data bit 14 set on
    ~``switched_fraction`` of pixels (all photon-hit pixels first), bit 15 clear.
    jungfrau: gain bits 00/01/11 in 90/8/2 % and 0.01 % of the invalid pattern 10.
    """
    rng = np.random.default_rng(seed)
    shp = spec.shape
    raw = out if out is not None else np.empty((nframes,) + shp, dtype=np.uint16)
    peds = consts["peds"]
    if spec.kind == "epix10ka":
        cfg5 = (consts["cfg"].astype(np.uint16) & 0o34)
        m0 = EPIX10KA_MODE_LUT[cfg5]            # mode when data bit 14 = 0
        m1 = EPIX10KA_MODE_LUT[cfg5 | 32]       # mode when data bit 14 = 1
        idx = np.indices(shp)
        def ped_of(m):
            return np.where(m == 255, np.float32(3000.0), peds[np.minimum(m, 6), idx[0], idx[1], idx[2]])
        p0, p1 = ped_of(m0), ped_of(m1)
        for f in range(nframes):
            hit = rng.random(size=shp, dtype=np.float32) < photon_fraction
            photons = np.where(hit, rng.exponential(200.0, size=shp).astype(np.float32), np.float32(0))
            sw = hit | (rng.random(size=shp, dtype=np.float32) < max(0.0, switched_fraction - photon_fraction))
            ped = np.where(sw, p1, p0)
            scale = np.where(sw, np.float32(0.03), np.float32(1.0))       # low gain: fewer ADU per photon
            v = np.rint(ped + 3.0 * rng.standard_normal(size=shp, dtype=np.float32) + photons * scale)
            v = np.clip(v, 0, 16383).astype(np.uint16)
            raw[f] = v | (sw.astype(np.uint16) << 14)
    else:
        for f in range(nframes):
            u = rng.random(size=shp, dtype=np.float32)
            gbits = np.zeros(shp, dtype=np.uint16)
            gbits[u >= 0.90] = 1
            gbits[u >= 0.98] = 3
            gbits[u >= 0.9999] = 2                                          # invalid pattern 10
            mode = np.array([0, 1, 0, 2], dtype=np.intp)[gbits]
            idx = np.indices(shp)
            ped = peds[mode, idx[0], idx[1], idx[2]]
            sig = np.where(gbits == 0, rng.exponential(30.0, size=shp).astype(np.float32) *
                           (rng.random(size=shp, dtype=np.float32) < 0.05), np.float32(0))
            v = np.rint(ped + 3.0 * rng.standard_normal(size=shp, dtype=np.float32) + sig)
            v = np.clip(v, 0, 16383).astype(np.uint16)
            raw[f] = v | (gbits << 14)
    return raw


# --- geometry ---------------------------------------------------------------------------------

def make_pixel_index_arrays(spec: DetectorSpec, gap: int = 10, with_duplicates: bool = True):
    """Integer image-index arrays ``iX, iY`` of shape (nseg, rows, cols), int32.

    epix10k2M-like: 4 quads x 4 panels, quad q rotated by q*90 degrees and placed pinwheel-wise
    around the beam centre, with ``gap``-pixel gaps; with ``with_duplicates`` the last two panels
    of each quad overlap by one pixel line, so some image pixels have two sources (the scatter is
    last-writer-wins).  Other segment counts: panels stacked in two
    columns, every second one rotated by 180 degrees.
    """
    R, C = spec.rows, spec.cols
    r, c = np.meshgrid(np.arange(R, dtype=np.int64), np.arange(C, dtype=np.int64), indexing="ij")
    iX = np.empty(spec.shape, dtype=np.int64)
    iY = np.empty(spec.shape, dtype=np.int64)
    if spec.nseg % 4 == 0 and spec.kind == "epix10ka":
        nquads = spec.nseg // 4
        qh, qw = 2 * R + gap, 2 * C + gap                 # quad extent before rotation
        side = max(qh, qw)
        for s in range(spec.nseg):
            q, p = divmod(s, 4)
            pr, pc = divmod(p, 2)
            x = r + pr * (R + gap)
            y = c + pc * (C + gap)
            if with_duplicates and p == 3:
                x = x - (gap + 1)                          # overlaps panel 1 by one line
            rot = q % 4
            if rot == 1:   x, y = y, (qh - 1) - x
            elif rot == 2: x, y = (qh - 1) - x, (qw - 1) - y
            elif rot == 3: x, y = (qw - 1) - y, x
            ox = (q % 4 in (2, 3)) * (side + gap) + (q // 4) * 2 * (side + gap)
            oy = (q % 4 in (1, 2)) * (side + gap)
            iX[s], iY[s] = x + ox, y + oy
        del nquads
    else:
        for s in range(spec.nseg):
            pr, pc = divmod(s, 2)
            x, y = r, c
            if s % 2 == 1:
                x, y = (R - 1) - x, (C - 1) - y
            ox = pr * (R + gap)
            if with_duplicates and s == spec.nseg - 1 and pr > 0:
                ox -= gap + 1
            iX[s], iY[s] = x + ox, y + pc * (C + gap)
    iX -= iX.min(); iY -= iY.min()
    return iX.astype(np.int32), iY.astype(np.int32)
