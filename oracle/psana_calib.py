"""CPU oracle for the psana-ray hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product (``psana-ray_b200/``)
never does: it fails loudly when the CUDA library is missing.

PARITY UNPINNED.  The arithmetic on this path is not in ``/root/reference`` at all.  The
reference reaches it through ``psana_wrapper.PsanaWrapperSmd.iter_events(mode=...)``
(ref: psana_ray/producer.py:11,88,150-159) -> psana ``Detector.calib(evt)`` /
``Detector.image(evt)``.  ``psana-wrapper`` is declared un-pinned
(ref: project.toml:9, ``git+https://github.com/carbonscott/psana-wrapper``, no commit), psana
(LCLS-I ``psana-conda``) is not declared at all, neither is installable offline, and the
reference holds no test or golden vector for the path.  What follows is a NumPy restatement of
the published psana algorithm (upstream files named per function), kept in psana's own NumPy
idiom (``np.select``, ``np.ma.median``, ``np.hsplit``/``np.hstack``, fancy-index assignment) so
rounding and tie behaviour are NumPy's.  It is pinned only against the hand-derived
known-answer vectors under ``tests/golden/``, not against psana output.

Restated uncertainties are parameters, never constants: the Jungfrau common-mode-eligible gain
set (``jf_cm_eligible``), the ``cmpars`` tuple (psana reads its default from the run's
``common_mode`` calibration constant), and the mask keyword set.
"""
from __future__ import annotations

from math import fabs

import numpy as np

# ---------------------------------------------------------------------------------------------
# Detector/GlobalUtils.py :: divide_protected
# ---------------------------------------------------------------------------------------------

M14 = 0x3FFF       # 14-bit ADC value mask (UtilsEpix10ka.py / UtilsJungfrau.py ``M14``)
B14 = 0o40000      # data bit 14: epix10ka auto-ranging "switched to low gain" flag
B04 = 0o20         # cfg bit 4: per-ASIC trbit
B05 = 0o40         # where data bit 14 lands after ``>> 9``

EPIX10KA_MODES = ("FH", "FM", "FL", "AHL-H", "AML-M", "AHL-L", "AML-L")
JUNGFRAU_MODES = ("g0", "g1", "g2")


def divide_protected(num, den, vsub_zero=0):
    """num/den with ``vsub_zero`` where ``den == 0``.

    Upstream: psana ``Detector/GlobalUtils.py::divide_protected`` [EXT]; used once per run as
    ``gfac = divide_protected(np.ones_like(gain), gain)`` in ``calib_epix10ka_any`` /
    ``calib_jungfrau``.
    """
    pro_num = np.select((den != 0,), (num,), default=vsub_zero)
    pro_den = np.select((den != 0,), (den,), default=1)
    return pro_num / pro_den


# ---------------------------------------------------------------------------------------------
# Detector/UtilsEpix10ka.py :: gain_maps_epix10ka_any, calib_epix10ka_any
# ---------------------------------------------------------------------------------------------

def cbits_epix10ka(raw, cfg):
    """Per-pixel control bits: config bits 4..2 plus data bit 14 moved to bit 5.

    ``cfg`` is the per-pixel config byte: bits 3,2 = pixel gain configuration, bit 4 = the
    ASIC's trbit already broadcast over the ASIC quadrant, bits 1,0 (mask/test) ignored by the
    compares below.  SURVEY.md section 8a-EXT row e2.
    """
    cfg = np.asarray(cfg)
    return np.bitwise_or(cfg.astype(np.uint16) & 0o34, (np.asarray(raw) & B14) >> 9)


def gain_maps_epix10ka_any(raw, cfg):
    """Seven boolean maps in calibration-file order FH, FM, FL, AHL-H, AML-M, AHL-L, AML-L.

    Upstream: ``Detector/UtilsEpix10ka.py::gain_maps_epix10ka_any`` [EXT] (e2).  Config bits
    (3,2) = 01 match no map.
    """
    cbits = cbits_epix10ka(raw, cfg)
    return ((cbits & 28) == 28,
            (cbits & 28) == 12,
            (cbits & 12) == 8,
            (cbits & 60) == 16,
            (cbits & 60) == 0,
            (cbits & 60) == 48,
            (cbits & 60) == 32)


def gain_mode_index_epix10ka(raw, cfg):
    """uint8 mode index 0..6 per pixel, 255 where no map matches (decode known-answer form)."""
    gmaps = gain_maps_epix10ka_any(raw, cfg)
    return np.select(gmaps, [np.uint8(i) for i in range(7)], default=np.uint8(255)).astype(np.uint8)


def calib_epix10ka_any(raw, cfg, peds, gain, mask=None, cmpars=None, gfac=None):
    """Calibrated epix10ka frame, float32, same shape as ``raw`` ((nseg, 352, 384)).

    Upstream: ``Detector/UtilsEpix10ka.py::calib_epix10ka_any`` [EXT] (e3); BASELINE.md
    section 2.2.  ``peds``/``gain`` are (7, nseg, 352, 384) float32; ``mask`` uint8 1=good;
    ``cmpars`` = (alg, mode, cormax[, npix_min]) or None for no common mode.
    """
    raw = np.asarray(raw)
    if gfac is None:
        gfac = divide_protected(np.ones_like(gain), gain)
    gmaps = gain_maps_epix10ka_any(raw, cfg)
    gr0, gr1, gr2, gr3, gr4, gr5, gr6 = gmaps

    factor = np.select(gmaps, tuple(gfac[i] for i in range(7)), default=1)
    pedest = np.select(gmaps, tuple(peds[i] for i in range(7)), default=0)

    arrf = np.array(raw & M14, dtype=np.float32) - pedest

    if cmpars is not None:
        mode = int(cmpars[1])
        cormax = cmpars[2]
        npixmin = cmpars[3] if len(cmpars) > 3 else 10
        if mode > 0:
            arr1 = np.ones_like(raw, dtype=np.uint8)
            grhm = np.select((gr0, gr1, gr3, gr4), (arr1, arr1, arr1, arr1), default=0)
            gmask = np.bitwise_and(grhm, mask) if mask is not None else grhm
            if gmask.ndim == 2:
                gmask = gmask.reshape((1,) + gmask.shape)
            a3 = arrf.reshape((-1,) + arrf.shape[-2:])
            hrows = a3.shape[1] // 2      # 176
            for s in range(a3.shape[0]):
                if mode & 4:    # 2-d banks of (176, 48)
                    common_mode_2d_hsplit_nbanks(a3[s, :hrows, :], mask=gmask[s, :hrows, :], nbanks=8, cormax=cormax, npix_min=npixmin)
                    common_mode_2d_hsplit_nbanks(a3[s, hrows:, :], mask=gmask[s, hrows:, :], nbanks=8, cormax=cormax, npix_min=npixmin)
                if mode & 1:    # rows of 48 px per bank over the full segment height
                    common_mode_rows_hsplit_nbanks(a3[s], mask=gmask[s], nbanks=8, cormax=cormax, npix_min=npixmin)
                if mode & 2:    # columns of 176 px per half
                    common_mode_cols(a3[s, :hrows, :], mask=gmask[s, :hrows, :], cormax=cormax, npix_min=npixmin)
                    common_mode_cols(a3[s, hrows:, :], mask=gmask[s, hrows:, :], cormax=cormax, npix_min=npixmin)

    return arrf * factor if mask is None else arrf * factor * mask


# ---------------------------------------------------------------------------------------------
# Detector/UtilsJungfrau.py :: calib_jungfrau
# ---------------------------------------------------------------------------------------------

BW1 = 0o40000   # 1 << 14
BW2 = 0o100000  # 2 << 14
BW3 = 0o140000  # 3 << 14


def gain_mode_index_jungfrau(raw):
    """uint8 index 0,1,2 for gain bits 00,01,11; 255 for the invalid pattern 10."""
    raw = np.asarray(raw)
    gr0 = raw < BW1
    gr1 = (raw >= BW1) & (raw < BW2)
    gr2 = raw >= BW3
    return np.select((gr0, gr1, gr2), [np.uint8(0), np.uint8(1), np.uint8(2)], default=np.uint8(255)).astype(np.uint8)


def calib_jungfrau(raw, peds, gain=None, offs=None, mask=None, cmpars=None, gfac=None,
                   jf_cm_eligible=(0, 1)):
    """Calibrated Jungfrau frame, float32, shape of ``raw`` ((nseg, 512, 1024)).

    Upstream: ``Detector/UtilsJungfrau.py::calib_jungfrau`` [EXT] (e4); BASELINE.md section
    2.3.  ``peds``/``gain``/``offs`` (3, nseg, 512, 1024) float32; ``gain None -> ones``,
    ``offs None -> zeros``.  ``jf_cm_eligible`` = which gain ranges take part in common mode
    (restated as {g0, g1}; a stated uncertainty, hence a parameter).
    """
    raw = np.asarray(raw)
    if gain is None and gfac is None:
        gain = np.ones_like(peds)
    if offs is None:
        offs = np.zeros_like(peds)
    if gfac is None:
        gfac = divide_protected(np.ones_like(gain), gain)

    gr0 = raw < BW1
    gr1 = (raw >= BW1) & (raw < BW2)
    gr2 = raw >= BW3

    arrf = np.array(raw & M14, dtype=np.float32)
    arrf[gr0] -= peds[0][gr0]
    arrf[gr1] -= peds[1][gr1]
    arrf[gr2] -= peds[2][gr2]

    factor = np.select((gr0, gr1, gr2), (gfac[0], gfac[1], gfac[2]), default=1)
    offset = np.select((gr0, gr1, gr2), (offs[0], offs[1], offs[2]), default=0)

    arrf -= offset

    if cmpars is not None:
        mode = int(cmpars[1])
        cormax = cmpars[2]
        npixmin = cmpars[3] if len(cmpars) > 3 else 10
        if mode > 0:
            arr1 = np.ones_like(raw, dtype=np.uint8)
            grs = (gr0, gr1, gr2)
            sel = tuple(grs[i] for i in jf_cm_eligible)
            grhm = np.select(sel, (arr1,) * len(sel), default=0)
            gmask = np.bitwise_and(grhm, mask) if mask is not None else grhm
            a3 = arrf.reshape((-1,) + arrf.shape[-2:])
            g3 = gmask.reshape(a3.shape)
            hrows = a3.shape[1] // 2      # 256
            for s in range(a3.shape[0]):
                if mode & 4:    # 2-d banks of (256, 64)
                    common_mode_2d_hsplit_nbanks(a3[s, :hrows, :], mask=g3[s, :hrows, :], nbanks=16, cormax=cormax, npix_min=npixmin)
                    common_mode_2d_hsplit_nbanks(a3[s, hrows:, :], mask=g3[s, hrows:, :], nbanks=16, cormax=cormax, npix_min=npixmin)
                if mode & 1:    # rows of 64 px per bank
                    common_mode_rows_hsplit_nbanks(a3[s], mask=g3[s], nbanks=16, cormax=cormax, npix_min=npixmin)
                if mode & 2:    # columns of 256 px per half
                    common_mode_cols(a3[s, :hrows, :], mask=g3[s, :hrows, :], cormax=cormax, npix_min=npixmin)
                    common_mode_cols(a3[s, hrows:, :], mask=g3[s, hrows:, :], cormax=cormax, npix_min=npixmin)

    return arrf * factor if mask is None else arrf * factor * mask


# ---------------------------------------------------------------------------------------------
# Detector/UtilsCommonMode.py
# ---------------------------------------------------------------------------------------------

def common_mode_2d(arr, mask=None, cormax=None, npix_min=10):
    """In-place median subtraction over a whole 2-d block, eligible pixels only.

    Upstream: ``Detector/UtilsCommonMode.py::common_mode_2d`` [EXT] (e5); BASELINE.md section
    2.4: skip if ``npix < npix_min``; skip if ``|cm| > cormax``; ``a[b] -= cm``.
    """
    if mask is None:
        cmode = np.median(arr)
        if cormax is not None and fabs(cmode) > cormax:
            return
        arr -= cmode
        return
    bmask = mask > 0
    npix = int(bmask.sum())
    if npix < npix_min:
        return
    cmode = np.median(arr[bmask])
    if cormax is not None and fabs(cmode) > cormax:
        return
    arr[bmask] -= cmode


def common_mode_rows(arr, mask=None, cormax=None, npix_min=10):
    """In-place per-row median subtraction.

    Upstream: ``UtilsCommonMode.py::common_mode_rows`` [EXT] (e5): applied only where
    ``npix > npix_min`` and ``|cm| < cormax`` (both strict).
    """
    rows, cols = arr.shape
    if mask is None:
        cmode = np.median(arr, axis=1)
        if cormax is not None:
            cmode = np.select((np.fabs(cmode) < cormax,), (cmode,), default=0)
        _, m2 = np.meshgrid(np.zeros(cols, dtype=np.int16), cmode)
        arr -= m2
        return
    marr = np.ma.array(arr, mask=mask < 1)
    cmode = np.ma.median(marr, axis=1)
    npix = mask.sum(axis=1)
    cmode = np.select((npix > npix_min,), (np.ma.getdata(cmode),), default=0)
    if cormax is not None:
        cmode = np.select((np.fabs(cmode) < cormax,), (cmode,), default=0)
    _, m2 = np.meshgrid(np.zeros(cols, dtype=np.int16), cmode)
    bmask = mask > 0
    arr[bmask] -= m2[bmask]


def common_mode_cols(arr, mask=None, cormax=None, npix_min=10):
    """In-place per-column median subtraction (same guards as rows, along axis 0)."""
    rows, cols = arr.shape
    if mask is None:
        cmode = np.median(arr, axis=0)
        if cormax is not None:
            cmode = np.select((np.fabs(cmode) < cormax,), (cmode,), default=0)
        m1, _ = np.meshgrid(cmode, np.zeros(rows, dtype=np.int16))
        arr -= m1
        return
    marr = np.ma.array(arr, mask=mask < 1)
    cmode = np.ma.median(marr, axis=0)
    npix = mask.sum(axis=0)
    cmode = np.select((npix > npix_min,), (np.ma.getdata(cmode),), default=0)
    if cormax is not None:
        cmode = np.select((np.fabs(cmode) < cormax,), (cmode,), default=0)
    m1, _ = np.meshgrid(cmode, np.zeros(rows, dtype=np.int16))
    bmask = mask > 0
    arr[bmask] -= m1[bmask]


def common_mode_2d_hsplit_nbanks(data, mask, nbanks=4, cormax=None, npix_min=10):
    """Split into ``nbanks`` column banks, ``common_mode_2d`` on each, stack back in place."""
    bdata = np.hsplit(data, nbanks)
    bmask = np.hsplit(mask, nbanks) if mask is not None else [None] * nbanks
    for b, m in zip(bdata, bmask):
        common_mode_2d(b, m, cormax, npix_min=npix_min)
    data[:] = np.hstack(bdata)[:]


def common_mode_rows_hsplit_nbanks(data, mask, nbanks=4, cormax=None, npix_min=10):
    """Split into ``nbanks`` column banks, ``common_mode_rows`` on each, stack back in place."""
    bdata = np.hsplit(data, nbanks)
    bmask = np.hsplit(mask, nbanks) if mask is not None else [None] * nbanks
    for b, m in zip(bdata, bmask):
        common_mode_rows(b, m, cormax, npix_min=npix_min)
    data[:] = np.hstack(bdata)[:]


# ---------------------------------------------------------------------------------------------
# PSCalib/GeometryAccess.py :: img_from_pixel_arrays ; AreaDetector.image
# ---------------------------------------------------------------------------------------------

def img_from_pixel_arrays(iX, iY, W=None, dtype=np.float32, vbase=0):
    """2-d image from pixel index arrays; duplicates are last-writer-wins in flat C order.

    Upstream: ``PSCalib/GeometryAccess.py::img_from_pixel_arrays`` [EXT] (e6); BASELINE.md
    section 2.5.
    """
    iX = np.asarray(iX)
    iY = np.asarray(iY)
    xsize = int(iX.max()) + 1
    ysize = int(iY.max()) + 1
    weights = np.ones_like(iX, dtype=dtype) if W is None else np.asarray(W)
    img = vbase * np.ones((xsize, ysize), dtype=dtype) if vbase else np.zeros((xsize, ysize), dtype=dtype)
    img[iX.ravel(), iY.ravel()] = weights.ravel()
    return img


# ---------------------------------------------------------------------------------------------
# psana_ray/producer.py:92-97 (this part IS in the reference)
# ---------------------------------------------------------------------------------------------

def producer_masks(data, bad_pixel_mask=None, manual_pixel_mask=None):
    """ref: psana_ray/producer.py:92-97 -- bad mask, then manual mask, then 2-D -> (1,H,W)."""
    if bad_pixel_mask is not None:
        data = np.where(bad_pixel_mask, data, 0)
    if manual_pixel_mask is not None:
        data = np.where(manual_pixel_mask, data, 0)
    if data.ndim == 2:
        data = data[None,]
    return data


# ---------------------------------------------------------------------------------------------
# Whole-event conveniences used by the tests and the CPU baseline
# ---------------------------------------------------------------------------------------------

def det_calib(kind, raw, consts, cmpars=None):
    """``det.calib(evt)`` for one event.  ``consts`` is the dict the synthetic generators make."""
    if kind == "epix10ka":
        return calib_epix10ka_any(raw, consts["cfg"], consts["peds"], consts["gain"],
                                  mask=consts.get("mask"), cmpars=cmpars, gfac=consts.get("gfac"))
    if kind == "jungfrau":
        return calib_jungfrau(raw, consts["peds"], consts.get("gain"), consts.get("offs"),
                              mask=consts.get("mask"), cmpars=cmpars, gfac=consts.get("gfac"))
    raise ValueError(kind)


def det_image(kind, raw, consts, iX, iY, cmpars=None):
    """``det.image(evt)`` = calib + geometry scatter (AreaDetector.image [EXT], e6)."""
    return img_from_pixel_arrays(iX, iY, W=det_calib(kind, raw, consts, cmpars))
