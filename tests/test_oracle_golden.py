"""The NumPy oracle against the hand-derived known-answer vectors (tests/golden/kat_v1.json).

Parity is unpinned against psana itself (not installable offline); these
vectors pin the oracle to the definitions in SURVEY.md 8a-EXT / BASELINE.md section 2.
"""
import numpy as np
import pytest

import kat
from oracle import psana_calib as oc

G = kat.load()


def test_epix10ka_decode_table_all_64_cbits():
    for row in G["epix10ka_decode"]:
        raw = np.array([[(row["d14"] << 14) | 0x155]], dtype=np.uint16)
        cfg = np.array([[row["cfg"]]], dtype=np.uint8)
        assert int(oc.cbits_epix10ka(raw, cfg)[0, 0]) == (row["cbits"] & 0o74)
        assert int(oc.gain_mode_index_epix10ka(raw, cfg)[0, 0]) == row["mode"], row
        gm = oc.gain_maps_epix10ka_any(raw, cfg)
        assert sum(int(g[0, 0]) for g in gm) == (0 if row["mode"] == 255 else 1)
        elig = bool(gm[0][0, 0] | gm[1][0, 0] | gm[3][0, 0] | gm[4][0, 0])
        assert elig == row["cm_eligible"], row


def test_jungfrau_decode_table():
    for row in G["jungfrau_decode"]:
        for v in (0, 1, 0x3FFF):
            raw = np.array([[(row["gain_bits"] << 14) | v]], dtype=np.uint16)
            assert int(oc.gain_mode_index_jungfrau(raw)[0, 0]) == row["mode"]


def test_divide_protected():
    gain = kat.f32(G["divide_protected"]["gain"])
    gfac = oc.divide_protected(np.ones_like(gain), gain)
    assert gfac.dtype == np.float32
    kat.assert_same_bits(gfac, kat.f32(G["divide_protected"]["gfac"]))


def test_epix10ka_single_pixels():
    for case in G["epix10ka_pixels"]:
        shape = (1, 2, 8)
        raw = np.full(shape, case["raw"], dtype=np.uint16)
        cfg = np.full(shape, case["cfg"], dtype=np.uint8)
        peds = np.full((7,) + shape, 12345.0, dtype=np.float32)   # wrong rows must not be picked
        gain = np.full((7,) + shape, 0.001, dtype=np.float32)
        if case["mode"] != 255:
            peds[case["mode"]] = kat.f32(case["ped"])
            gain[case["mode"]] = kat.f32(case["gain"])
        else:
            peds[:] = kat.f32(case["ped"]); gain[:] = kat.f32(case["gain"])
        mask = np.full(shape, case["mask"], dtype=np.uint8)
        out = oc.calib_epix10ka_any(raw, cfg, peds, gain, mask)
        assert out.dtype == np.float32
        assert int(oc.gain_mode_index_epix10ka(raw, cfg)[0, 0, 0]) == case["mode"]
        kat.assert_same_bits(out, np.full(shape, kat.f32(case["expect"]), dtype=np.float32), str(case))


def test_jungfrau_single_pixels():
    for case in G["jungfrau_pixels"]:
        shape = (1, 2, 16)
        raw = np.full(shape, case["raw"], dtype=np.uint16)
        peds = np.full((3,) + shape, 12345.0, dtype=np.float32)
        gain = np.full((3,) + shape, 0.001, dtype=np.float32)
        offs = np.full((3,) + shape, -777.0, dtype=np.float32)
        if case["mode"] != 255:
            peds[case["mode"]] = kat.f32(case["ped"]); gain[case["mode"]] = kat.f32(case["gain"])
            offs[case["mode"]] = kat.f32(case["offs"])
        else:
            peds[:] = kat.f32(case["ped"]); gain[:] = kat.f32(case["gain"]); offs[:] = kat.f32(case["offs"])
        mask = np.full(shape, case["mask"], dtype=np.uint8)
        out = oc.calib_jungfrau(raw, peds, gain, offs, mask)
        assert out.dtype == np.float32
        kat.assert_same_bits(out, np.full(shape, kat.f32(case["expect"]), dtype=np.float32), str(case))


def test_median_cases_2d_rows_cols():
    for case in G["median_cases"]:
        v = np.array(case["values"], dtype=np.float32)
        e = np.array(case["eligible"], dtype=np.uint8)
        med = kat.f32(case["median"])
        n = v.size
        a = v.reshape(1, n).copy(); oc.common_mode_2d(a, e.reshape(1, n), cormax=1e9, npix_min=0)
        kat.assert_same_bits(a.ravel()[e > 0], (v - med)[e > 0], case["name"])
        kat.assert_same_bits(a.ravel()[e == 0], v[e == 0], case["name"])
        a = v.reshape(1, n).copy(); oc.common_mode_rows(a, e.reshape(1, n), cormax=1e9, npix_min=0)
        kat.assert_same_bits(a.ravel()[e > 0], (v - med)[e > 0], case["name"])
        a = v.reshape(n, 1).copy(); oc.common_mode_cols(a, e.reshape(n, 1), cormax=1e9, npix_min=0)
        kat.assert_same_bits(a.ravel()[e > 0], (v - med)[e > 0], case["name"])


def test_guard_semantics_differ_between_2d_and_lines():
    gc = G["guard_cases"]
    cormax = float(kat.f32(gc["cormax"])); npix_min = gc["npix_min"]
    for case in gc["cases"]:
        v = np.array(case["values"], dtype=np.float32); e = np.array(case["eligible"], dtype=np.uint8); n = v.size
        a = v.reshape(1, n).copy(); oc.common_mode_2d(a, e.reshape(1, n), cormax=cormax, npix_min=npix_min)
        kat.assert_same_bits(a.ravel(), np.array(case["after_2d"], dtype=np.float32), case["name"] + " 2d")
        a = v.reshape(1, n).copy(); oc.common_mode_rows(a, e.reshape(1, n), cormax=cormax, npix_min=npix_min)
        kat.assert_same_bits(a.ravel(), np.array(case["after_line"], dtype=np.float32), case["name"] + " rows")
        a = v.reshape(n, 1).copy(); oc.common_mode_cols(a, e.reshape(n, 1), cormax=cormax, npix_min=npix_min)
        kat.assert_same_bits(a.ravel(), np.array(case["after_line"], dtype=np.float32), case["name"] + " cols")


def test_scatter_last_writer_wins_and_gaps_zero():
    s = G["scatter"]
    img = oc.img_from_pixel_arrays(np.array(s["iX"]), np.array(s["iY"]), np.array(s["W"], dtype=np.float32))
    assert img.dtype == np.float32
    kat.assert_same_bits(img, np.array(s["img"], dtype=np.float32))


@pytest.mark.parametrize("idx", range(len(G["tiny_pipelines"])))
def test_tiny_pipelines_all_common_mode_words(idx):
    t = G["tiny_pipelines"][idx]
    shape = tuple(t["shape"])
    raw = np.array(t["raw"], dtype=np.uint16).reshape(shape)
    mask = np.array(t["mask"], dtype=np.uint8).reshape(shape)
    nm = 7 if t["kind"] == "epix10ka" else 3
    peds = kat.f32(t["peds"]).reshape((nm,) + shape)
    gain = kat.f32(t["gain"]).reshape((nm,) + shape)
    cormax = float(kat.f32(t["cormax"]))
    for m, exp in t["expect_by_mode"].items():
        cmpars = (t["alg"], int(m), cormax, t["npix_min"])
        if t["kind"] == "epix10ka":
            cfg = np.array(t["cfg"], dtype=np.uint8).reshape(shape)
            out = oc.calib_epix10ka_any(raw, cfg, peds, gain, mask, cmpars)
        else:
            offs = kat.f32(t["offs"]).reshape((3,) + shape)
            out = oc.calib_jungfrau(raw, peds, gain, offs, mask, cmpars)
        kat.assert_same_bits(out, kat.f32(exp).reshape(shape), f"{t['kind']} mode {m}")
    if t["kind"] == "epix10ka":
        out = oc.calib_epix10ka_any(raw, cfg, peds, gain, None, (7, 7, cormax, t["npix_min"]))
        kat.assert_same_bits(out, kat.f32(t["expect_nomask_mode7"]).reshape(shape), "nomask")
    else:
        out = oc.calib_jungfrau(raw, peds, None, None, mask, (7, 5, cormax, t["npix_min"]))
        kat.assert_same_bits(out, kat.f32(t["expect_nogain_nooffs_mode5"]).reshape(shape), "nogain")


def test_producer_masks_order_and_ndim():
    # ref: psana_ray/producer.py:92-97
    d = np.arange(12, dtype=np.float32).reshape(3, 4) - 5
    bad = (np.arange(12).reshape(3, 4) % 2).astype(np.uint8)
    man = (np.arange(12).reshape(3, 4) % 3 > 0)
    out = oc.producer_masks(d, bad, man)
    assert out.shape == (1, 3, 4)
    exp = np.where(man, np.where(bad, d, 0), 0)[None,]
    kat.assert_same_bits(out, exp)
    d3 = np.ones((2, 3, 4), dtype=np.float32)
    assert oc.producer_masks(d3).shape == (2, 3, 4)
