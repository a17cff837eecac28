"""Property tests of the oracle's common-mode functions against an independent, deliberately naive definition
(sort the eligible values of a block / row / column in pure Python, take the two middle ones).  The reference has
no tests at this boundary, so besides the hand-made known-answer vectors the oracle is held to the
textbook definition on random shapes, masks, ties and guard values, and to a few invariances of a median."""
import numpy as np
from hypothesis import given, settings, strategies as st

from oracle import psana_calib as oc


def _median32(values):
    """NumPy's median of float32 values, spelled out: mean of the two middle order statistics, in float32."""
    v = sorted(float(x) for x in values)
    n = len(v)
    return np.float32((np.float32(v[(n - 1) // 2]) + np.float32(v[n // 2])) / np.float32(2))


def _naive_lines(a, m, cormax, npix_min, axis):
    out = a.copy()
    lines = a.shape[1 - axis]
    for j in range(lines):
        sl = (j, slice(None)) if axis == 1 else (slice(None), j)
        elig = m[sl] > 0
        n = int(elig.sum())
        if n == 0 or not n > npix_min:
            continue
        cm = _median32(a[sl][elig])
        if not abs(cm) < cormax:
            continue
        line = out[sl]
        line[elig] = a[sl][elig] - cm
    return out


def _naive_2d(a, m, cormax, npix_min):
    out = a.copy()
    elig = m > 0
    n = int(elig.sum())
    if n < npix_min or n == 0:
        return out
    cm = _median32(a[elig])
    if abs(cm) > cormax:
        return out
    out[elig] = a[elig] - cm
    return out


@st.composite
def blocks(draw):
    rows = draw(st.integers(1, 9)); cols = draw(st.integers(1, 9))
    # few distinct values (ties), halves and integers (even-count means), both signs
    vals = draw(st.lists(st.sampled_from([-7.5, -2.0, -0.25, 0.0, 0.5, 1.0, 1.5, 3.0, 40.0, 1000.0]), min_size=rows * cols, max_size=rows * cols))
    mask = draw(st.lists(st.integers(0, 1), min_size=rows * cols, max_size=rows * cols))
    a = np.array(vals, dtype=np.float32).reshape(rows, cols)
    m = np.array(mask, dtype=np.uint8).reshape(rows, cols)
    cormax = draw(st.sampled_from([0.5, 1.0, 3.0, 50.0, 1e9]))
    npix_min = draw(st.integers(0, 6))
    return a, m, cormax, npix_min


@settings(max_examples=300, deadline=None)
@given(blocks())
def test_common_mode_2d_is_the_median_of_the_eligible_pixels(b):
    a, m, cormax, npix_min = b
    got = a.copy()
    oc.common_mode_2d(got, m, cormax, npix_min)
    assert np.array_equal(got.view(np.uint32), _naive_2d(a, m, cormax, npix_min).view(np.uint32))
    assert np.array_equal(got[m == 0], a[m == 0])                      # ineligible pixels are never touched


@settings(max_examples=300, deadline=None)
@given(blocks())
def test_common_mode_rows_and_cols_follow_the_strict_guards(b):
    a, m, cormax, npix_min = b
    if not m.any():                 # np.ma.median of an all-masked array warns; the producer path never calls it that way either
        m = m.copy(); m[0, 0] = 1
    for axis, fn in ((1, oc.common_mode_rows), (0, oc.common_mode_cols)):
        got = a.copy()
        fn(got, m, cormax, npix_min)
        exp = _naive_lines(a, m, cormax, npix_min, axis)
        assert np.array_equal(got.view(np.uint32), exp.view(np.uint32)), (axis, a, m, cormax, npix_min)


@settings(max_examples=100, deadline=None)
@given(blocks(), st.randoms(use_true_random=False))
def test_median_does_not_depend_on_pixel_order(b, rnd):
    a, m, cormax, npix_min = b
    perm = list(range(a.size)); rnd.shuffle(perm)
    ap = a.ravel()[perm].reshape(a.shape); mp = m.ravel()[perm].reshape(a.shape)
    x, y = a.copy(), ap.copy()
    oc.common_mode_2d(x, m, cormax, npix_min); oc.common_mode_2d(y, mp, cormax, npix_min)
    assert np.array_equal(x.ravel()[perm].view(np.uint32), y.ravel().view(np.uint32))


def test_banks_split_is_per_bank_and_in_place():
    rng = np.random.default_rng(3)
    a = rng.normal(0, 3, (6, 16)).astype(np.float32); a[:, 8:] += 5          # second pair of banks sits 5 ADU higher
    m = np.ones_like(a, dtype=np.uint8)
    got = a.copy()
    oc.common_mode_2d_hsplit_nbanks(got, m, nbanks=4, cormax=100.0, npix_min=1)
    for b in range(4):
        blk = got[:, 4 * b:4 * b + 4]
        assert abs(float(np.median(blk))) < 1e-6                            # every bank is centred on its own median
        exp = a[:, 4 * b:4 * b + 4] - np.median(a[:, 4 * b:4 * b + 4])
        assert np.array_equal(blk.view(np.uint32), exp.astype(np.float32).view(np.uint32))
