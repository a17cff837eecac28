"""Loader for tests/golden/kat_v1.json and builders that embed the cases in real detector shapes."""
import json
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kat_v1.json")


def load():
    with open(_PATH) as fh:
        return json.load(fh)


def f32(bits):
    return np.array(bits, dtype=np.uint32).view(np.float32)


def same_bits(a, b):
    """Bit-exact float32 equality (so -0.0 != +0.0 and NaN payloads count)."""
    a = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    b = np.ascontiguousarray(b, dtype=np.float32).view(np.uint32)
    return a.shape == b.shape and bool(np.array_equal(a, b))


def assert_same_bits(a, b, what=""):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    bad = a.view(np.uint32) != b.view(np.uint32)
    if bad.any():
        i = np.argwhere(bad)[0]
        raise AssertionError(f"{what}: {int(bad.sum())} of {bad.size} values differ; first at {tuple(i)}: "
                             f"{a[tuple(i)]!r} vs {b[tuple(i)]!r}")
