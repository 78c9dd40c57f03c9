"""Shared helpers for the tests (test infrastructure)."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
MODELS = os.path.join(ROOT, "assets", "models")


def with_planes(sc, oracle):
    """Fill planes/squares of a synthetic scene with the oracle's Triangle-constructor restatement."""
    if sc.get("planes") is not None:
        return sc
    planes, squares = oracle.triangle_setup(sc["verts"])
    out = dict(sc)
    out["planes"], out["squares"] = planes, squares
    return out


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def assert_bits_equal(a, b, what=""):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    bad = np.flatnonzero(bits(a).ravel() != bits(b).ravel())
    assert bad.size == 0, f"{what}: {bad.size} of {a.size} differ, first at {bad[:5]}: {a.ravel()[bad[:5]]} vs {b.ravel()[bad[:5]]}"
