"""The coefficient table of the matrix-sign iteration (opencv-dlco_b200/csrc/pj_sign.cu, c_sign_coef) against
its generator (tools/sign_poly.py: one LP per step) and against the scalar properties the kernel relies on."""
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def table_from_source(design=0):
    """rows of c_sign_coef[design] in pj_sign.cu: design 0 = start interval [1e-5, 1] (10 quintics), design 1 =
    [1e-4, 1] (8 quintics); the last row of each is the closing Newton-Schulz cubic."""
    src = open(os.path.join(ROOT, "opencv-dlco_b200", "csrc", "pj_sign.cu")).read()
    body = src[src.index("c_sign_coef[2][SIGN_MAX_STEPS + 1][3] = {"):]
    body = body[:body.index("};")]
    rows = re.findall(r"\{\s*(-?[\d.]+)f,\s*(-?[\d.]+)f,\s*(-?[\d.]+)f\s*\}", body)
    t = np.array(rows, dtype=np.float64)
    assert t.shape == (24, 3)
    return t[12 * design:12 * design + 12]


DESIGNS = [(0, 1e-5, 10), (1, 1e-4, 8)]


@pytest.mark.parametrize("design,lo,nq", DESIGNS)
def test_table_shape_and_last_steps(design, lo, nq):
    t = table_from_source(design)
    assert np.allclose(t[11], [1.5, -0.5, 0.0])
    assert np.all(t[:nq, 0] > 1.8) and np.all(t[:nq, 1] < 0) and np.all(t[:nq, 2] > 0)
    # the host copy used by the fused-epilogue path holds the same numbers
    src = open(os.path.join(ROOT, "opencv-dlco_b200", "csrc", "pj_sign.cu")).read()
    body = src[src.index("h_sign_coef[2][SIGN_MAX_STEPS][3] = {"):]
    body = body[:body.index("};")]
    rows = np.array(re.findall(r"\{\s*(-?[\d.]+)f,\s*(-?[\d.]+)f,\s*(-?[\d.]+)f\s*\}", body), dtype=np.float64)
    assert rows.shape == (22, 3) and np.array_equal(rows[11 * design:11 * design + nq], t[:nq])


@pytest.mark.parametrize("design,lo,nq", DESIGNS)
def test_table_matches_generator(design, lo, nq):
    pytest.importorskip("scipy")
    from sign_check import make_seq
    seq = np.array(make_seq(lo, n_polish=0))
    t = table_from_source(design)
    assert len(seq) == nq
    assert np.abs(seq - t[:nq]).max() <= 2e-5, np.abs(seq - t[:nq]).max()


@pytest.mark.parametrize("design,lo,nq", DESIGNS)
def test_composition_maps_design_interval_to_one(design, lo, nq):
    """the quintics + the cubic polish, as the kernel applies them (rows 0..nq-1, then row 11): every x in
    [lo, 1] ends within 5e-6 of 1, smaller x stay in [0, 1] and keep their order (so an eigenvalue below the
    design interval contributes at most |lambda| / 2 of error, never a wrong sign)."""
    t = table_from_source(design)
    steps = list(t[:nq]) + [t[11]]
    x = np.concatenate([np.geomspace(1e-10, lo, 400, endpoint=False), np.geomspace(lo, 1.0, 4000)])
    y = x.copy()
    for a, b, c in steps:
        y = a * y + b * y**3 + c * y**5
        assert np.all(np.isfinite(y)) and y.max() < 2.1
    inside = x >= lo
    assert np.abs(1 - y[inside]).max() <= 5e-6
    below = ~inside
    assert np.all(y[below] > 0) and np.all(y[below] <= 1 + 5e-6)
    assert np.all(np.diff(y[below]) >= -2e-6)          # monotone up to the equi-ripple of the converged end
    # odd polynomials: the negative half mirrors
    yn = -x
    for a, b, c in steps:
        yn = a * yn + b * yn**3 + c * yn**5
    assert np.allclose(yn, -y)
