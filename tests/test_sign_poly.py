"""The coefficient table of the matrix-sign iteration (opencv-dlco_b200/csrc/pj_sign.cu, c_sign_coef) against
its generator (tools/sign_poly.py: one LP per step) and against the scalar properties the kernel relies on."""
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def table_from_source():
    src = open(os.path.join(ROOT, "opencv-dlco_b200", "csrc", "pj_sign.cu")).read()
    body = src[src.index("c_sign_coef[SIGN_STEPS + 1][3] = {"):]
    body = body[:body.index("};")]
    rows = re.findall(r"\{\s*(-?[\d.]+)f,\s*(-?[\d.]+)f,\s*(-?[\d.]+)f\s*\}", body)
    return np.array(rows, dtype=np.float64)


def test_table_shape_and_last_steps():
    t = table_from_source()
    assert t.shape == (10, 3)
    assert np.allclose(t[8], [15 / 8, -10 / 8, 3 / 8]) and np.allclose(t[9], [1.5, -0.5, 0.0])


def test_table_matches_generator():
    pytest.importorskip("scipy")
    from sign_check import make_seq
    seq = np.array(make_seq(1e-4, n_polish=0))
    t = table_from_source()
    assert len(seq) == 8
    assert np.abs(seq - t[:8]).max() <= 2e-5, np.abs(seq - t[:8]).max()


def test_composition_maps_design_interval_to_one():
    """8 minimax quintics + the cubic polish, as the kernel applies them (steps 0..7, then row 9): every
    x in [1e-4, 1] ends within 5e-6 of 1, smaller x stay in [0, 1] and keep their order (so an
    eigenvalue below the design interval contributes at most |lambda| / 2 of error, never a wrong sign)."""
    t = table_from_source()
    x = np.concatenate([np.geomspace(1e-9, 1e-4, 400, endpoint=False), np.geomspace(1e-4, 1.0, 4000)])
    y = x.copy()
    for a, b, c in list(t[:8]) + [t[9]]:
        y = a * y + b * y**3 + c * y**5
        assert np.all(np.isfinite(y)) and y.max() < 2.1
    inside = x >= 1e-4
    assert np.abs(1 - y[inside]).max() <= 5e-6
    below = ~inside
    assert np.all(y[below] > 0) and np.all(y[below] <= 1 + 5e-6)
    assert np.all(np.diff(y[below]) >= -2e-6)          # monotone up to the equi-ripple of the converged end
    # odd polynomials: the negative half mirrors
    yn = -x
    for a, b, c in list(t[:8]) + [t[9]]:
        yn = a * yn + b * yn**3 + c * yn**5
    assert np.allclose(yn, -y)
