import sys, numpy as np
import os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from sign_poly import design
f32 = np.float32

def make_seq(l0, n_polish=1):
    lo, hi, seq = l0, 1.0, []
    for k in range(30):
        co, t = design(lo, hi)
        xx = np.unique(np.concatenate([np.geomspace(lo, hi, 20001), np.linspace(lo, hi, 20001)]))
        yy = co[0] * xx + co[1] * xx**3 + co[2] * xx**5
        lo, hi = yy.min() / 1.02, yy.max() * 1.005
        seq.append(co)
        if t < 2e-3: break
    seq += [(15 / 8, -10 / 8, 3 / 8)] * n_polish
    return seq

def psd_part_sign(A, seq, dt=f32):
    A = A.astype(dt)
    alpha = dt(np.sqrt((A.astype(np.float64) ** 2).sum()))
    X = (A / alpha).astype(dt)
    I = np.eye(len(A), dtype=dt)
    for a, b, c in seq:
        Y = (X @ X).astype(dt)
        Z = (dt(a) * I + dt(b) * Y + dt(c) * (Y @ Y)).astype(dt)
        X = (X @ Z).astype(dt)
        X = (dt(0.5) * (X + X.T)).astype(dt)
    Ap = (dt(0.5) * (A + A @ X)).astype(dt)
    Ap = dt(0.5) * (Ap + Ap.T)
    return Ap, X

def psd_part_eig(A):
    w, V = np.linalg.eigh(A.astype(np.float64))
    return (V * np.maximum(w, 0)) @ V.T, w

if __name__ == "__main__":
    seq = make_seq(float(sys.argv[1]) if len(sys.argv) > 1 else 1e-5)
    print(len(seq), "steps")
    rng = np.random.default_rng(0)
    for D in (160, 544):
        for trial in range(3):
            # pj-learn-like: A = -c (dfAvg + mu I), dfAvg = X^T diag(c) X / b^2
            Xb = rng.standard_normal((400 * (trial + 1), D)) * 0.1
            c = rng.integers(-200, 200, size=len(Xb)).astype(np.float64)
            df = (Xb.T * c) @ Xb / 40000
            A = -(np.sqrt(trial * 50 + 1.0) / 0.1) * (df + 0.001 * np.eye(D))
            A = 0.5 * (A + A.T)
            ref, w = psd_part_eig(A)
            for dt in (np.float64, np.float32):
                Ap, S = psd_part_sign(A, seq, dt)
                err = np.abs(Ap - ref).max() / np.abs(ref).max()
                rank = (D + np.trace(S.astype(np.float64))) / 2
                print("D=%d trial %d %s: relerr %.2e  rank(sign) %.3f  rank(eig) %d  min|w|/||A||F %.1e" % (
                    D, trial, np.dtype(dt).name, err, rank, (w > 0).sum(), np.abs(w).min() / np.sqrt((A**2).sum())))
