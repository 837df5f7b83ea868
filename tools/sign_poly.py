"""Offline design of the odd-quintic minimax sequence for the matrix sign iteration:
p_k(x) = a x + b x^3 + c x^5 minimising max |1 - p(x)| on the current interval [lo, hi] (LP)."""
import numpy as np
from scipy.optimize import linprog

def design(lo, hi, n=4000):
    x = np.unique(np.concatenate([np.geomspace(lo, hi, n), np.linspace(lo, hi, n)]))
    V = np.stack([x, x**3, x**5], 1)
    # variables a,b,c,t ; minimise t ; 1 - V@c <= t ; V@c - 1 <= t
    A = np.block([[-V, -np.ones((len(x), 1))], [V, -np.ones((len(x), 1))]])
    b = np.concatenate([-np.ones(len(x)), np.ones(len(x))])
    r = linprog([0, 0, 0, 1], A_ub=A, b_ub=b, bounds=[(None, None)] * 3 + [(0, None)], method="highs")
    a, bb, c, t = r.x
    return (a, bb, c), t

def sequence(l0, tol=3e-8, maxit=20):
    lo, hi = l0, 1.0
    seq = []
    for k in range(maxit):
        co, t = design(lo, hi)
        seq.append(co)
        # monotone on [0, lo]? image of [lo,hi] = [1-t, 1+t]
        xx = np.unique(np.concatenate([np.geomspace(lo, hi, 20001), np.linspace(lo, hi, 20001)]))
        yy = co[0] * xx + co[1] * xx**3 + co[2] * xx**5
        lo, hi = yy.min() / 1.02, yy.max() * 1.005          # safety margins against fp32 rounding
        print(k, co, "t=%.3e" % t, "image [%.3e, %.6f]" % (lo, hi))
        if t < tol:
            break
    return seq

if __name__ == "__main__":
    import sys
    l0 = float(sys.argv[1]) if len(sys.argv) > 1 else 1e-6
    seq = sequence(l0)
    print(len(seq), "steps")
    # scalar check across [1e-9, 1]
    x = np.geomspace(1e-9, 1, 2001)
    y = x.copy()
    for a, b, c in seq:
        y = a * y + b * y**3 + c * y**5
    print("max |1-y| on [l0,1]:", np.abs(1 - y[x >= l0]).max(), " y at 1e-7,1e-8:", y[np.searchsorted(x, 1e-7)], y[np.searchsorted(x,1e-8)], "max y overall", y.max())
    np.save("/tmp/sign_seq.npy", np.array(seq))
