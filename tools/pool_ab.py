"""A/B of a pooling-kernel variant selected by an environment variable read once per process (DLCO_POOL_SPLIT=1|2,
DLCO_POOL_IMPL, ...): run once per setting; each run saves its distances (compare with numpy.array_equal: a variant
must not change a bit) and prints the time of 16 384 pairs through dlco_fulldist_build.
    DLCO_POOL_SPLIT=1 python tools/pool_ab.py a.npy;  python tools/pool_ab.py b.npy
The first process on a cold GPU measures ~5 % slower than the following ones: interleave the settings (a b a b)."""
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = importlib.import_module("opencv-dlco_b200")


def main():
    ctx = B.Context(0)
    bank = ctx.bank_analytic(B.ring_params())
    rng = np.random.default_rng(3)
    quick = len(sys.argv) > 2 and sys.argv[2] == "--quick"        # distances of 600 pairs only (tests/test_gpu_kernels.py)
    n_patches, n_pairs = 4096, (600 if quick else 8192 * 2 + 300)
    patches = torch.from_numpy(rng.integers(0, 256, (n_patches, 64, 64), dtype=np.uint8)).to(ctx.device)
    idx = np.zeros((n_pairs, 4), np.int32)
    idx[:, 0] = rng.integers(0, n_patches, n_pairs)
    idx[:, 2] = rng.integers(0, n_patches, n_pairs)
    d_i = torch.from_numpy(idx).to(ctx.device)
    out = ctx.fulldist_build(bank, patches, d_i)
    torch.cuda.synchronize()
    np.save(sys.argv[1], out.cpu().numpy())
    if quick:
        return
    ts = []
    for _ in range(5):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.fulldist_build(bank, patches, d_i, 0, 8192 * 2)
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    env = " ".join("%s=%s" % kv for kv in sorted(os.environ.items()) if kv[0].startswith("DLCO_"))
    print("%s: %.3f ms per 16384 pairs (min of 5), %.1fk pairs/s" % (env or "defaults", min(ts) * 1e3, 16.384 / min(ts)))


if __name__ == "__main__":
    main()
