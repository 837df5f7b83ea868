"""tcgen05 3xTF32 accuracy probe (not a pytest file)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import dlco_b200
ctx = dlco_b200.Context(0)
dev = ctx.device
rng = np.random.default_rng(0)
for K in (64, 512, 4096):
    for kind in ("nonneg", "signed"):
        if kind == "nonneg":
            a = (rng.random((256, K)) / K * 2).astype(np.float32)          # rows sum to ~1 (filters)
            b = (rng.random((128, K)) * (rng.random((128, K)) < 0.25) * 4).astype(np.float32)
        else:
            a = rng.normal(0, 1, (256, K)).astype(np.float32)
            b = rng.normal(0, 1, (128, K)).astype(np.float32)
        want = a.astype(np.float64) @ b.astype(np.float64).T
        got = ctx.gemm3x_nt(torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)).cpu().numpy().astype(np.float64)
        f32 = (torch.from_numpy(a).to(dev) @ torch.from_numpy(b).to(dev).T).cpu().numpy().astype(np.float64)
        err = (got - want)
        scale = np.abs(want).max()
        print("K=%5d %-6s scale %.3g | tc: max %.3e mean(signed) %.3e rms %.3e | torch fp32: max %.3e mean %.3e rms %.3e" % (
            K, kind, scale, np.abs(err).max() / scale, err.mean() / scale, np.sqrt((err ** 2).mean()) / scale,
            np.abs(f32 - want).max() / scale, (f32 - want).mean() / scale, np.sqrt(((f32 - want) ** 2).mean()) / scale))
