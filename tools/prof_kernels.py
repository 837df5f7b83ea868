#!/usr/bin/env python
"""A handful of launches of each hot-path kernel of rows b/c/d at BASELINE sizes, for ncu:
  ncu --set full --clock-control none --import-source on -k regex:^k_ -o out python tools/prof_kernels.py
Each kernel is launched twice (the second launch is the one to read: warm caches, allocated scratch)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import dlco_b200
    import bench_paths as bp
    ctx = dlco_b200.Context(0)
    dev = ctx.device
    F, n = 5120, 500000
    dists, labels = bp.synth_dists(torch, dev, n, F, 11)
    w = torch.zeros(F, dtype=torch.float32, device=dev)
    w[torch.randperm(F, device=dev)[:12]] = 0.05
    os.environ["DLCO_GEMV_SPARSE"] = "0"
    for _ in range(2):
        ctx.pr_patchdist(dists, w)                                   # k_gemv_rows, 10.24 GB
    os.environ.pop("DLCO_GEMV_SPARSE")
    for _ in range(2):
        pd = ctx.pr_patchdist(dists, w)                              # gather path (12 non-zeros)
    d_lab = torch.from_numpy(labels[:100000]).to(dev)
    for _ in range(2):
        ctx.roc_stats(pd[:100000].contiguous(), d_lab)               # sort + k_roc
    from dlco_b200 import learn
    ip, ineg, npt, nnt = learn.split_train_val(labels)
    d_ip, d_in = torch.from_numpy(ip).to(dev), torch.from_numpy(ineg).to(dev)
    for n_cfg in (1, 256):
        mus = np.full(n_cfg, 0.035, np.float32)
        gam = np.linspace(0.05, 1.0, n_cfg).astype(np.float32) if n_cfg > 1 else np.array([0.25], np.float32)
        wv = torch.zeros((n_cfg, F), dtype=torch.float32, device=dev)
        df = torch.zeros_like(wv)
        s = 2215
        for k in range(2):
            s = ctx.pr_rda_run(dists, d_ip, npt, d_in, nnt, mus, gam, k * 5000, 5000, s, wv, df)   # k_pr_rda
    pos_val, neg_val = d_ip[npt:].contiguous(), d_in[nnt:].contiguous()
    for _ in range(2):
        ctx.pr_validate(dists, pos_val, neg_val, w, 0.035)           # gathered GEMV x2 + k_hinge_partial 50k x 50k
    del dists, wv, df
    torch.cuda.empty_cache()

    D = 544
    ud, ulabels = bp.synth_uprj(torch, dev, n, D, 12)
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    for r in (80, 128):
        W = torch.randn((D, D), generator=g, device=dev) * 0.3
        for _ in range(2):
            ctx.pj_patchdist(ud, W, r)                               # k_sqnorm_tc, 1.09 GB
    split = learn.split_train_val(ulabels)
    mus = np.repeat(np.array([0.0001, 0.0005, 0.001, 0.002, 0.003], np.float32), 9)
    gms = np.tile(np.array([0.025, 0.05, 0.1, 0.15, 0.2, 0.25, 0.5, 0.75, 1.0], np.float32), 5)
    S = learn.PjSweep(ctx, ud, ulabels, mus, gms, split=split)
    for _ in range(3):
        S.step()                                                     # grouped SYRK (45 configs) + grid kernels
    S.ranks()
    L = learn.PjLearner(ctx, ud, ulabels, 0.001, 0.1, split=split)
    for _ in range(3):
        L.step()                                                     # single-config step
    L.validate()
    ctx.sync()
    print("prof_kernels: done, launches", ctx.launch_count())


if __name__ == "__main__":
    main()
