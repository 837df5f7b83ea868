#!/usr/bin/env python
"""Launches for the ncu capture of the kernels added late in round 1: the sign-prox sweep iteration
(k_gemm3x_nt<true/false>, k_sign_*, k_quad_coldot, k_rows_planes) and the tensor-path comp-uprjdists
builder (k_desc<2>, k_pool_fulldist<1>, k_uprj_finish).  Run by tools/profile_new.sh."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import dlco_b200
    import bench as B
    import bench_paths as bp
    from dlco_b200 import learn
    ctx = dlco_b200.Context(0)
    dev = ctx.device
    # comp-uprjdists: 3 batches of 256 pairs x 68 filters
    full = B.synth_bank(torch, dev)
    sel = full[(torch.arange(68, device=dev) * 601 + 16000) % full.shape[0]].contiguous()
    del full
    bank = ctx.bank(sel)
    patches = B.synth_patches(torch, dev, 700, 3)
    idx = B.synth_pairs(torch, dev, 700, 3, 768)
    for _ in range(2):
        ctx.uprjdist_build(bank, patches, idx)
    torch.cuda.synchronize()
    # sign-prox sweep: 45 configurations at D = 544, 3 iterations
    D, n = 544, 100000
    ud, ulabels = bp.synth_uprj(torch, dev, n, D, 12)
    split = learn.split_train_val(ulabels)
    mus = np.repeat(np.array([0.0001, 0.0005, 0.001, 0.002, 0.003], np.float32), 9)
    gms = np.tile(np.array([0.025, 0.05, 0.1, 0.15, 0.2, 0.25, 0.5, 0.75, 1.0], np.float32), 5)
    S = learn.PjSweep(ctx, ud, ulabels, mus, gms, split=split, prox="sign")
    for _ in range(3):
        S.step()
    S.ranks()
    ctx.sync()
    print("prof_new: done, launches", ctx.launch_count())


if __name__ == "__main__":
    main()
