#!/usr/bin/env python
"""bench_paths.py -- device timings of the hot-path rows other than the full-distance build
(SURVEY 8 rows b, c, d): one JSON object per sub-path with its roofline fraction.

  python bench_paths.py [--n-train 500000] [--n-test 100000] [--quick]

Used by bench.py (`secondary` block) and on its own for the ncu captures under profiles/.
All inputs are synthetic and resident in HBM; timings are CUDA events on the launching stream,
after warm-up.  The GPU legs never touch oracle/; `cpu_leg` (the `cpu_baseline` of these rows,
--cpu) times the oracle's CPU port of each row on a bounded sample, on the box's host cores.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def _peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), float(p["bf16_tflops_sustained"]), "measured"
    except (OSError, ValueError, KeyError):
        return 6500.0, 1400.0, "fallback (B200_PROFILING.md)"


def timeit(torch, fn, warm=3, rep=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(rep):
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.mean(ts)), float(np.min(ts))


def synth_dists(torch, dev, n, f, seed, sep=0.3):
    """Distance-like matrix: non-negative, positives (even rows) smaller on average."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    out = torch.empty((n, f), dtype=torch.float32, device=dev)
    step = 65536
    for a in range(0, n, step):
        b = min(n, a + step)
        x = torch.rand((b - a, f), generator=g, device=dev)
        lab = ((torch.arange(a, b, device=dev) % 2) == 0).float()[:, None]
        out[a:b] = x * (1.0 + sep * (1.0 - lab))
    labels = ((np.arange(n) % 2) == 0).astype(np.uint8)
    return out, labels


def synth_uprj(torch, dev, n, d, seed, sep=0.6):
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    lab = ((torch.arange(n, device=dev) % 2) == 0).float()[:, None]
    x = torch.randn((n, d), generator=g, device=dev) * (0.08 + 0.08 * sep * (1.0 - lab))
    labels = ((np.arange(n) % 2) == 0).astype(np.uint8)
    return x.contiguous(), labels


def run(ctx, n_train=500000, n_test=100000, quick=False, D=544, n_pj_cfg=45, n_pr_cfg=256):
    import torch
    import dlco_b200
    from dlco_b200 import learn
    dev = ctx.device
    hbm, bf16, src = _peaks()
    tf32x3 = bf16 / 2.0 / 3.0                    # fp32-equivalent rate of a 3xTF32 product
    res = {"peaks": {"hbm_gbs": hbm, "bf16_tflops_sustained": bf16, "source": src}}
    F = 5120

    # ---- (c)/(d1): GEMV over the full-distance matrix (ComputePRStats, src/misc.cpp:226) -----
    dists, labels = synth_dists(torch, dev, n_train, F, 11)
    w = torch.zeros(F, dtype=torch.float32, device=dev)
    w[torch.randperm(F, device=dev)[:12]] = 0.05
    os.environ["DLCO_GEMV_SPARSE"] = "0"         # dense streaming path: every byte of the matrix is read
    mean_ms, best_ms = timeit(torch, lambda: ctx.pr_patchdist(dists, w), rep=5 if quick else 10)
    os.environ.pop("DLCO_GEMV_SPARSE")
    gb = n_train * F * 4 / 1e9
    res["pr_patchdist_gemv"] = {"rows": n_train, "cols": F, "ms": mean_ms, "best_ms": best_ms,
                                "achieved_gbs": gb / (mean_ms * 1e-3), "frac_hbm": gb / (mean_ms * 1e-3) / hbm,
                                "bytes": n_train * F * 4}
    # the same call as shipped: w has 12 non-zeros (the reference's models have 4-21), only their
    # columns are gathered -- one 32-byte sector per non-zero and row
    mean_ms, best_ms = timeit(torch, lambda: ctx.pr_patchdist(dists, w), rep=5 if quick else 10)
    sect = n_train * 12 * 32 / 1e9
    res["pr_patchdist_sparse"] = {"rows": n_train, "cols": F, "nnz": 12, "ms": mean_ms, "best_ms": best_ms,
                                  "sector_gbs": sect / (mean_ms * 1e-3), "frac_hbm_sectors": sect / (mean_ms * 1e-3) / hbm,
                                  "dense_equiv_gbs": gb / (mean_ms * 1e-3)}

    # ---- (d1) ComputePRStats end to end from a HOST matrix: n_test x 5120 pinned floats in, FPR95 / AUC out --------
    import time
    h_d = dists[:n_test].cpu().pin_memory()
    d_d = torch.empty_like(dists[:n_test])
    d_lab_t = torch.from_numpy(labels[:n_test]).to(dev)

    def stats_e2e():
        d_d.copy_(h_d, non_blocking=True)
        return ctx.roc_stats(ctx.pr_patchdist(d_d, w), d_lab_t)
    stats_e2e()
    t0 = time.perf_counter()
    for _ in range(3):
        stats_e2e()
    dt = (time.perf_counter() - t0) / 3
    res["pr_stats_e2e_host"] = {"rows": n_test, "cols": F, "ms": dt * 1e3, "models_per_s": 1.0 / dt,
                                "h2d_bytes": int(h_d.numel() * 4), "d2h_bytes": 12,
                                "note": "PCIe-bound: the matrix upload is the whole cost; a -prj list amortises it"}
    del h_d, d_d

    # ---- (d1) ROC / FPR95 on n_test pairs -----------------------------------------------
    pd = ctx.pr_patchdist(dists[:n_test], w)
    d_lab = torch.from_numpy(labels[:n_test]).to(dev)
    mean_ms, best_ms = timeit(torch, lambda: ctx.roc_stats(pd, d_lab))
    res["roc_stats"] = {"n": n_test, "ms": mean_ms, "best_ms": best_ms}

    # ---- (c1) pr-learn RDA chain -----------------------------------------------------------
    ip, ineg, npt, nnt = learn.split_train_val(labels)
    d_ip, d_in = torch.from_numpy(ip).to(dev), torch.from_numpy(ineg).to(dev)
    for n_cfg in (1, n_pr_cfg):
        mus = np.full(n_cfg, 0.035, np.float32)
        gammas = np.linspace(0.05, 1.0, n_cfg).astype(np.float32) if n_cfg > 1 else np.array([0.25], np.float32)
        wv = torch.zeros((n_cfg, F), dtype=torch.float32, device=dev)
        df = torch.zeros_like(wv)
        steps = 20000 if quick else 200000
        st = {"s": 2215, "t": 0}

        def chain():
            st["s"] = ctx.pr_rda_run(dists, d_ip, npt, d_in, nnt, mus, gammas, st["t"], steps, st["s"], wv, df)
            st["t"] += steps
        mean_ms, best_ms = timeit(torch, chain, warm=1, rep=3)
        res["pr_rda_%dcfg" % n_cfg] = {"steps": steps, "ms": mean_ms,
                                       "steps_per_s_per_cfg": steps / (mean_ms * 1e-3),
                                       "cfg_steps_per_s": n_cfg * steps / (mean_ms * 1e-3),
                                       "algorithmic_bytes_per_step": 2 * F * 4,
                                       "achieved_gbs": n_cfg * steps * 2 * F * 4 / (mean_ms * 1e-3) / 1e9,
                                       "nnz": int((wv[0] != 0).sum().item())}

    # ---- (c2) pr validation: 2 gathered GEMVs + 50k x 50k hinge ------------------------------
    pos_val, neg_val = d_ip[npt:].contiguous(), d_in[nnt:].contiguous()
    mean_ms, best_ms = timeit(torch, lambda: ctx.pr_validate(dists, pos_val, neg_val, w, 0.035))
    vb = (pos_val.numel() + neg_val.numel()) * F * 4
    res["pr_validate"] = {"n_pos": pos_val.numel(), "n_neg": neg_val.numel(), "ms": mean_ms,
                          "achieved_gbs": vb / (mean_ms * 1e-3) / 1e9, "frac_hbm": vb / (mean_ms * 1e-3) / 1e9 / hbm}
    a = torch.rand(50000, device=dev)
    b = torch.rand(50000, device=dev) + 0.3
    mean_ms, best_ms = timeit(torch, lambda: ctx.hinge_pairsum(a, b))
    res["hinge_pairsum_50k_x_50k"] = {"ms": mean_ms, "gterms_per_s": 2.5e9 / (mean_ms * 1e-3) / 1e9}
    del dists, wv, df
    torch.cuda.empty_cache()

    # ---- (b) pj-learn inner step at D ---------------------------------------------------------
    ud, ulabels = synth_uprj(torch, dev, n_train, D, 12)
    split = learn.split_train_val(ulabels)
    L = learn.PjLearner(ctx, ud, ulabels, 0.001, 0.1, split=split)
    for _ in range(30):
        L.step()                                 # let the rank settle
    torch.cuda.synchronize()
    rp, rn = L.sample_batch()
    d_rp, d_rn = torch.from_numpy(rp).to(dev), torch.from_numpy(rn).to(dev)
    parts = {}
    parts["grad_ms"] = timeit(torch, lambda: ctx.pj_rda_grad(ud, d_rp, d_rn, L.W, L.rank, out=L.dLoss))[0]
    dfa = L.dfavg.clone()
    parts["update_ms"] = timeit(torch, lambda: ctx.pj_rda_update(40, 200, 0.001, 0.1, L.dLoss, dfa, L.A))[0]
    A0 = L.A.clone()
    parts["eig_ms"] = timeit(torch, lambda: ctx.pj_eig(A0, L.evals, L.evec), rep=5)[0]
    Ax, Wx = L.A.clone(), L.W.clone()
    parts["prox_ms"] = timeit(torch, lambda: ctx.pj_prox_apply(L.evals, L.evec, Ax, Wx))[0]
    n_it = 20 if quick else 100
    mean_ms, best_ms = timeit(torch, lambda: [L.step() for _ in range(n_it)], warm=1, rep=3)
    ctx.profile(True)
    for _ in range(20):
        ctx.pj_rda_grad(ud, d_rp, d_rn, L.W, L.rank, out=L.dLoss)
    torch.cuda.synchronize()
    syrk_ms, syrk_n = ctx.profile_read(2)
    ctx.profile(False)
    syrk_flop = 2.0 * 400 * D * D
    syrk_avg = syrk_ms / max(syrk_n, 1)
    res["pj_step"] = dict(parts, D=D, rank=L.rank, iters_per_s=n_it / (mean_ms * 1e-3),
                          syrk_ms=syrk_avg, syrk_tflops_fp32_equiv=syrk_flop / (syrk_avg * 1e-3) / 1e12 if syrk_n else None,
                          syrk_frac_of_3xtf32_peak=(syrk_flop / (syrk_avg * 1e-3) / 1e12 / tf32x3) if syrk_n else None)

    # grouped SYRK over the configurations of a sweep sharing one batch
    Wg = L.W[None, :max(L.rank, 1)].repeat(n_pj_cfg, 1, 1).contiguous()
    ranks = [L.rank] * n_pj_cfg
    outg = torch.empty((n_pj_cfg, D, D), dtype=torch.float32, device=dev)
    ctx.profile(True)
    mean_ms, _ = timeit(torch, lambda: ctx.pj_rda_grad_grid(ud, d_rp, d_rn, Wg, ranks, out=outg))
    syrk_ms, syrk_n = ctx.profile_read(2)
    ctx.profile(False)
    syrk_avg = syrk_ms / max(syrk_n, 1)
    res["pj_grad_grid"] = {"n_cfg": n_pj_cfg, "ms": mean_ms, "syrk_ms": syrk_avg,
                           "syrk_tflops_fp32_equiv": n_pj_cfg * syrk_flop / (syrk_avg * 1e-3) / 1e12 if syrk_n else None,
                           "syrk_frac_of_3xtf32_peak": (n_pj_cfg * syrk_flop / (syrk_avg * 1e-3) / 1e12 / tf32x3) if syrk_n else None,
                           "cfg_grads_per_s": n_pj_cfg / (mean_ms * 1e-3)}

    # ---- (b) the whole lock-step sweep iteration: 45 (mu, gamma) configurations of one data set ----
    mus = np.repeat(np.array([0.0001, 0.0005, 0.001, 0.002, 0.003], np.float32), 9)[:n_pj_cfg]
    gms = np.tile(np.array([0.025, 0.05, 0.1, 0.15, 0.2, 0.25, 0.5, 0.75, 1.0], np.float32), 5)[:n_pj_cfg]
    S = learn.PjSweep(ctx, ud, ulabels, mus, gms, split=split)
    for _ in range(10):
        S.step()
    S.ranks()
    n_it = 10 if quick else 40
    mean_ms, best_ms = timeit(torch, lambda: [S.step() for _ in range(n_it)], warm=1, rep=3)
    rk = S.ranks()
    res["pj_sweep_step"] = {"n_cfg": int(len(mus)), "D": D, "ms_per_iter": mean_ms / n_it,
                            "cfg_iters_per_s": len(mus) * n_it / (mean_ms * 1e-3),
                            "rank_min": int(rk.min()), "rank_max": int(rk.max()),
                            "note": "sample + margins + hinge + grouped SYRK + update + batched eig + prox, no host sync"}
    del S
    torch.cuda.empty_cache()
    # the same sweep with the prox's projection by the tensor-core matrix-sign iteration (no eig per iteration)
    for tag, mm, gg in (("pj_sweep_step_sign", mus, gms), ("pj_step_sign_1cfg", mus[2:3], gms[2:3])):
        S = learn.PjSweep(ctx, ud, ulabels, mm, gg, split=split, prox="sign")
        for _ in range(5):
            S.step()
        ctx.profile(True)
        mean_ms, best_ms = timeit(torch, lambda: [S.step() for _ in range(n_it)], warm=1, rep=3)
        gemm_ms, gemm_n = ctx.profile_read(2)
        ctx.profile(False)
        mean_ms, best_ms = timeit(torch, lambda: [S.step() for _ in range(n_it)], warm=1, rep=3)
        rk = S.ranks()
        # 29 GEMMs per iteration: 26 (sign) + 1 (A sign(A)) at K = D, margins at K = D, SYRK at K = 2 * batch
        flop = len(mm) * (2.0 * D * D * D * 27 + 2.0 * 400 * D * D * 2)
        res[tag] = {"n_cfg": int(len(mm)), "D": D, "ms_per_iter": mean_ms / n_it,
                    "cfg_iters_per_s": len(mm) * n_it / (mean_ms * 1e-3),
                    "rank_min": int(rk.min()), "rank_max": int(rk.max()),
                    "gemm_launches_per_iter": gemm_n / (4.0 * n_it), "gemm_ms_per_iter_profiled": gemm_ms / (4.0 * n_it),
                    "gemm_tflops_fp32_equiv": flop / (mean_ms / n_it * 1e-3) / 1e12,
                    "gemm_frac_of_3xtf32_peak": flop / (mean_ms / n_it * 1e-3) / 1e12 / tf32x3,
                    "note": "sample + x^T A x margins + hinge + grouped SYRK + update + 27-GEMM matrix-sign projection, no host sync"}
        del S
        torch.cuda.empty_cache()

    # ---- (d2) ComputePJStats distances + (b6) pj validation -----------------------------------
    mean_ms, best_ms = timeit(torch, lambda: ctx.pj_patchdist(ud, L.W, L.rank))
    res["pj_patchdist"] = {"rows": n_train, "D": D, "rank": L.rank, "ms": mean_ms,
                           "achieved_gbs": n_train * D * 4 / (mean_ms * 1e-3) / 1e9,
                           "frac_hbm": n_train * D * 4 / (mean_ms * 1e-3) / 1e9 / hbm,
                           "tflops": 2.0 * n_train * D * max(L.rank, 1) / (mean_ms * 1e-3) / 1e12}
    # the ranks the reference's exported models have (48 / 64 / 80 / 120, workspace/11-opencv-export.sh:3-6)
    for r in (48, 80, 128, 256):
        r = min(r, max(L.rank, 1))
        Wr = L.W[L.rank - r:L.rank].contiguous() if L.rank >= r else L.W
        mean_ms, best_ms = timeit(torch, lambda: ctx.pj_patchdist(ud, Wr, r))
        res["pj_patchdist_rank%d" % r] = {"rows": n_train, "D": D, "rank": r, "ms": mean_ms, "best_ms": best_ms,
                                           "achieved_gbs": n_train * D * 4 / (mean_ms * 1e-3) / 1e9,
                                           "frac_hbm": n_train * D * 4 / (mean_ms * 1e-3) / 1e9 / hbm,
                                           "tflops": 2.0 * n_train * D * r / (mean_ms * 1e-3) / 1e12}
    mean_ms, best_ms = timeit(torch, lambda: L.validate(), rep=5)
    res["pj_validate"] = {"ms": mean_ms, "n_pos": L.pos_val.numel(), "n_neg": L.neg_val.numel()}
    del ud, L
    torch.cuda.empty_cache()

    # ---- (b1) comp-uprjdists builder: get_desc x2 + pooling with the nSel selected filters -------
    res["uprjdist_build"] = run_uprj_builder(ctx, n_train if not quick else min(n_train, 200000), D // 8, hbm)
    return res


def run_uprj_builder(ctx, n_pairs, n_sel, hbm):
    """src/comp-uprjdists.cpp:298-335 over n_pairs synthetic pairs with n_sel filters picked from the
    gen-poolregion-shaped bank.  Algorithmic bytes per pair: 8192 in + 32 * n_sel out."""
    import torch
    import bench as B
    dev = ctx.device
    full = B.synth_bank(torch, dev)
    rows = (torch.arange(n_sel, device=dev) * 601 + 8 * 2000) % full.shape[0]
    sel = full[rows].contiguous()
    del full
    torch.cuda.empty_cache()
    bank = ctx.bank(sel)
    per_point = 3
    n_points = max(64, int(n_pairs * 0.9) // per_point)
    patches = B.synth_patches(torch, dev, n_points, per_point)
    idx = B.synth_pairs(torch, dev, n_points, per_point, n_pairs)
    torch.cuda.synchronize()
    mean_ms, best_ms = timeit(torch, lambda: ctx.uprjdist_build(bank, patches, idx), warm=1, rep=3)
    ctx.profile(True)                 # per-kernel shares from a separate pass (events serialise the launches)
    ctx.uprjdist_build(bank, patches, idx)
    torch.cuda.synchronize()
    desc_ms, desc_n = ctx.profile_read(0)
    pool_ms, pool_n = ctx.profile_read(1)
    fin_ms, fin_n = ctx.profile_read(6)
    ctx.profile(False)
    # end to end with host buffers: pinned patches / pairs in, the N x 8 nSel matrix out to pinned host memory
    import time
    h_p, h_i = patches.cpu().pin_memory(), idx.cpu().pin_memory()
    h_o = torch.empty((n_pairs, 8 * n_sel), dtype=torch.float32).pin_memory()
    d_p, d_i = torch.empty_like(patches), torch.empty_like(idx)

    def e2e_once():
        d_p.copy_(h_p, non_blocking=True)
        d_i.copy_(h_i, non_blocking=True)
        h_o.copy_(ctx.uprjdist_build(bank, d_p, d_i), non_blocking=True)
        torch.cuda.synchronize()
    e2e_once()
    t0 = time.perf_counter()
    for _ in range(3):
        e2e_once()
    e2e_s = (time.perf_counter() - t0) / 3
    by = n_pairs * (8192 + 32 * n_sel)
    out = {"pairs": n_pairs, "n_sel": n_sel, "ms": mean_ms, "best_ms": best_ms,
           "e2e_host": {"pairs_per_s": n_pairs / e2e_s, "ms": e2e_s * 1e3, "h2d_bytes": int(h_p.numel() + h_i.numel() * 4),
                        "d2h_bytes": int(h_o.numel() * 4)},
           "pairs_per_s": n_pairs / (mean_ms * 1e-3), "desc_per_s": 2 * n_pairs / (mean_ms * 1e-3),
           "algorithmic_bytes_per_pair": 8192 + 32 * n_sel,
           "achieved_gbs": by / (mean_ms * 1e-3) / 1e9, "frac_hbm": by / (mean_ms * 1e-3) / 1e9 / hbm,
           "desc_kernel_ms_per_pass": desc_ms if desc_n else None, "pool_kernel_ms_per_pass": pool_ms if pool_n else None,
           "finish_kernel_ms_per_pass": fin_ms if fin_n else None, "launches_per_pass": desc_n + pool_n + fin_n,
           "pool_gflops": 2.0 * 2 * n_sel * 4096 * 8 * n_pairs / (mean_ms * 1e-3) / 1e9}
    bank.close()
    return out


def _uprj_chunk(sel, patches, idx):
    from oracle import dlco_oracle as O
    from threadpoolctl import threadpool_limits
    with threadpool_limits(1):          # one pair per process, as the reference's OMP loop: no nested BLAS threads
        return O.uprj_distances(sel, patches, idx, dt=np.float32)


def cpu_leg(D=544, quick=False):
    """CPU baselines of rows b, c, d: the oracle's port of the reference's host path
    (src/pr-learn.cpp:303-358, src/pj-learn.cpp:311-527, src/misc.cpp:171-333) on bounded samples.
    numpy -> OpenBLAS with every host thread where the reference calls cv::gemm / cblas_sgemm; the
    serial pr-learn chain is the C restatement (one core, as the reference's `ordered` section)."""
    import ctypes
    import subprocess
    import time
    from oracle import dlco_oracle as O
    lib_path = os.path.join(ROOT, "oracle", "libdlco_oracle.so")
    if not os.path.exists(lib_path):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    lib = ctypes.CDLL(lib_path)
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(3)
    res = {"cores": cores, "kind": "port"}
    F, n = 5120, 20000 if quick else 40000
    labels = (np.arange(n) % 2 == 0).astype(np.uint8)
    dists = rng.random((n, F), dtype=np.float32) * (1.0 + 0.3 * (1 - labels[:, None])).astype(np.float32)
    w = np.zeros(F, np.float32)
    w[rng.permutation(F)[:12]] = 0.05

    def best(fn, rep=3):
        ts = []
        for _ in range(rep):
            t = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t)
        return min(ts)
    # (c2)/(d1) w * Dists^T (src/misc.cpp:226, src/pr-learn.cpp:343-344): SGEMV over the matrix
    dt = best(lambda: dists @ w)
    res["pr_patchdist_gemv"] = {"rows": n, "ms": dt * 1e3, "achieved_gbs": n * F * 4 / dt / 1e9,
                                "sample": "%d x %d f32 rows, numpy/OpenBLAS sgemv, %d threads" % (n, F, cores)}
    pd = (dists @ w).astype(np.float32)
    dt = best(lambda: O.roc_stats(pd, labels))
    res["roc_stats"] = {"n": n, "ms": dt * 1e3, "sample": "sort + ROC + polygon area on %d pairs (numpy)" % n}
    t = time.perf_counter()
    O.pr_stats(O.pr_params(O.ring_params()), dists, labels, w, dt=np.float32)
    dt = time.perf_counter() - t
    res["pr_stats_e2e_host"] = {"rows": n, "ms": dt * 1e3, "models_per_s": 1.0 / dt,
                                "sample": "ComputePRStats on %d x %d (bookkeeping + sgemv + sort + ROC)" % (n, F)}
    # (b1) comp-uprjdists: get_desc x 2 + (nSel x 4096)(4096 x 8) x 2 per pair, the reference's OMP-over-pairs loop
    import multiprocessing as mp
    n_sel = D // 8
    pt_patches, pid = O.synth_patches(40, per_point=3, seed=5)
    sel = O.pr_filters_for_rings(O.ring_params()[np.arange(n_sel // 8 + 1) * 601 % 5120])[0][:n_sel]
    idxu = O.synth_pairs(pid, 256 if quick else 1024, seed=9)
    with mp.get_context("fork").Pool(min(cores, 32)) as pool:
        t = time.perf_counter()
        pool.starmap(_uprj_chunk, [(sel, pt_patches, idxu[c:c + 16]) for c in range(0, len(idxu), 16)])
        dt = time.perf_counter() - t
    res["uprjdist_build"] = {"pairs": len(idxu), "n_sel": n_sel, "ms": dt * 1e3, "pairs_per_s": len(idxu) / dt,
                             "sample": "%d pairs x %d filters, oracle get_desc + sgemm per pair on %d processes" % (len(idxu), n_sel, min(cores, 32))}
    # (c1) the serial RDA chain
    ip, ineg, npt, nnt = O.split_train_val(labels)
    ip, ineg = np.asarray(ip, np.int32), np.asarray(ineg, np.int32)
    vp = ctypes.c_void_p
    lib.dlco_oracle_pr_rda.restype = None
    lib.dlco_oracle_pr_rda.argtypes = [vp, ctypes.c_long, ctypes.c_int, vp, ctypes.c_int, vp, ctypes.c_int,
                                       ctypes.c_float, ctypes.c_float, ctypes.c_ulonglong, ctypes.c_ulonglong,
                                       ctypes.POINTER(ctypes.c_ulonglong), vp, vp]
    steps = 50000 if quick else 200000
    wv, df, st = np.zeros(F, np.float32), np.zeros(F, np.float32), ctypes.c_ulonglong(2215)
    t = time.perf_counter()
    lib.dlco_oracle_pr_rda(dists.ctypes.data, F, F, ip.ctypes.data, npt, ineg.ctypes.data, nnt, 0.035, 0.25, 0,
                           steps, ctypes.byref(st), wv.ctypes.data, df.ctypes.data)
    dt = time.perf_counter() - t
    res["pr_rda_1cfg"] = {"steps": steps, "ms": dt * 1e3, "steps_per_s_per_cfg": steps / dt, "cores_used": 1,
                          "sample": "%d steps over a %d-row matrix, C restatement, 1 core (serial chain)" % (steps, n)}
    # (c3) all-pairs hinge, as the reference's kernel computes it (brute force), OpenMP over host cores
    a = rng.random(50000, dtype=np.float32)
    b = rng.random(50000, dtype=np.float32) + np.float32(0.3)
    lib.dlco_oracle_hinge_pairsum.restype = ctypes.c_double
    lib.dlco_oracle_hinge_pairsum.argtypes = [vp, ctypes.c_int, vp, ctypes.c_int, ctypes.c_double]
    nh = 10000 if quick else 50000
    dt = best(lambda: lib.dlco_oracle_hinge_pairsum(a.ctypes.data, nh, b.ctypes.data, 50000, 1.0), rep=2)
    res["hinge_pairsum_50k_x_50k"] = {"ms": dt * 1e3 * (50000 / nh), "gterms_per_s": nh * 50000 / dt / 1e9,
                                      "sample": "%d x 50000 terms, C brute force, OpenMP %d threads" % (nh, cores)}
    del dists
    # (b) pj-learn: whole RDA iterations (margins, hinge, sum of outer products, update, ssyevr prox)
    nu = 20000
    ulab = (np.arange(nu) % 2 == 0).astype(np.uint8)
    ud = (rng.standard_normal((nu, D)) * (0.08 + 0.048 * (1 - ulab[:, None]))).astype(np.float32)
    up, un, upt, unt = O.split_train_val(ulab)
    state = O.pj_rda_steps(ud, up, un, upt, unt, 0.001, 0.1, 3)
    n_it = 5 if quick else 20
    t = time.perf_counter()
    state = O.pj_rda_steps(ud, up, un, upt, unt, 0.001, 0.1, n_it, state=state)
    dt = time.perf_counter() - t
    res["pj_step"] = {"D": D, "iters_per_s": n_it / dt, "ms_per_iter": dt / n_it * 1e3, "rank": int(state["W"].shape[0]),
                      "sample": "%d iterations at D=%d, numpy/OpenBLAS sgemm + LAPACK ssyevr, %d threads" % (n_it, D, cores)}
    # (d2) ComputePJStats distances
    Wm = state["W"].astype(np.float32)
    dt = best(lambda: O.pj_patch_dist(Wm, ud, np.float32))
    res["pj_patchdist"] = {"rows": nu, "D": D, "rank": int(Wm.shape[0]), "ms": dt * 1e3,
                           "achieved_gbs": nu * D * 4 / dt / 1e9,
                           "sample": "%d x %d rows, numpy/OpenBLAS sgemm, %d threads" % (nu, D, cores)}
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n-train", type=int, default=500000)
    ap.add_argument("--n-test", type=int, default=100000)
    ap.add_argument("--dim", type=int, default=544)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--cpu", action="store_true", help="also time the oracle's CPU port of each row")
    ap.add_argument("--cpu-only", action="store_true")
    ap.add_argument("--only-uprj", action="store_true", help="only the comp-uprjdists builder leg (row b1)")
    args = ap.parse_args()
    if args.cpu_only:
        print(json.dumps({"cpu_baseline": cpu_leg(args.dim, args.quick)}))
        return
    import torch
    import dlco_b200
    ctx = dlco_b200.Context(0)
    if args.only_uprj:
        print(json.dumps({"uprjdist_build": run_uprj_builder(ctx, args.n_train, args.dim // 8, _peaks()[0])}))
        return
    res = run(ctx, args.n_train, args.n_test, args.quick, D=args.dim)
    if args.cpu:
        res["cpu_baseline"] = cpu_leg(args.dim, args.quick)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
