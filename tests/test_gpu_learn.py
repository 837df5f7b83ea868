"""GPU parity tests for the unprojected-distance builder, the hinge pair-sum, the pr-learn and
pj-learn steps and the FPR95 evaluation -- all through the C ABI, against the CPU oracle."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
from conftest import make_bank  # noqa: E402
from oracle import dlco_oracle as O  # noqa: E402


def dev_t(ctx, a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(ctx.device)


# ------------------------------------------------------------------------------------------
# comp-uprjdists: src/comp-uprjdists.cpp:298-335
# ------------------------------------------------------------------------------------------
def test_uprjdist_matches_oracle(gpu_ctx, synth, oracle_lib):
    patches, _, idx = synth
    rings = O.ring_params()
    w = np.zeros(len(rings), np.float32)
    sel_rings = [3, 40, 900, 2600, 4100]
    w[sel_rings] = 1.0
    filt, _ = make_bank(oracle_lib, rings[sel_rings])
    full = np.zeros((len(rings) * 8, 4096), np.float32)
    for k, r in enumerate(sel_rings):
        full[8 * r:8 * r + 8] = filt[8 * k:8 * k + 8]
    sel = O.select_pr_filters(full, w)                         # SelectPRFilters: unique + sortrows
    assert 0 < sel.shape[0] <= 40
    d_p, d_i = dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)
    ids = torch.arange(patches.shape[0], dtype=torch.int32, device=gpu_ctx.device)
    pt = gpu_ctx.get_desc(d_p, ids).cpu().numpy()
    n_sel = sel.shape[0]
    bank = gpu_ctx.bank(sel)
    got = gpu_ctx.uprjdist_build(bank, d_p, d_i).cpu().numpy()
    bank.close()
    want = np.stack([O.uprj_distance(sel, pt[r[0]], pt[r[2]]) for r in idx])
    assert got.shape == want.shape
    # signed differences of values clamped to [0, 1]: absolute tolerance on that scale
    assert np.abs(got - want).max() <= 2e-6, np.abs(got - want).max()


@pytest.mark.parametrize("n_sel,n_pairs,batch", [(68, 700, 256), (68, 1700, 0), (60, 17, 0), (130, 300, 128), (5, 256, 0)])
def test_uprjdist_tensor_path_matches_simt_path(gpu_ctx, oracle_lib, n_sel, n_pairs, batch, monkeypatch):
    """K-split tcgen05 pooling (k_pool_fulldist<1> + k_uprj_finish) against the fp32 SIMT cross-check
    kernel over ragged multi-batch ranges, banks of one and two 128-row tiles and duplicate filters;
    plus the domain's own invariants: swapping the pair negates the row, a patch against itself is 0."""
    if batch:
        monkeypatch.setenv("DLCO_UPRJDIST_BATCH", str(batch))   # default 768: force several batches on few pairs
    rings = O.ring_params()
    pick = (np.arange((n_sel + 7) // 8) * 577 + 900) % len(rings)
    filt, _ = make_bank(oracle_lib, rings[pick])
    sel = filt[:n_sel].copy()
    if n_sel > 8:
        sel[7] = sel[2]                                       # a duplicate row: both outputs must agree
    patches, pid = O.synth_patches(40, per_point=3, seed=3)
    idx = O.synth_pairs(pid, n_pairs, seed=4)
    idx[0, 2] = idx[0, 0]                                     # a patch against itself
    d_p, d_i = dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)
    swapped = idx[:, [2, 3, 0, 1]].copy()
    bank = gpu_ctx.bank(sel)
    got = gpu_ctx.uprjdist_build(bank, d_p, d_i).cpu().numpy()
    got_sw = gpu_ctx.uprjdist_build(bank, d_p, dev_t(gpu_ctx, swapped)).cpu().numpy()
    part = gpu_ctx.uprjdist_build(bank, d_p, d_i, lo=3, hi=min(n_pairs, 270)).cpu().numpy()
    monkeypatch.setenv("DLCO_POOL_IMPL", "simt")
    ref = gpu_ctx.uprjdist_build(bank, d_p, d_i).cpu().numpy()
    monkeypatch.delenv("DLCO_POOL_IMPL")
    bank.close()
    assert got.shape == (n_pairs, 8 * n_sel)
    assert np.abs(got - ref).max() <= 2e-6, np.abs(got - ref).max()
    assert np.abs(got).max() > 0.05
    assert np.array_equal(got_sw, -got)
    assert not got[0].any()
    assert np.array_equal(part, got[3:min(n_pairs, 270)])
    if n_sel > 8:
        assert np.array_equal(got[:, 56:64], got[:, 16:24])


# ------------------------------------------------------------------------------------------
# SubtractVectorsByRows + sum: src/kernelop-opencv.cu:49-80
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n1,n2", [(1, 1), (5000, 7001), (257, 4096), (4097, 3)])
def test_hinge_pairsum(gpu_ctx, n1, n2):
    rng = np.random.default_rng(n1 * 7 + n2)
    pos = rng.normal(0, 1, n1).astype(np.float32)
    neg = rng.normal(0.5, 1, n2).astype(np.float32)
    got = gpu_ctx.hinge_pairsum(dev_t(gpu_ctx, pos), dev_t(gpu_ctx, neg))
    want = O.hinge_pairsum(pos, neg)
    assert abs(got - want) <= 1e-6 * max(abs(want), 1.0)


def test_hinge_pairsum_all_inactive(gpu_ctx):
    pos = np.full(100, -5.0, np.float32)
    neg = np.full(333, 5.0, np.float32)
    assert gpu_ctx.hinge_pairsum(dev_t(gpu_ctx, pos), dev_t(gpu_ctx, neg)) == 0.0


def test_hinge_pairsum_reference_size(gpu_ctx, oracle_lib):
    """50 000 x 50 000 (the reference's validation size) against the C brute force."""
    import ctypes
    rng = np.random.default_rng(3)
    pos = rng.gamma(2.0, 1.0, 50000).astype(np.float32)
    neg = rng.gamma(3.0, 1.5, 50000).astype(np.float32)
    got = gpu_ctx.hinge_pairsum(dev_t(gpu_ctx, pos), dev_t(gpu_ctx, neg))
    want = oracle_lib.dlco_oracle_hinge_pairsum(pos.ctypes.data_as(ctypes.c_void_p), 50000,
                                                neg.ctypes.data_as(ctypes.c_void_p), 50000, 1.0)
    assert abs(got - want) <= 1e-6 * want
    assert abs(O.hinge_pairsum(pos, neg) - want) <= 1e-9 * want


# ------------------------------------------------------------------------------------------
# synthetic "Distance" matrices for the learners
# ------------------------------------------------------------------------------------------
def synth_fulldist(n, f, seed):
    rng = np.random.default_rng(seed)
    lab = (rng.random(n) < 0.5).astype(np.uint8)
    base = rng.gamma(2.0, 0.5, (n, f)).astype(np.float32)
    informative = rng.random(f) < 0.3
    scale = np.where(lab[:, None] == 1, np.where(informative, 0.35, 1.0), 1.0).astype(np.float32)
    return (base * scale).astype(np.float32), lab


def synth_uprj(n, d, seed):
    rng = np.random.default_rng(seed)
    lab = (rng.random(n) < 0.5).astype(np.uint8)
    x = rng.normal(0, 0.25, (n, d)).astype(np.float32)
    x[lab == 1, : d // 2] *= 0.6
    return x, lab


# ------------------------------------------------------------------------------------------
# pr-learn: src/pr-learn.cpp:303-358
# ------------------------------------------------------------------------------------------
def test_pr_rda_matches_oracle(gpu_ctx):
    import dlco_b200
    dists, lab = synth_fulldist(3000, 64, 1)
    d_d = dev_t(gpu_ctx, dists)
    L = dlco_b200.learn.PrLearner(gpu_ctx, d_d, lab, 0.035, 0.25)
    ip, ineg, npt, nnt = O.split_train_val(lab)
    assert np.array_equal(ip, L.idx_pos.cpu().numpy()) and np.array_equal(ineg, L.idx_neg.cpu().numpy())
    assert (npt, nnt) == (L.n_pos_trn, L.n_neg_trn)
    n1, n2 = 700, 1300                                   # two calls: state and RNG carry over
    L.run_steps(n1)
    L.run_steps(n2)
    w_o, df_o, rng, gap = O.pr_rda_steps(dists, ip, ineg, npt, nnt, 0.035, 0.25, n1 + n2)
    assert L.rng_state == rng.state                       # identical sample stream consumed
    w_g, df_g = L.w[0].cpu().numpy(), L.dfavg[0].cpu().numpy()
    assert gap > 1e-6, "oracle run has a near-tie hinge decision; pick another seed"
    assert np.abs(df_g - df_o).max() <= 1e-4 * np.abs(df_o).max()
    assert np.abs(w_g - w_o).max() <= 1e-4 * max(np.abs(w_o).max(), 1e-12)
    assert np.array_equal(w_g > 0, w_o > 0)               # same support
    # validation objective
    loss, regul = L.validate(0)
    lo, ro = O.pr_validate(dists[ip[npt:]], dists[ineg[nnt:]], w_g, 0.035)
    assert abs(loss - lo) <= 1e-5 * max(lo, 1e-6) and abs(regul - ro) <= 1e-6 * max(ro, 1e-6)


def test_pr_rda_grid_equals_single_runs(gpu_ctx):
    """one CTA per (mu, gamma): the grid launch must reproduce the single-config runs bit for bit"""
    import dlco_b200
    dists, lab = synth_fulldist(2000, 128, 2)
    d_d = dev_t(gpu_ctx, dists)
    mus, gammas = [0.01, 0.035, 0.05], [0.1, 0.25, 1.0]
    G = dlco_b200.learn.PrLearner(gpu_ctx, d_d, lab, mus, gammas)
    G.run_steps(1500)
    for c in range(3):
        S = dlco_b200.learn.PrLearner(gpu_ctx, d_d, lab, mus[c], gammas[c])
        S.run_steps(1500)
        assert torch.equal(S.w[0], G.w[c]) and torch.equal(S.dfavg[0], G.dfavg[c])


def test_pr_train_log_format(gpu_ctx):
    import re
    import dlco_b200
    dists, lab = synth_fulldist(2000, 64, 5)
    lines = []
    L = dlco_b200.learn.PrLearner(gpu_ctx, dev_t(gpu_ctx, dists), lab, 0.035, 0.25, log_step=500,
                                  emit=lines.append)
    L.train(1500)
    heads = [ln.split(":")[0] for ln in lines]
    assert heads[0] == "Best" and heads[1] == "Stat"
    steps = [int(re.match(r"(Best|Step): (\d+)  Loss", ln).group(2)) for ln in lines if ln[:4] in ("Best", "Step")]
    assert steps == [500, 1000, 1500]
    pat = r"^(Best|Step): \d+  Loss: \d+\.\d{6} Regul: \d+\.\d{6} Obj: \d+\.\d{6} \(\d+\.\d{6}\)  NNZ: \d+ \(\d+\)  Ttime: \d+\.\d{4} Vtime: \d+\.\d{4}$"
    assert all(re.match(pat, ln) for ln in lines if ln[:4] in ("Best", "Step"))


# ------------------------------------------------------------------------------------------
# FPR95 evaluation: src/misc.cpp:171-333
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1000, 4097, 100000])
def test_roc_stats(gpu_ctx, n):
    rng = np.random.default_rng(n)
    lab = (rng.random(n) < 0.4).astype(np.uint8)
    pd = (rng.gamma(2.0, 1.0, n) * np.where(lab == 1, 0.5, 1.0)).astype(np.float32)
    pd[::7] = pd[3]                                           # ties: stable order by index
    f, a = gpu_ctx.roc_stats(dev_t(gpu_ctx, pd), dev_t(gpu_ctx, lab))
    fo, ao = O.roc_stats(pd, lab)
    assert f == float(fo)                                     # same float counters: bit-exact
    assert abs(a - ao) <= 1e-12


@pytest.mark.parametrize("n", [1, 31, 4096, 4100, 500003])
def test_roc_stats_sort_is_stable_and_total(gpu_ctx, n):
    """The hand-written radix sort behind dlco_roc_stats (csrc/stats.cu): heavy ties (keys quantised to 64 levels),
    negative keys, -0 / +0, tile and warp edges -- the ROC counters must equal the oracle's, whose argsort is stable."""
    rng = np.random.default_rng(n)
    lab = (rng.random(n) < 0.5).astype(np.uint8)
    pd = (np.round(rng.normal(0, 4, n) + lab * -1.5) / 8).astype(np.float32)    # ~64 distinct values, both signs
    pd[rng.random(n) < 0.05] = np.float32(-0.0)
    pd[rng.random(n) < 0.05] = np.float32(0.0)
    if lab.min() == lab.max():
        lab[0] = 1 - lab[0] if n > 1 else lab[0]
    f, a = gpu_ctx.roc_stats(dev_t(gpu_ctx, pd), dev_t(gpu_ctx, lab))
    fo, ao = O.roc_stats(pd, lab)
    if n > 1:
        assert f == float(fo) and abs(a - ao) <= 1e-12, (f, float(fo), a, ao)


def test_pr_patchdist_and_stats(gpu_ctx):
    dists, lab = synth_fulldist(5000, 256, 9)
    rng = np.random.default_rng(1)
    w = np.where(rng.random(256) < 0.1, rng.random(256), 0).astype(np.float32)
    pd = gpu_ctx.pr_patchdist(dev_t(gpu_ctx, dists), dev_t(gpu_ctx, w)).cpu().numpy()
    want = dists.astype(np.float64) @ w.astype(np.float64)
    assert np.abs(pd - want).max() <= 1e-6 * np.abs(want).max()
    f, a = gpu_ctx.roc_stats(dev_t(gpu_ctx, pd), dev_t(gpu_ctx, lab))
    fo, ao = O.roc_stats(want.astype(np.float32), lab)
    assert abs(f - float(fo)) <= 1e-3 and abs(a - ao) <= 1e-5   # FPR95 within 0.1 pp


# ------------------------------------------------------------------------------------------
# tcgen05 3xTF32 GEMM used by the pj-learn SYRK
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("m,n,k", [(128, 64, 32), (544, 544, 400), (100, 70, 50), (640, 65, 1000), (1500, 1300, 96)])
def test_gemm3x_nt_fp32_accuracy(gpu_ctx, m, n, k):
    rng = np.random.default_rng(m + n + k)
    a = rng.normal(0, 1, (m, k)).astype(np.float32)
    b = rng.normal(0, 1, (n, k)).astype(np.float32)
    got = gpu_ctx.gemm3x_nt(dev_t(gpu_ctx, a), dev_t(gpu_ctx, b)).cpu().numpy()
    want = a.astype(np.float64) @ b.astype(np.float64).T
    # the tensor core accumulates with truncation: the error grows with K (about 4e-6 of the
    # largest entry at K = 512, measured); the SYRK it serves has K = 400
    assert np.abs(got - want).max() <= 1e-5 * np.abs(want).max(), np.abs(got - want).max()


# ------------------------------------------------------------------------------------------
# pj-learn: src/pj-learn.cpp:311-527
# ------------------------------------------------------------------------------------------
def test_pj_grad_matches_oracle(gpu_ctx):
    x, lab = synth_uprj(3000, 48, 11)
    rng = np.random.default_rng(0)
    b = 50
    rp = rng.integers(0, 3000, b).astype(np.int32)
    rn = rng.integers(0, 3000, b).astype(np.int32)
    W = rng.normal(0, 1.0, (20, 48)).astype(np.float32)
    got = gpu_ctx.pj_rda_grad(dev_t(gpu_ctx, x), dev_t(gpu_ctx, rp), dev_t(gpu_ctx, rn), dev_t(gpu_ctx, W), 20)
    got = got.cpu().numpy()
    pd = O.pj_margins(W, x[rp]).astype(np.float32)
    nd = O.pj_margins(W, x[rn]).astype(np.float32)
    gap = np.abs((pd[:, None] + np.float32(1.0)) - nd[None, :]).min()
    assert gap > 1e-5, "near-tie hinge in the test data; change the seed"
    want, c_pos, c_neg = O.pj_grad(x[rp], x[rn], pd, nd)
    assert c_pos.sum() > 0 and c_pos.sum() < b * b            # hinge partly active
    assert np.abs(got - want).max() <= 1e-5 * np.abs(want).max()
    # rank 0 (t = 0: W is all zeros -> every pair violates)
    got0 = gpu_ctx.pj_rda_grad(dev_t(gpu_ctx, x), dev_t(gpu_ctx, rp), dev_t(gpu_ctx, rn), None, 0).cpu().numpy()
    z = np.zeros(b, np.float32)
    want0, _, _ = O.pj_grad(x[rp], x[rn], z, z)
    assert np.abs(got0 - want0).max() <= 1e-5 * np.abs(want0).max()


def test_pj_update_eig_prox(gpu_ctx):
    rng = np.random.default_rng(5)
    D = 40
    dl = rng.normal(0, 100, (D, D)).astype(np.float32)
    dl = (dl + dl.T).astype(np.float32)
    df = rng.normal(0, 0.01, (D, D)).astype(np.float32)
    t, b, mu, gamma = 7, 200, 0.001, 0.1
    d_df, d_A = dev_t(gpu_ctx, df), torch.zeros((D, D), device=gpu_ctx.device)
    gpu_ctx.pj_rda_update(t, b, mu, gamma, dev_t(gpu_ctx, dl), d_df, d_A)
    f32 = np.float32
    beta = f32(1.0) / f32(b * b * (t + 1))
    df_o = (df * f32(t / (t + 1)) + dl * beta).astype(f32)
    A_o = ((df_o + f32(mu) * np.eye(D, dtype=f32)) * f32(-np.sqrt(t + 1.0) / float(f32(gamma)))).astype(f32)
    A_o = (f32(0.5) * (A_o + A_o.T)).astype(f32)
    assert np.array_equal(d_df.cpu().numpy(), df_o)           # same fp32 operation order: exact
    assert np.array_equal(d_A.cpu().numpy(), A_o)
    ev, evec = gpu_ctx.pj_eig(d_A)
    W = torch.zeros((D, D), device=gpu_ctx.device)
    rank = gpu_ctx.pj_prox_apply(ev, evec, d_A, W)
    A_ref, W_ref, e = O.pj_prox(A_o.astype(np.float64), np.float64)
    assert rank == W_ref.shape[0] and 0 < rank < D
    A_g, W_g = d_A.cpu().numpy(), W.cpu().numpy()
    assert np.abs(A_g - A_ref).max() <= 1e-5 * np.abs(A_ref).max()
    assert np.abs(W_g[:rank].T @ W_g[:rank] - A_ref).max() <= 1e-5 * np.abs(A_ref).max()
    assert np.abs(W_g[rank:]).max() == 0.0
    # rows ordered by ascending eigenvalue
    norms = (W_g[:rank] ** 2).sum(axis=1)
    assert np.all(np.diff(norms) >= -1e-6 * norms.max())


def test_pj_rda_matches_oracle(gpu_ctx):
    import dlco_b200
    x, lab = synth_uprj(3000, 32, 4)
    mu, gamma, b, steps = 0.001, 0.1, 50, 40
    L = dlco_b200.learn.PjLearner(gpu_ctx, dev_t(gpu_ctx, x), lab, mu, gamma, sz_batch=b)
    for _ in range(steps):
        L.step()
    ip, ineg, npt, nnt = O.split_train_val(lab)
    st = O.pj_rda_steps(x, ip, ineg, npt, nnt, mu, gamma, steps, sz_batch=b, dt=np.float64)
    assert L.rng.state == st["rng"].state
    A_g = L.A.cpu().numpy()
    assert np.abs(A_g - st["A"]).max() <= 1e-4 * np.abs(st["A"]).max()      # north_star: A within 1e-4
    assert L.rank == st["W"].shape[0]
    W_g = L.W.cpu().numpy()[:L.rank]
    assert np.abs(W_g.T @ W_g - A_g).max() <= 1e-5 * np.abs(A_g).max()      # A = W^T W
    loss, regul = L.validate()
    lo, ro = O.pj_validate(x[ip[npt:]], x[ineg[nnt:]], st["W"], st["A"], mu)
    assert abs(loss - lo) <= 1e-4 * lo and abs(regul - ro) <= 1e-4 * ro
    f, a = L.stats()
    dim, fo, ao = O.pj_stats(st["W"], x, lab)
    assert dim == L.rank and abs(f - float(fo)) <= 1e-3 and abs(a - ao) <= 1e-4


def test_pj_grad_grid_equals_single(gpu_ctx):
    x, lab = synth_uprj(2000, 64, 12)
    rng = np.random.default_rng(1)
    b = 40
    rp = dev_t(gpu_ctx, rng.integers(0, 2000, b).astype(np.int32))
    rn = dev_t(gpu_ctx, rng.integers(0, 2000, b).astype(np.int32))
    d_x = dev_t(gpu_ctx, x)
    Ws = rng.normal(0, 0.8, (3, 64, 64)).astype(np.float32)
    ranks = [10, 64, 0]
    d_W = dev_t(gpu_ctx, Ws)
    grid = gpu_ctx.pj_rda_grad_grid(d_x, rp, rn, d_W, ranks)
    for g in range(3):
        single = gpu_ctx.pj_rda_grad(d_x, rp, rn, d_W[g], ranks[g])
        assert torch.equal(single, grid[g])


# ------------------------------------------------------------------------------------------
# staged gradient entry points + shard.py driven by the real Context (world of 1)
# ------------------------------------------------------------------------------------------
def test_pj_staged_grad_equals_fused(gpu_ctx):
    x, lab = synth_uprj(3000, 72, 21)
    rng = np.random.default_rng(3)
    b = 64
    rp = rng.integers(0, 3000, b).astype(np.int32)
    rn = rng.integers(0, 3000, b).astype(np.int32)
    W = rng.normal(0, 0.8, (30, 72)).astype(np.float32)
    d_x, d_W = dev_t(gpu_ctx, x), dev_t(gpu_ctx, W)
    fused = gpu_ctx.pj_rda_grad(d_x, dev_t(gpu_ctx, rp), dev_t(gpu_ctx, rn), d_W, 30).cpu().numpy()
    rows = dev_t(gpu_ctx, np.concatenate([rp, rn]))
    m = gpu_ctx.pj_margins(d_x, rows, d_W, 30)
    c = gpu_ctx.pj_hinge_weights(m)
    staged = gpu_ctx.pj_wsyrk(d_x, rows, c).cpu().numpy()
    # staged: general weights, 3xTF32 planes; fused: hinge counts, fp16 hi/lo planes (DLCO_SYRK_PLANES=tf32 makes
    # them bitwise equal) -- the same 22-bit operands, different rounding of the sums
    assert np.abs(fused - staged).max() <= 2e-6 * np.abs(fused).max()
    # split the batch rows in two "ranks": partial SYRKs add up to the whole
    half = rows.numel() // 3
    part = gpu_ctx.pj_wsyrk(d_x, rows[:half].contiguous(), c[:half].contiguous()) + \
        gpu_ctx.pj_wsyrk(d_x, rows[half:].contiguous(), c[half:].contiguous())
    assert np.abs(part.cpu().numpy() - fused).max() <= 2e-6 * np.abs(fused).max()
    # empty share
    z = gpu_ctx.pj_wsyrk(d_x, rows[:0].contiguous(), c[:0].contiguous())
    assert float(z.abs().max()) == 0.0


def test_sharded_driver_world1_matches_direct(gpu_ctx):
    import dlco_b200
    from dlco_b200 import shard
    x, lab = synth_uprj(4000, 48, 5)
    d_x, d_lab = dev_t(gpu_ctx, x), dev_t(gpu_ctx, lab)
    rng = np.random.default_rng(8)
    W = rng.normal(0, 1.0, (16, 48)).astype(np.float32)
    d_W = dev_t(gpu_ctx, W)
    A = dev_t(gpu_ctx, (W.T @ W).astype(np.float32))
    ip, ineg, npt, nnt = dlco_b200.learn.split_train_val(lab)
    coll = shard.Collective()
    sp = shard.ShardedPj(gpu_ctx, coll, d_x, d_lab, 0, 4000)
    rp, rn = ip[:40], ineg[:40]
    g = sp.grad(rp, rn, d_W, 16).cpu().numpy()
    want = gpu_ctx.pj_rda_grad(d_x, dev_t(gpu_ctx, rp), dev_t(gpu_ctx, rn), d_W, 16).cpu().numpy()
    # the staged SYRK keeps general weights on 3xTF32 planes, the fused one runs fp16 hi/lo planes: same
    # 22-bit operands, different rounding of the sums
    assert np.abs(g - want).max() <= 2e-6 * np.abs(want).max()
    loss, regul = sp.validate(ip[npt:], ineg[nnt:], d_W, 16, A, 0.001)
    l2, r2 = gpu_ctx.pj_validate(d_x, dev_t(gpu_ctx, ip[npt:]), dev_t(gpu_ctx, ineg[nnt:]), d_W, 16, A, 0.001)
    assert abs(loss - l2) <= 1e-6 * abs(l2) and abs(regul - r2) <= 1e-6 * abs(r2)
    f, a = sp.stats(d_W, 16)
    f2, a2 = gpu_ctx.roc_stats(gpu_ctx.pj_patchdist(d_x, d_W, 16), d_lab)
    assert (f, a) == (f2, a2)


@pytest.mark.parametrize("nnz", [1, 12, 300, 400])
def test_pr_patchdist_sparse_and_dense_paths(gpu_ctx, nnz):
    """the gather path (few non-zeros in w) and the streaming path give the same distances"""
    import os
    dists, lab = synth_fulldist(3000, 5120, 4)
    rng = np.random.default_rng(nnz)
    w = np.zeros(5120, np.float32)
    w[rng.choice(5120, nnz, replace=False)] = rng.random(nnz).astype(np.float32) + 0.01
    d_d, d_w = dev_t(gpu_ctx, dists), dev_t(gpu_ctx, w)
    want = dists.astype(np.float64) @ w.astype(np.float64)
    got = gpu_ctx.pr_patchdist(d_d, d_w).cpu().numpy()           # sparse when nnz <= 5120 / 16
    os.environ["DLCO_GEMV_SPARSE"] = "0"
    try:
        dense = gpu_ctx.pr_patchdist(d_d, d_w).cpu().numpy()
    finally:
        os.environ.pop("DLCO_GEMV_SPARSE")
    assert np.abs(got - want).max() <= 1e-6 * np.abs(want).max()
    assert np.abs(dense - want).max() <= 1e-6 * np.abs(want).max()
    rows = dev_t(gpu_ctx, rng.integers(0, 3000, 777).astype(np.int32))
    m = gpu_ctx.pr_margins(d_d, rows, d_w).cpu().numpy()
    assert np.abs(m - want[rows.cpu().numpy()]).max() <= 1e-6 * np.abs(want).max()


# ------------------------------------------------------------------------------------------
# lock-step (mu, gamma) sweep: device-resident ranks / RNG, batched eigen-decomposition
# ------------------------------------------------------------------------------------------
def test_pj_sample_batch_matches_host_rng(gpu_ctx):
    import dlco_b200
    x, lab = synth_uprj(3000, 16, 4)
    L = dlco_b200.learn.PjLearner(gpu_ctx, dev_t(gpu_ctx, x), lab, 0.001, 0.1, sz_batch=50)
    S = dlco_b200.learn.PjSweep(gpu_ctx, dev_t(gpu_ctx, x), lab, [0.001], [0.1], sz_batch=50)
    rp_d = torch.empty(50, dtype=torch.int32, device=gpu_ctx.device)
    rn_d = torch.empty_like(rp_d)
    for _ in range(3):
        rp, rn = L.sample_batch()
        gpu_ctx.pj_sample_batch(S.idx_pos, S.n_pos_trn, S.idx_neg, S.n_neg_trn, 50, S.rng_state, rp_d, rn_d)
        gpu_ctx.sync()
        np.testing.assert_array_equal(rp_d.cpu().numpy(), rp)
        np.testing.assert_array_equal(rn_d.cpu().numpy(), rn)
    assert int(S.rng_state.cpu().numpy().view(np.uint64)[0]) == L.rng.state


def test_pj_eig_batched_reconstructs(gpu_ctx):
    rng = np.random.default_rng(2)
    n, D = 5, 96
    M = rng.normal(0, 1, (n, D, D)).astype(np.float32)
    A = ((M + M.transpose(0, 2, 1)) * 0.5).astype(np.float32)
    ev, evec = gpu_ctx.pj_eig_batched(dev_t(gpu_ctx, A))
    ev, evec = ev.cpu().numpy().astype(np.float64), evec.cpu().numpy().astype(np.float64)
    for g in range(n):
        assert np.all(np.diff(ev[g]) >= 0)
        want = np.linalg.eigvalsh(A[g].astype(np.float64))
        assert np.abs(ev[g] - want).max() <= 2e-5 * np.abs(want).max()
        rec = evec[g].T @ np.diag(ev[g]) @ evec[g]              # row k = eigenvector k
        assert np.abs(rec - A[g]).max() <= 2e-5 * np.abs(A[g]).max()


def teacher_forced_worst(gpu_ctx, x, lab, mu, gamma, steps, sz_batch, prox):
    """Worst per-iteration relative error of (dfAvg, A) when every GPU iteration is replayed by the fp64
    oracle from the GPU's own state.  Free-running comparisons fall back on this when one of the
    sz_batch^2 hinge decisions per iteration sits within fp32 rounding of a tie and flips (the
    trajectories then part by ~1e-4..1e-3 although every kernel is within 1e-6)."""
    import dlco_b200
    S = dlco_b200.learn.PjSweep(gpu_ctx, dev_t(gpu_ctx, x), lab, [mu], [gamma], sz_batch=sz_batch, prox=prox)
    ip, ineg, npt, nnt = O.split_train_val(lab)
    rel = lambda a, b: float(np.abs(a - b).max() / np.abs(b).max())          # noqa: E731
    worst = 0.0
    for t in range(steps):
        rank0 = int(S.ranks()[0]) if t else 0
        rng = O.CvRNG(2215)
        rng.state = int(S.rng_state.cpu().numpy().view(np.uint64)[0])
        S.refresh_w()
        W0 = S.W[0].cpu().numpy().astype(np.float64)
        snap = dict(W=W0[:rank0] if rank0 else np.zeros_like(W0), A=S.A[0].cpu().numpy().astype(np.float64),
                    dfAvg=S.dfavg[0].cpu().numpy().astype(np.float64), t=t, rng=rng)
        S.step()
        tf = O.pj_rda_steps(x, ip, ineg, npt, nnt, mu, gamma, 1, sz_batch=sz_batch, dt=np.float64, state=snap)
        worst = max(worst, rel(S.dfavg[0].cpu().numpy(), tf["dfAvg"]), rel(S.A[0].cpu().numpy(), tf["A"]))
    return worst


def assert_a_matches_oracle(gpu_ctx, A_g, st, x, lab, mu, gamma, steps, sz_batch, prox):
    """north_star: learnt A within 1e-4 relative after a fixed number of iterations.  Free-running when no
    hinge decision flips; otherwise per iteration (teacher-forced) with the free-run drift bounded."""
    err = float(np.abs(A_g - st["A"]).max() / np.abs(st["A"]).max())
    if err <= 1e-4:
        return True
    assert err <= 5e-3, err
    assert teacher_forced_worst(gpu_ctx, x, lab, mu, gamma, steps, sz_batch, prox) <= 1e-4
    return False


def test_pj_sweep_matches_single_runs_and_oracle(gpu_ctx):
    import dlco_b200
    x, lab = synth_uprj(3000, 32, 4)
    b, steps = 50, 40
    mus, gammas = [0.001, 0.0005, 0.002], [0.1, 0.05, 0.2]
    d_x = dev_t(gpu_ctx, x)
    S = dlco_b200.learn.PjSweep(gpu_ctx, d_x, lab, mus, gammas, sz_batch=b)
    for _ in range(steps):
        S.step()
    ranks = S.ranks()
    ip, ineg, npt, nnt = O.split_train_val(lab)
    for c in range(3):
        st = O.pj_rda_steps(x, ip, ineg, npt, nnt, mus[c], gammas[c], steps, sz_batch=b, dt=np.float64)
        A_g = S.A[c].cpu().numpy()
        same_path = assert_a_matches_oracle(gpu_ctx, A_g, st, x, lab, mus[c], gammas[c], steps, b, "eig")
        assert abs(int(ranks[c]) - st["W"].shape[0]) <= (0 if same_path else 2)
        W_g = S.W[c].cpu().numpy()[:ranks[c]]
        assert np.abs(W_g.T @ W_g - A_g).max() <= 1e-5 * np.abs(A_g).max()
        assert not S.W[c, ranks[c]:].any().item()
        L = dlco_b200.learn.PjLearner(gpu_ctx, d_x, lab, mus[c], gammas[c], sz_batch=b)
        for _ in range(steps):
            L.step()
        assert abs(L.rank - int(ranks[c])) <= (0 if same_path else 2)
        assert (L.A - S.A[c]).abs().max().item() <= (2e-5 if same_path else 5e-3) * L.A.abs().max().item()
        loss, regul = S.validate(c, ranks[c])
        lo, ro = O.pj_validate(x[ip[npt:]], x[ineg[nnt:]], st["W"], st["A"], mus[c])
        tol = 1e-4 if same_path else 5e-3
        assert abs(loss - lo) <= tol * lo and abs(regul - ro) <= tol * ro
    assert int(S.rng_state.cpu().numpy().view(np.uint64)[0]) == st["rng"].state


def test_pj_sweep_log_format(gpu_ctx):
    import dlco_b200
    import re
    x, lab = synth_uprj(3000, 32, 9)
    lines = []
    S = dlco_b200.learn.PjSweep(gpu_ctx, dev_t(gpu_ctx, x), lab, [0.001, 0.002], [0.1, 0.2], sz_batch=40,
                                log_step=10, emit=lines.append)
    S.train(25)
    assert sum(s.startswith("[0.001/0.1] Best: ") for s in lines) >= 1
    pat = re.compile(r"^\[[0-9.e-]+/[0-9.e-]+\] (Best|Step): \d+  Loss: [0-9.]+ Regul: [0-9.]+ Obj: [0-9.]+ "
                     r"\([0-9.]+\) Rank: \d+ \(\d+\) Ttime: [0-9.]+ Vtime: [0-9.]+$")
    assert all(pat.match(s) for s in lines if " Best: " in s or " Step: " in s)
    assert S.W_save[0] is not None and S.A_save[0] is not None


# ------------------------------------------------------------------------------------------
# tensor-core margins / ComputePJStats distances (sqnorm_tc.cu) against fp64 and the SIMT kernel
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,D,rank,gather", [
    (1000, 544, 80, False), (4097, 480, 120, True), (400, 544, 544, True), (129, 64, 3, False),
    (70000, 544, 48, False), (300, 36, 36, False), (2500, 544, 300, True), (128, 256, 256, False)])
def test_sqnorm_tc_matches_fp64_and_simt(gpu_ctx, n, D, rank, gather):
    import os
    rng = np.random.default_rng(n + D + rank)
    nx = n * 2 if gather else n
    x = rng.uniform(-1, 1, (nx, D)).astype(np.float32) * rng.uniform(0.01, 1, (nx, 1)).astype(np.float32)
    x[rng.random((nx, D)) < 0.1] = 0
    W = np.zeros((D, D), np.float32)
    W[:rank] = rng.normal(0, 1, (rank, D)).astype(np.float32) * np.geomspace(1e-3, 3.0, rank)[:, None].astype(np.float32)
    rows = rng.permutation(nx)[:n].astype(np.int32) if gather else None
    d_x, d_W = dev_t(gpu_ctx, x), dev_t(gpu_ctx, W)
    d_rows = dev_t(gpu_ctx, rows) if gather else None
    xs = x[rows] if gather else x
    want = ((xs.astype(np.float64) @ W[:rank].astype(np.float64).T) ** 2).sum(axis=1)
    got = {}
    for mode in ("0", "1"):
        os.environ["DLCO_SQNORM_SIMT"] = mode
        try:
            if gather:
                out = gpu_ctx.pj_margins(d_x, d_rows, d_W, rank)
            else:
                out = gpu_ctx.pj_patchdist(d_x, d_W, rank)
            gpu_ctx.sync()
        finally:
            os.environ.pop("DLCO_SQNORM_SIMT")
        got[mode] = out.cpu().numpy().astype(np.float64)
        assert np.abs(got[mode] - want).max() <= 1e-5 * np.abs(want).max(), mode
    # element-wise, relative to each row's own value
    nz = want > 1e-6 * want.max()
    assert (np.abs(got["0"][nz] - want[nz]) / want[nz]).max() <= 2e-5


def test_sqnorm_tc_saturates_out_of_contract_inputs(gpu_ctx):
    """|x| > 127 is outside the contract of the tensor path: it must stay finite (saturation)."""
    x = np.full((256, 64), 1000.0, np.float32)
    W = np.eye(64, dtype=np.float32)
    out = gpu_ctx.pj_patchdist(dev_t(gpu_ctx, x), dev_t(gpu_ctx, W), 64)
    assert torch.isfinite(out).all().item()


def test_pr_rda_200k_steps_full_width_matches_c_oracle(gpu_ctx, oracle_lib):
    """pr-learn parity after a fixed 200 000 steps at the real width (5120 rings),
    (mu, gamma) = (0.035, 0.25), against the C restatement of src/pr-learn.cpp:303-329 (fp32
    reference order); north_star: learnt w within 1e-4 relative."""
    import ctypes
    import dlco_b200
    n, F, steps = 12000, 5120, 200000
    rng = np.random.default_rng(77)
    lab = (np.arange(n) % 2 == 0).astype(np.uint8)
    dists = rng.random((n, F), dtype=np.float32)
    dists *= (1.0 + 0.3 * (1 - lab[:, None])).astype(np.float32)
    L = dlco_b200.learn.PrLearner(gpu_ctx, dev_t(gpu_ctx, dists), lab, 0.035, 0.25)
    L.run_steps(steps)
    gpu_ctx.sync()
    ip, ineg = L.idx_pos.cpu().numpy().astype(np.int32), L.idx_neg.cpu().numpy().astype(np.int32)
    w, df, st = np.zeros(F, np.float32), np.zeros(F, np.float32), ctypes.c_ulonglong(2215)
    vp = ctypes.c_void_p
    oracle_lib.dlco_oracle_pr_rda.restype = None
    oracle_lib.dlco_oracle_pr_rda.argtypes = [vp, ctypes.c_long, ctypes.c_int, vp, ctypes.c_int, vp, ctypes.c_int,
                                              ctypes.c_float, ctypes.c_float, ctypes.c_ulonglong,
                                              ctypes.c_ulonglong, ctypes.POINTER(ctypes.c_ulonglong), vp, vp]
    oracle_lib.dlco_oracle_pr_rda(dists.ctypes.data, F, F, ip.ctypes.data, L.n_pos_trn, ineg.ctypes.data,
                                  L.n_neg_trn, 0.035, 0.25, 0, steps, ctypes.byref(st), w.ctypes.data, df.ctypes.data)
    assert L.rng_state == st.value
    w_g, df_g = L.w[0].cpu().numpy(), L.dfavg[0].cpu().numpy()
    assert np.abs(df_g - df).max() <= 1e-4 * np.abs(df).max()
    assert np.abs(w_g - w).max() <= 1e-4 * max(np.abs(w).max(), 1e-12)
    assert (w > 0).sum() > 0


@pytest.mark.parametrize("prox", ["eig", "sign"])
def test_pj_sweep_batch200_150_steps_matches_oracle(gpu_ctx, prox):
    """the reference's batch size (200 + 200) over 150 iterations at D = 160.

    (1) teacher-forced: every iteration is replayed by the fp64 oracle FROM THE GPU'S OWN STATE
        (W, A, dfAvg, t, RNG) and the resulting dfAvg and A must agree within 1e-4 relative
        (north_star's bar) -- this holds the kernels (sample, margins, hinge, SYRK, update, eig,
        prox) to the bar on each of the 150 iterations, on the states a real run visits.
    (2) free-running against the fp64 oracle from t = 0: the recursion amplifies fp32 rounding (40 000
        hinge decisions per iteration, some within rounding of a tie; eigenvalues crossing zero), so
        the reference's own fp32 arithmetic drifts from fp64 as well (oracle dt=float32 vs float64).
        The GPU must stay within max(1e-4, 4 x that drift): as close to the exact recursion as the
        reference's own arithmetic is."""
    import dlco_b200
    x, lab = synth_uprj(6000, 160, 31)
    mu, gamma, steps = 0.001, 0.1, 150
    S = dlco_b200.learn.PjSweep(gpu_ctx, dev_t(gpu_ctx, x), lab, [mu], [gamma], prox=prox)
    ip, ineg, npt, nnt = O.split_train_val(lab)
    st64 = st32 = None
    worst_tf_df = worst_tf_a = 0.0
    rel = lambda a, b: float(np.abs(a - b).max() / np.abs(b).max())          # noqa: E731
    for t in range(steps):
        rank0 = int(S.ranks()[0]) if t else 0
        rng = O.CvRNG(2215)
        rng.state = int(S.rng_state.cpu().numpy().view(np.uint64)[0])
        S.refresh_w()                                      # prox="sign": W only exists on demand
        W0 = S.W[0].cpu().numpy().astype(np.float64)
        snap = dict(W=W0[:rank0] if rank0 else np.zeros_like(W0), A=S.A[0].cpu().numpy().astype(np.float64),
                    dfAvg=S.dfavg[0].cpu().numpy().astype(np.float64), t=t, rng=rng)
        S.step()
        tf = O.pj_rda_steps(x, ip, ineg, npt, nnt, mu, gamma, 1, dt=np.float64, state=snap)
        rank1 = int(S.ranks()[0])
        worst_tf_df = max(worst_tf_df, rel(S.dfavg[0].cpu().numpy(), tf["dfAvg"]))
        worst_tf_a = max(worst_tf_a, rel(S.A[0].cpu().numpy(), tf["A"]))
        assert abs(rank1 - tf["W"].shape[0]) <= 1                           # an eigenvalue within rounding of 0
        assert int(S.rng_state.cpu().numpy().view(np.uint64)[0]) == tf["rng"].state
        st64 = O.pj_rda_steps(x, ip, ineg, npt, nnt, mu, gamma, 1, dt=np.float64, state=st64)
        st32 = O.pj_rda_steps(x, ip, ineg, npt, nnt, mu, gamma, 1, dt=np.float32, state=st32)
        if t % 10 == 9:
            ref_drift = rel(st32["A"].astype(np.float64), st64["A"])
            gpu_drift = rel(S.A[0].cpu().numpy(), st64["A"])
            # floor 5e-4: one flipped hinge decision moves a short run by ~1e-4; at the BASELINE configurations the
            # fp32 ORACLE itself drifts 2.5e-3 from its fp64 twin (tests/test_gpu_baseline_configs.py prints both)
            assert gpu_drift <= max(5e-4, 4 * ref_drift), (t, gpu_drift, ref_drift)
    assert worst_tf_df <= 1e-4 and worst_tf_a <= 1e-4, (worst_tf_df, worst_tf_a)
    assert int(S.rng_state.cpu().numpy().view(np.uint64)[0]) == st64["rng"].state


@pytest.mark.parametrize("D,n_cfg", [(32, 3), (160, 2), (544, 5), (100, 1), (1400, 1)])
def test_pj_psd_project_matches_eigh(gpu_ctx, D, n_cfg):
    """the prox's projection (src/pj-learn.cpp:434-470: ssyevr, clip, V L V^T) by the tensor-core sign
    iteration: A+ within 1e-5 relative of the fp64 eigen-decomposition route (the fp32 LAPACK route of
    the reference is itself ~1e-6 away), exact rank, S^2 = I, and the trivial cases (PSD input is a
    fixed point, NSD input projects to zero, the zero matrix has rank 0)."""
    rng = np.random.default_rng(D)
    mats = []
    for g in range(n_cfg):                                  # pj-learn-shaped: -c (X^T diag(c) X / b^2 + mu I)
        xb = rng.standard_normal((400, D)) * 0.1
        c = rng.integers(-200, 200, size=400).astype(np.float64)
        a = -(np.sqrt(50.0 * g + 1.0) / 0.1) * ((xb.T * c) @ xb / 40000 + 0.001 * np.eye(D))
        mats.append(0.5 * (a + a.T))
    A = np.stack(mats).astype(np.float32)
    d_A = dev_t(gpu_ctx, A.copy())
    rank, S = gpu_ctx.pj_psd_project(d_A, want_sign=True)
    got, S = d_A.cpu().numpy(), S.cpu().numpy().astype(np.float64)
    for g in range(n_cfg):
        w, V = np.linalg.eigh(A[g].astype(np.float64))
        ref = (V * np.maximum(w, 0)) @ V.T
        # measured 4e-6 .. 8e-6 (3xTF32 products summed on the tensor core; numpy fp32 reaches 8e-7)
        assert np.abs(got[g] - ref).max() <= 2e-5 * np.abs(ref).max(), np.abs(got[g] - ref).max() / np.abs(ref).max()
        assert int(rank[g]) == int((w > 0).sum())
        assert np.abs(got[g] - got[g].T).max() == 0.0
        if np.abs(w).min() > 1e-3 * np.sqrt((w * w).sum()):          # every eigenvalue inside the design interval
            assert np.abs(S[g] @ S[g] - np.eye(D)).max() <= 1e-4
    # fixed point / annihilation / zero
    psd = np.stack([got[0], -got[0], np.zeros_like(got[0])])
    d_B = dev_t(gpu_ctx, psd.copy())
    rk, _ = gpu_ctx.pj_psd_project(d_B)
    out = d_B.cpu().numpy()
    assert np.abs(out[0] - got[0]).max() <= 3e-5 * np.abs(got[0]).max()
    assert np.abs(out[1]).max() <= 3e-5 * np.abs(got[0]).max()
    assert not out[2].any() and int(rk[2]) == 0


@pytest.mark.parametrize("D,n_cfg", [(544, 3), (480, 2), (100, 2), (72, 1)])
def test_pj_psd_project_fused_epilogues_are_bit_identical(gpu_ctx, D, n_cfg):
    """DLCO_SIGN_FUSED=1: the Y and Z operand planes of the sign iteration are written by the GEMM epilogues (directly and,
    for the tiles below the diagonal that are never computed, transposed through shared memory) instead of by separate
    passes.  Same arithmetic, so the projected matrices, ranks and sign(A) must equal the unfused route bit for bit --
    including D that is not a multiple of 16 (ragged 16-row runs of the transposed stores)."""
    rng = np.random.default_rng(7 * D)
    xb = rng.standard_normal((n_cfg, 400, D)) * 0.1
    c = rng.integers(-200, 200, size=(n_cfg, 400, 1)).astype(np.float64)
    A = -(10.0) * (np.matmul((xb * c).transpose(0, 2, 1), xb) / 40000 + 0.001 * np.eye(D))
    A = (0.5 * (A + A.transpose(0, 2, 1))).astype(np.float32)
    out = {}
    for fused in ("0", "1"):
        os.environ["DLCO_SIGN_FUSED"] = fused
        try:
            d_A = dev_t(gpu_ctx, A.copy())
            rank, S = gpu_ctx.pj_psd_project(d_A, want_sign=True)
            rank = rank.cpu().numpy() if hasattr(rank, "cpu") else np.asarray(rank)
            out[fused] = (d_A.cpu().numpy(), rank.copy(), S.cpu().numpy())
        finally:
            os.environ.pop("DLCO_SIGN_FUSED", None)
    assert np.array_equal(out["0"][1], out["1"][1])
    assert np.array_equal(out["0"][0], out["1"][0])
    assert np.array_equal(out["0"][2], out["1"][2])


def test_pj_sweep_sign_prox_matches_oracle_and_eig_route(gpu_ctx):
    """prox by the matrix-sign iteration (no eigen-decomposition per iteration) against the fp64 oracle
    (north_star bar: A within 1e-4) and against the eigen-decomposition route on the same GPU; W rebuilt on
    demand factors A, validation and FPR95 agree with the oracle."""
    import dlco_b200
    x, lab = synth_uprj(3000, 32, 4)
    b, steps = 50, 40
    mus, gammas = [0.001, 0.0005, 0.002], [0.1, 0.05, 0.2]
    d_x = dev_t(gpu_ctx, x)
    S = dlco_b200.learn.PjSweep(gpu_ctx, d_x, lab, mus, gammas, sz_batch=b, prox="sign")
    E = dlco_b200.learn.PjSweep(gpu_ctx, d_x, lab, mus, gammas, sz_batch=b, prox="eig")
    for _ in range(steps):
        S.step()
        E.step()
    ranks = S.ranks()
    ip, ineg, npt, nnt = O.split_train_val(lab)
    for c in range(3):
        st = O.pj_rda_steps(x, ip, ineg, npt, nnt, mus[c], gammas[c], steps, sz_batch=b, dt=np.float64)
        A_g = S.A[c].cpu().numpy()
        same_path = assert_a_matches_oracle(gpu_ctx, A_g, st, x, lab, mus[c], gammas[c], steps, b, "sign")
        assert_a_matches_oracle(gpu_ctx, E.A[c].cpu().numpy(), st, x, lab, mus[c], gammas[c], steps, b, "eig")
        assert abs(int(ranks[c]) - st["W"].shape[0]) <= (0 if same_path else 2)
        loss, regul = S.validate(c, ranks[c])              # rebuilds W from A
        W_g = S.W[c].cpu().numpy()[:ranks[c]]
        assert np.abs(W_g.T @ W_g - A_g).max() <= 3e-5 * np.abs(A_g).max()
        assert not S.W[c, ranks[c]:].any().item()
        lo, ro = O.pj_validate(x[ip[npt:]], x[ineg[nnt:]], st["W"], st["A"], mus[c])
        tol = 1e-4 if same_path else 5e-3
        assert abs(loss - lo) <= tol * lo and abs(regul - ro) <= tol * ro
        fpr_s, auc_s = S.stats(c, ranks[c])
        fpr_e, auc_e = E.stats(c, int(E.ranks()[c]))
        assert abs(fpr_s - fpr_e) <= 1e-3 and abs(auc_s - auc_e) <= 1e-4      # FPR95 within 0.1 percentage points
    assert int(S.rng_state.cpu().numpy().view(np.uint64)[0]) == st["rng"].state


def test_sign_sym_kernel_matches_the_generic_prep(gpu_ctx, tmp_path):
    """k_sign_sym_planes (csrc/pj_sign.cu: sym(T) written as planes block and mirror block at once) against the generic
    k_sign_prep<PREP_SYM> it replaced (DLCO_SIGN_SYM_GENERIC=1, read once per process): the projected matrices, sign(A)
    and the ranks bit for bit, fused and unfused sequences, D a multiple of 128 or not."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = {}
    for name, env in (("new", {}), ("generic", {"DLCO_SIGN_SYM_GENERIC": "1"})):
        path = str(tmp_path / (name + ".npz"))
        e = dict(os.environ)
        e.update(env)
        r = subprocess.run([sys.executable, os.path.join(root, "tools", "diag", "psd_dump.py"), path], env=e,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=300)
        assert r.returncode == 0, r.stdout.decode()[-2000:]
        res[name] = np.load(path)
    assert sorted(res["new"].files) == sorted(res["generic"].files) and len(res["new"].files) == 12
    for k in res["new"].files:
        assert np.isfinite(res["new"][k]).all(), k
        assert np.array_equal(res["new"][k], res["generic"][k]), k
