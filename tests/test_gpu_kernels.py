"""GPU parity tests: every call goes through the C ABI (libdlco_b200.so via ctypes) and is
compared with the CPU oracle (oracle/dlco_oracle.py, restating the reference file:line cited
in each oracle function).  Tolerances follow BASELINE.json's north_star: distances 1e-5
relative (to the row's largest distance -- SURVEY.md H3), learnt w / A 1e-4 relative, FPR95
0.1 percentage points; integer / index results are exact."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
from conftest import make_bank  # noqa: E402
from oracle import dlco_oracle as O  # noqa: E402

DIST_RTOL = 1e-5


def dev_t(ctx, a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(ctx.device)


# ------------------------------------------------------------------------------------------
# get_desc: src/vgg-desc.cpp:41-152
# ------------------------------------------------------------------------------------------
def test_get_desc_matches_oracle(gpu_ctx, synth):
    patches, _, _ = synth
    ids = torch.arange(patches.shape[0], dtype=torch.int32, device=gpu_ctx.device)
    pt = gpu_ctx.get_desc(dev_t(gpu_ctx, patches), ids).cpu().numpy()
    ref64 = np.stack([O.get_desc(p, dt=np.float64) for p in patches])
    ref32 = np.stack([O.get_desc(p) for p in patches])
    assert pt.shape == ref64.shape == (patches.shape[0], 4096, 8)
    # at most two non-zero bins per pixel, non-negative
    assert (np.count_nonzero(pt, axis=2) <= 2).all() and (pt >= 0).all()
    # the fp32 evaluation of the reference differs from exact arithmetic by the conditioning of
    # atan2 / the soft-bin offset: the GPU must be as close to fp64 as the fp32 oracle is
    scale = ref64.max()
    err_gpu = np.abs(pt - ref64).max() / scale
    err_o32 = np.abs(ref32 - ref64).max() / scale
    assert err_gpu <= max(2.0 * err_o32, 2e-5), (err_gpu, err_o32)
    # quantile normalisation: the 0.8-quantile of the magnitude maps to n_bins = 8
    mag = pt.sum(axis=2)
    q = np.array([O.quantile_threshold(m.astype(np.float32)) for m in mag])
    np.testing.assert_allclose(q, 8.0, rtol=1e-5)


def test_get_desc_subset_and_repeat_ids(gpu_ctx, synth):
    patches, _, _ = synth
    ids = torch.tensor([5, 0, 5, 17], dtype=torch.int32, device=gpu_ctx.device)
    pt = gpu_ctx.get_desc(dev_t(gpu_ctx, patches), ids).cpu().numpy()
    assert np.array_equal(pt[0], pt[2])
    ref = O.get_desc(patches[17], dt=np.float64)
    assert np.abs(pt[3] - ref).max() <= 5e-5 * ref.max()


def test_get_desc_constant_patch_is_zero(gpu_ctx):
    """T == 0 branch of src/vgg-desc.cpp:131: a flat patch has zero gradient everywhere."""
    patches = np.full((1, 64, 64), 77, np.uint8)
    ids = torch.zeros(1, dtype=torch.int32, device=gpu_ctx.device)
    pt = gpu_ctx.get_desc(dev_t(gpu_ctx, patches), ids).cpu().numpy()
    assert np.isfinite(pt).all() and np.abs(pt).max() == 0.0


def test_get_desc_matches_cv2_driven_fixture(gpu_ctx):
    """k_desc against OpenCV's own get_desc (tests/golden/desc_kat.npz: cv2 4.13 driven through the
    reference's call sequence).  k_desc follows OpenCV's evaluation order (fused row taps, symmetric fused
    column taps, sqrt(fma) magnitude, double atan2 + CV_PI, fused scaled converts), so the only freedom left
    is CUDA's double atan2 against glibc's in the last bit of a double: bins identical, values bit-exact
    except where that bit decides a float rounding."""
    from test_oracle_golden import kat_patchtrans, load_desc_kat
    patches, v1, v2, b1, _ = load_desc_kat("pt")
    want = kat_patchtrans(v1, v2, b1)
    ids = torch.arange(patches.shape[0], dtype=torch.int32, device=gpu_ctx.device)
    got = gpu_ctx.get_desc(dev_t(gpu_ctx, patches), ids).cpu().numpy()
    assert got.shape == want.shape
    assert np.array_equal(got != 0, want != 0)                            # same bins, same zeros
    np.testing.assert_allclose(got, want, rtol=3e-7, atol=0)              # never more than 2 ulp
    exact = (got.view(np.uint32) == want.view(np.uint32)).mean()
    assert exact >= 0.9999, exact
    # the two other variants of the fixture (atan2f overload; IPP's approximate magnitude) bracket what
    # another build of the reference would produce: a few ulp, no structural change
    for variant in ("pt_atan2f", "pt_ipp"):
        _, a1, a2, ab1, _ = load_desc_kat(variant)
        alt = kat_patchtrans(a1, a2, ab1)
        assert np.abs(got - alt).max() <= 1e-6 * max(1.0, np.abs(alt).max())


# ------------------------------------------------------------------------------------------
# comp-fulldists: src/comp-fulldists.cpp:294-356
# ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def small_bank(oracle_lib):
    rings = O.ring_params()
    sel = np.r_[0:8, 700:708, 2500:2508, 5112:5120]          # 32 rings = 256 filters = 2 M tiles
    filt, prp = make_bank(oracle_lib, rings[sel])
    return filt, prp


def _want_fulldist(filt, pt, idx):
    want = np.zeros((len(idx), filt.shape[0] // 8))
    for k, row in enumerate(idx):
        want[k] = O.full_distance(filt, pt[row[0]], pt[row[2]])
    return want


@pytest.mark.parametrize("impl", ["tc", "simt"])
def test_fulldist_matches_oracle(gpu_ctx, synth, small_bank, impl):
    patches, _, idx = synth
    filt, _ = small_bank
    d_p, d_i = dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)
    ids = torch.arange(patches.shape[0], dtype=torch.int32, device=gpu_ctx.device)
    pt = gpu_ctx.get_desc(d_p, ids).cpu().numpy()
    want = _want_fulldist(filt, pt, idx)                       # fp64 pooling of the GPU descriptors
    bank = gpu_ctx.bank(filt)
    os.environ["DLCO_POOL_IMPL"] = impl
    try:
        got = gpu_ctx.fulldist_build(bank, d_p, d_i).cpu().numpy()
    finally:
        os.environ.pop("DLCO_POOL_IMPL", None)
        bank.close()
    assert np.isfinite(got).all()
    rel = np.abs(got - want).max(axis=1) / want.max(axis=1)
    assert rel.max() <= DIST_RTOL, rel.max()
    # end to end against the oracle's own (cv2-pinned) fp32 descriptors: the same bar
    want_e2e = O.full_distances(filt, patches, idx)
    rel = np.abs(got - want_e2e).max(axis=1) / want_e2e.max(axis=1)
    assert rel.max() <= DIST_RTOL, rel.max()


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 33])
def test_fulldist_ragged_ranges(gpu_ctx, synth, small_bank, n):
    patches, _, idx = synth
    filt, _ = small_bank
    d_p, d_i = dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)
    bank = gpu_ctx.bank(filt)
    full = gpu_ctx.fulldist_build(bank, d_p, d_i).cpu().numpy()
    lo = 3
    got = gpu_ctx.fulldist_build(bank, d_p, d_i, lo, lo + n).cpu().numpy()
    bank.close()
    assert got.shape == (n, filt.shape[0] // 8)
    assert np.array_equal(got, full[lo:lo + n])               # batching must not change a single bit


def test_fulldist_host_entry_point(gpu_ctx, synth, small_bank):
    patches, _, idx = synth
    filt, _ = small_bank
    bank = gpu_ctx.bank(filt)
    dev = gpu_ctx.fulldist_build(bank, dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)).cpu().numpy()
    host = gpu_ctx.fulldist_build_host(bank, torch.from_numpy(patches), torch.from_numpy(idx)).numpy()
    bank.close()
    assert np.array_equal(dev, host)


def test_fulldist_host_async_blocks_in_flight(gpu_ctx, synth, small_bank):
    """dlco_fulldist_build_host_async + dlco_fulldist_host_wait: several blocks queued back to back (the staging
    slots keep alternating across calls), waited for one by one, bit-identical to the device entry point."""
    patches, _, idx0 = synth
    filt, _ = small_bank
    rng = np.random.default_rng(5)
    n = 1100
    idx = np.zeros((n, 4), np.int32)
    idx[:, 0] = rng.integers(0, patches.shape[0], n)
    idx[:, 2] = rng.integers(0, patches.shape[0], n)
    bank = gpu_ctx.bank(filt)
    want = gpu_ctx.fulldist_build(bank, dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)).cpu().numpy()
    h_p, h_i = torch.from_numpy(patches).pin_memory(), torch.from_numpy(idx).pin_memory()
    out = torch.zeros((n, filt.shape[0] // 8), dtype=torch.float32).pin_memory()
    cuts = [0, 300, 301, 301, 900, n]                        # ragged blocks, an empty one in between
    tickets = [gpu_ctx.fulldist_build_host_async(bank, h_p, h_i, lo, hi, out[lo:hi]) for lo, hi in zip(cuts[:-1], cuts[1:])]
    for (lo, hi), t in zip(zip(cuts[:-1], cuts[1:]), tickets):
        gpu_ctx.fulldist_host_wait(t)
        assert np.array_equal(out[lo:hi].numpy(), want[lo:hi])           # complete although later blocks are in flight
    gpu_ctx.sync()
    bank.close()
    assert np.array_equal(out.numpy(), want)


def test_fulldist_full_bank_three_batches(gpu_ctx, synth, full_bank):
    """The gen-poolregion bank at full size and more pairs than two batches: exercises the block-sparse bank order
    (8 x 8 pixel-block K chunks, clustered rows), the cached work schedule of the pooling kernel and the
    three-stream host path with both staging slots reused.  Checked against the fp64 oracle on a spread of rings,
    bit for bit between the device and host entry points and between one call and ragged sub-ranges."""
    patches, _, idx0 = synth
    _, filt, _ = full_bank
    rng = np.random.default_rng(11)
    n = 600                                                    # 256 + 256 + 88
    idx = np.zeros((n, 4), np.int32)
    idx[:, 0] = rng.integers(0, patches.shape[0], n)
    idx[:, 2] = rng.integers(0, patches.shape[0], n)
    idx[:40] = idx0
    d_p, d_i = dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)
    bank = gpu_ctx.bank(filt)
    info = bank.info()
    assert info["filters"] == 40960 and info["rows"] == 31776 and info["tiles"] == 249
    assert info["k_chunks"] <= 9800, info                      # x*64+y order sorted by row range: 11 571
    got = gpu_ctx.fulldist_build(bank, d_p, d_i).cpu().numpy()
    host = gpu_ctx.fulldist_build_host(bank, torch.from_numpy(patches), torch.from_numpy(idx)).numpy()
    part = gpu_ctx.fulldist_build(bank, d_p, d_i, 250, 530).cpu().numpy()
    bank.close()
    assert np.isfinite(got).all()
    assert np.array_equal(got, host)
    assert np.array_equal(part, got[250:530])
    ids = torch.arange(patches.shape[0], dtype=torch.int32, device=gpu_ctx.device)
    pt = gpu_ctx.get_desc(d_p, ids).cpu().numpy()
    rings = np.r_[0:5120:97, 5119]                             # all sigmas, radii and angles show up
    rows = (rings[:, None] * 8 + np.arange(8)[None, :]).reshape(-1)
    for k in (0, 7, 255, 256, 511, 512, 599):
        want = O.full_distance(filt[rows], pt[idx[k, 0]], pt[idx[k, 2]])
        scale = max(got[k].max(), 1e-30)
        assert np.abs(got[k, rings] - want).max() / scale <= DIST_RTOL, k


def test_pool_kernel_variants_are_bit_identical(gpu_ctx, tmp_path):
    """The pooling kernel's tuning knobs are read once per process: whole-chunk stages (DLCO_POOL_SPLIT=1), other
    pair-tile groupings of the work schedule (DLCO_POOL_GROUP_TILES), get_desc over several batches
    (DLCO_FULLDIST_DESC_GROUP) and the SIMT cross-check path must all give the default's distances -- the first three
    bit for bit (same products, same accumulation order), the SIMT path within the distance bar."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = {}
    for name, env in (("default", {}), ("split1", {"DLCO_POOL_SPLIT": "1"}), ("group3", {"DLCO_POOL_GROUP_TILES": "3"}),
                      ("group16", {"DLCO_POOL_GROUP_TILES": "16"}), ("desc3", {"DLCO_FULLDIST_DESC_GROUP": "3"})):
        path = str(tmp_path / (name + ".npy"))
        e = dict(os.environ)
        e.update(env)
        r = subprocess.run([sys.executable, os.path.join(root, "tools", "pool_ab.py"), path, "--quick"], env=e,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=300)
        assert r.returncode == 0, r.stdout.decode()[-2000:]
        outs[name] = np.load(path)
    assert np.isfinite(outs["default"]).all() and outs["default"].shape == (600, 5120)
    for name in ("split1", "group3", "group16", "desc3"):
        assert np.array_equal(outs[name], outs["default"]), name


def test_fulldist_identical_patches_give_zero(gpu_ctx, synth, small_bank):
    patches, pid, _ = synth
    filt, _ = small_bank
    idx = np.array([[4, pid[4], 4, pid[4]], [9, pid[9], 9, pid[9]]], np.int32)
    bank = gpu_ctx.bank(filt)
    got = gpu_ctx.fulldist_build(bank, dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)).cpu().numpy()
    bank.close()
    assert np.abs(got).max() == 0.0


def test_fulldist_symmetric_in_the_pair(gpu_ctx, synth, small_bank):
    patches, _, idx = synth
    filt, _ = small_bank
    swapped = idx[:, [2, 3, 0, 1]].copy()
    bank = gpu_ctx.bank(filt)
    a = gpu_ctx.fulldist_build(bank, dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)).cpu().numpy()
    b = gpu_ctx.fulldist_build(bank, dev_t(gpu_ctx, patches), dev_t(gpu_ctx, swapped)).cpu().numpy()
    bank.close()
    assert np.array_equal(a, b)                               # (p2 - p1)^2 == (p1 - p2)^2 bit for bit


def test_pair_labels(gpu_ctx, synth):
    _, _, idx = synth
    got = gpu_ctx.pair_labels(dev_t(gpu_ctx, idx)).cpu().numpy()
    assert np.array_equal(got, O.pair_labels(idx))


# ------------------------------------------------------------------------------------------
# gen-poolregion on the device: src/gen-poolregion.cpp:51-80,163-295
# ------------------------------------------------------------------------------------------
def test_gen_pr_filters_bit_exact_full_bank(gpu_ctx, full_bank):
    """All 40 960 filters generated on the device equal the C oracle's (pinned to workspace/opencv/*.i) bit for
    bit; PRParams likewise; the ring grid from the library's host loops equals the oracle's."""
    import dlco_b200
    rings, filt, prp = full_bank
    assert np.array_equal(dlco_b200.ring_params(), rings)
    d, p, fixed = gpu_ctx.gen_pr_filters(rings)
    got = d.cpu().numpy()
    assert fixed <= 64, fixed                                  # host-libm regenerations: rounding ties only
    assert np.array_equal(p, prp)
    bad = np.nonzero((got.view(np.uint32) != filt.view(np.uint32)).any(axis=1))[0]
    assert len(bad) == 0, (len(bad), bad[:8])


def test_bank_analytic_matches_uploaded_bank(gpu_ctx, synth, full_bank):
    patches, _, idx = synth
    rings, filt, _ = full_bank
    d_p, d_i = dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx)
    up = gpu_ctx.bank(filt)
    an = gpu_ctx.bank_analytic(rings)
    assert an.info() == up.info()
    a = gpu_ctx.fulldist_build(up, d_p, d_i).cpu().numpy()
    b = gpu_ctx.fulldist_build(an, d_p, d_i).cpu().numpy()
    up.close()
    an.close()
    assert np.array_equal(a, b)
    # a device-resident bank (dlco_bank_create_dev) goes the same way: no host copy of the filters
    sub = gpu_ctx.bank(torch.from_numpy(filt[:512]).to(gpu_ctx.device))
    sub2 = gpu_ctx.bank(filt[:512])
    assert sub.info() == sub2.info()
    a = gpu_ctx.fulldist_build(sub, d_p, d_i).cpu().numpy()
    b = gpu_ctx.fulldist_build(sub2, d_p, d_i).cpu().numpy()
    sub.close()
    sub2.close()
    assert np.array_equal(a, b)


# ------------------------------------------------------------------------------------------
# BASELINE configs[0]: 10 000 pairs x the full gen-poolregion bank, end to end at north_star's 1e-5
# ------------------------------------------------------------------------------------------
def _pool_fp64_torch(filt64, pt1, pt2, uprj=False):
    """The oracle's pooling formula (oracle/dlco_oracle.py full_distance / uprj_distance, i.e.
    src/comp-fulldists.cpp:328-342, src/comp-uprjdists.cpp:320-329) as fp64 torch matmuls: the checker for
    sizes numpy cannot finish in seconds; cross-checked against the numpy oracle inside the tests."""
    n = pt1.shape[0]
    a = torch.clamp_max(torch.matmul(filt64, pt1.permute(1, 0, 2).reshape(4096, n * 8)), 1.0)
    b = torch.clamp_max(torch.matmul(filt64, pt2.permute(1, 0, 2).reshape(4096, n * 8)), 1.0)
    if uprj:
        return (a - b).reshape(-1, n, 8).permute(1, 0, 2).reshape(n, -1)   # Desc1 - Desc2, [filter][bin]
    d = ((b - a) ** 2).reshape(-1, n, 8).sum(dim=2)                        # per filter
    return d.reshape(-1, 8, n).sum(dim=1).t().contiguous()                 # 8 members per ring


@pytest.fixture(scope="module")
def configs0(gpu_ctx):
    """10 000 synthetic pairs over 10 002 patches; descriptors from the cv2-pinned fp32 oracle."""
    patches, pid = O.synth_patches(3334, per_point=3, seed=20240601)
    idx = O.synth_pairs(pid, 10000, seed=20240602)
    used = np.unique(np.concatenate([idx[:, 0], idx[:, 2]]))
    pt = np.zeros((patches.shape[0], 4096, 8), np.float32)
    for i in used:
        pt[i] = O.get_desc(patches[i])
    return patches, idx, pt


def test_fulldist_configs0_end_to_end(gpu_ctx, configs0, full_bank):
    patches, idx, pt = configs0
    _, filt, _ = full_bank
    dev = gpu_ctx.device
    bank = gpu_ctx.bank(filt)
    got = gpu_ctx.fulldist_build(bank, dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx))
    torch.cuda.synchronize()
    bank.close()
    assert got.shape == (10000, 5120) and bool(torch.isfinite(got).all())
    f64 = torch.from_numpy(filt).to(dev).double()
    d_pt = torch.from_numpy(pt).to(dev)
    worst, worst_abs = 0.0, 0.0
    for lo in range(0, len(idx), 500):
        sl = idx[lo:lo + 500]
        want = _pool_fp64_torch(f64, d_pt[torch.from_numpy(sl[:, 0]).long().to(dev)].double(),
                                d_pt[torch.from_numpy(sl[:, 2]).long().to(dev)].double())
        if lo == 0:                                                        # the torch checker == the numpy oracle
            w0 = O.full_distance(filt[np.r_[0:8, 8 * 2600:8 * 2600 + 8]], pt[sl[3, 0]], pt[sl[3, 2]])
            np.testing.assert_allclose(want[3, [0, 2600]].cpu().numpy(), w0, rtol=1e-12)
        g = got[lo:lo + 500].double()
        rel = (g - want).abs().amax(dim=1) / want.amax(dim=1)
        worst = max(worst, float(rel.max()))
        worst_abs = max(worst_abs, float((g - want).abs().max()))
    print("configs[0] full distances: worst row-relative error %.3g (abs %.3g)" % (worst, worst_abs))
    assert worst <= DIST_RTOL, worst


def test_uprjdist_configs0_end_to_end(gpu_ctx, configs0, full_bank):
    """comp-uprjdists on the same 10 000 pairs with the 68 filters of the shipped liberty model
    (tests/golden/pr_select_kat.json), against fp64 pooling of the oracle's descriptors."""
    import json
    patches, idx, pt = configs0
    _, filt, _ = full_bank
    kat = [k for k in json.load(open(os.path.join(os.path.dirname(__file__), "golden", "pr_select_kat.json")))
           if k["rows"] == 68][0]
    w = np.zeros(5120, np.float32)
    w[kat["rings"]] = kat["w"]
    sel = O.select_pr_filters(filt, w)
    assert sel.shape == (68, 4096)
    dev = gpu_ctx.device
    bank = gpu_ctx.bank(sel)
    got = gpu_ctx.uprjdist_build(bank, dev_t(gpu_ctx, patches), dev_t(gpu_ctx, idx))
    torch.cuda.synchronize()
    bank.close()
    assert got.shape == (10000, 544)
    f64 = torch.from_numpy(sel).to(dev).double()
    d_pt = torch.from_numpy(pt).to(dev)
    i1, i2 = torch.from_numpy(idx[:, 0]).long().to(dev), torch.from_numpy(idx[:, 2]).long().to(dev)
    worst = 0.0
    for lo in range(0, len(idx), 2000):
        want = _pool_fp64_torch(f64, d_pt[i1[lo:lo + 2000]].double(), d_pt[i2[lo:lo + 2000]].double(), uprj=True)
        if lo == 0:
            np.testing.assert_allclose(want[5].cpu().numpy(), O.uprj_distance(sel, pt[idx[5, 0]], pt[idx[5, 2]]),
                                       rtol=1e-12, atol=1e-15)
        g = got[lo:lo + 2000].double()
        rel = (g - want).abs().amax(dim=1) / want.abs().amax(dim=1)
        worst = max(worst, float(rel.max()))
    print("configs[0] unprojected distances: worst row-relative error %.3g" % worst)
    assert worst <= DIST_RTOL, worst
