"""Pins the CPU oracle (oracle/dlco_oracle.py, oracle/dlco_oracle.c, oracle/h5lite.py) against the
reference tree's own committed run outputs (fixtures in tests/golden/, made by
tests/golden/make_golden.py from /root/reference/workspace) and against the OpenCV primitives
the reference calls (cv2 4.13 outputs, also fixtures).  CPU only."""
import hashlib
import json
import os
import re

import numpy as np
import pytest

from conftest import make_bank
from oracle import dlco_oracle as O
from oracle import h5lite

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---- gen-poolregion + SelectPRFilters vs workspace/opencv/vgg_generated_*.i (bit-exact) --------
@pytest.mark.parametrize("k", range(4))
def test_selected_filters_match_exported_headers(k, oracle_lib):
    kat = json.load(open(os.path.join(G, "pr_select_kat.json")))[k]
    rings = O.ring_params()
    assert rings.shape == (5120, 3)
    sel_rings = kat["rings"]
    filt, _ = make_bank(oracle_lib, rings[sel_rings])            # C oracle: 8 filters per ring
    full = np.zeros((5120 * 8, 4096), np.float32)
    w = np.zeros(5120, np.float32)
    for j, r in enumerate(sel_rings):
        full[8 * r:8 * r + 8] = filt[8 * j:8 * j + 8]
        w[r] = kat["w"][j]
    sel = O.select_pr_filters(full, w)
    assert sel.shape == (kat["rows"], 4096)
    assert hashlib.sha256(sel.tobytes()).hexdigest() == kat["sha256"]


def test_python_and_c_filter_restatements_agree(oracle_lib):
    rings = O.ring_params()[[0, 37, 1234, 5119]]
    f_c, p_c = make_bank(oracle_lib, rings)
    f_py, p_py = O.pr_filters_for_rings(rings)
    assert np.array_equal(f_c, f_py) and np.array_equal(p_c, p_py)
    np.testing.assert_allclose(f_c.sum(axis=1), 1.0, rtol=1e-5)   # unit-sum normalisation


def test_ring_grid():
    r = O.ring_params()
    assert np.array_equal(np.unique(r[:, 2]), np.arange(1, 33) * 0.5)       # 32 sigmas
    assert len(np.unique(r[:, 1])) == 5 and len(np.unique(r[:, 0])) == 32
    assert np.array_equal(r[0], [0.0, 0.0, 0.5]) and r[1, 0] == 1.0         # r fastest, then phi, then sigma


# ---- pr-learn outputs: w rows <-> log lines ----------------------------------------------------
def test_h5lite_reads_reference_pr_file_and_log_consistency():
    kat = json.load(open(os.path.join(G, "pr_w_rows.json")))
    f = h5lite.H5File(os.path.join(G, kat["file"]))
    info = f.info("w")
    assert info["shape"] == tuple(kat["shape"]) and info["dtype"] == np.float32
    assert info["chunk"] == (1, 5120) and info["filters"][0][0] == 1        # chunk 1 x 5120, deflate
    assert info["maxshape"][0] == 0xFFFFFFFFFFFFFFFF                        # unlimited rows
    w = f.read("w")
    for row, gold, lines in zip(w, kat["rows"], kat["saved_lines"]):
        assert np.array_equal(np.nonzero(row)[0], gold["idx"])
        assert np.array_equal(row[gold["idx"]], np.array(gold["val"], np.float32))
        m = re.match(r"Best: (\d+)  Loss: ([\d.]+) Regul: ([\d.]+) Obj: ([\d.]+) \(([\d.]+)\)  NNZ: (\d+) \((\d+)\)",
                     lines["best"])
        assert m, lines["best"]
        regul, nnz = float(m.group(3)), int(m.group(6))
        assert nnz == len(gold["idx"])
        # Regul = mu * sum|w| (src/pr-learn.cpp:358), printed with 6 decimals
        assert abs(kat["mu"] * np.abs(row.astype(np.float64)).sum() - regul) < 1e-6
        s = re.match(r"Stat: nPR #(\d+) \(#(\d+)\) Dim/MaxDim \[(\d+)/(\d+)\]", lines["stat"])
        n_pr, nz_dim, dim = int(s.group(1)), int(s.group(2)), int(s.group(3))
        # oracle's ComputePRStats bookkeeping (incl. the nPR quirk) reproduces the logged numbers
        prp = O.pr_params(O.ring_params())
        assert O.pr_dims(prp, row) == (n_pr, dim, nz_dim)


# ---- pj-learn outputs: W, A conventions ---------------------------------------------------------
def test_pj_model_conventions():
    z = np.load(os.path.join(G, "pj_model_kat.npz"))
    W, A_rows = z["W"], z["A_rows"]
    D = int(z["A_shape"][0])
    assert W.shape[1] == D
    A = W.astype(np.float64).T @ W.astype(np.float64)                      # A = W^T W
    assert np.abs(A[::8] - A_rows).max() <= 2e-6 * np.abs(A).max()
    assert float(z["A_sym_err"]) <= 1e-6 * np.abs(A).max()
    norms = (W.astype(np.float64) ** 2).sum(axis=1)                        # rows sqrt(e_k) v_k, ascending e_k
    assert np.all(np.diff(norms) >= -1e-5 * norms.max())
    G_ = W.astype(np.float64) @ W.astype(np.float64).T                     # orthogonal eigenvectors
    off = G_ - np.diag(np.diag(G_))
    assert np.abs(off).max() <= 1e-4 * norms.max()
    # the oracle's prox reproduces W from A up to the sign of each eigenvector
    A_full = A.astype(np.float32)
    A2, W2, e = O.pj_prox(A_full, np.float64)
    keep = e > 1e-6 * e.max()
    assert keep.sum() == W.shape[0]
    W2 = W2[-W.shape[0]:]
    sign = np.sign((W2 * W).sum(axis=1))
    assert np.abs(W2 * sign[:, None] - W).max() <= 2e-3 * np.abs(W).max()
    # log: Regul = mu * trace(A) of the last saved line, Dim = rows of W
    m = re.match(r"Stat: Dim \[(\d+)\]", str(z["last_saved"]))
    assert int(m.group(1)) == W.shape[0]


# ---- OpenCV primitives (cv2 outputs stored by make_golden.py) -----------------------------------
def test_gaussian_blur_and_gradients_match_cv2():
    z = np.load(os.path.join(G, "cv2_kat.npz"))
    gk = O.gaussian_kernel_1d(1.4)
    assert len(gk) == 13 and np.array_equal(gk, z["gk"])
    blur = O._sep_blur_replicate(z["img"], gk)
    assert np.abs(blur - z["blur"]).max() <= 3e-5 * 255                    # accumulation order only
    px = np.pad(z["blur"], ((0, 0), (1, 1)), mode="edge")
    py = np.pad(z["blur"], ((1, 1), (0, 0)), mode="edge")
    assert np.array_equal(px[:, 2:] - px[:, :-2], z["ix"])
    assert np.array_equal(py[2:, :] - py[:-2, :], z["iy"])
    mag = np.sqrt(z["ix"] * z["ix"] + z["iy"] * z["iy"]).astype(np.float32)
    assert np.abs(mag - z["mag"]).max() <= 1e-6 * mag.max()


def test_rand_shuffle_matches_cv2_fresh_process():
    z = np.load(os.path.join(G, "cv2_kat.npz"))
    g = O.CvRNG()
    a = O.rand_shuffle(list(range(1000)), g)
    b = O.rand_shuffle(list(range(777)), g)                                # same stream continues
    assert np.array_equal(a, z["shuf_a"]) and np.array_equal(b, z["shuf_b"])


def test_contour_area_matches_cv2():
    z = np.load(os.path.join(G, "cv2_kat.npz"))
    pts = z["pts"].astype(np.float64)
    x, y = pts[:, 0], pts[:, 1]
    xp, yp = np.roll(x, 1), np.roll(y, 1)
    assert abs(abs(np.sum(xp * y - yp * x)) * 0.5 - float(z["area"])) <= 1e-12


# ---- get_desc as a whole vs OpenCV itself (tests/golden/make_desc_golden.py drove cv2 4.13 through the
# reference's call sequence, src/vgg-desc.cpp:41-152) ------------------------------------------------
def load_desc_kat(variant="pt"):
    """-> patches, v1, v2, bin1, T of the fixture variant (pt | pt_atan2f | pt_ipp), rows p = x*64 + y."""
    z = np.load(os.path.join(G, "desc_kat.npz"))
    v1, v2, b1 = z["pt_v1"].copy(), z["pt_v2"].copy(), z["pt_bin1"].copy()
    if variant != "pt":
        for fld, a in (("v1", v1), ("v2", v2), ("bin1", b1)):
            a.reshape(-1)[z[variant + "_" + fld + "_idx"]] = z[variant + "_" + fld + "_val"]
    return z["patches"], v1, v2, b1, z[variant + "_T"]


def kat_patchtrans(v1, v2, b1):
    """(n, 4096, 8) PatchTrans from the fixture's sparse form (Bin2 assignment second, :143-146)."""
    n = v1.shape[0]
    pt = np.zeros((n, 4096, 8), np.float32)
    i, p = np.meshgrid(np.arange(n), np.arange(4096), indexing="ij")
    pt[i, p, b1] = v1
    pt[i, p, (b1 + 1) % 8] = v2
    return pt


@pytest.mark.parametrize("variant", ["pt", "pt_atan2f"])
def test_get_desc_bit_exact_vs_cv2_driven_fixture(variant):
    """The fp32 oracle reproduces OpenCV's get_desc to the bit: blur order, magnitude, angle, soft bins,
    cv::sort quantile, Mat/=s scaling and the transposed flatten -- on smooth, flat, ramp, saturated,
    single-pixel and noise patches."""
    patches, v1, v2, b1, T = load_desc_kat(variant)
    assert len(patches) >= 64
    for i, patch in enumerate(patches):
        o1, o2, ob1, ob2, t = O.get_desc_parts(patch, atan2_float=(variant == "pt_atan2f"))
        assert t == T[i], i
        assert np.array_equal(ob1.T.reshape(-1), b1[i]), i
        assert np.array_equal(o1.T.reshape(-1).view(np.uint32), v1[i].view(np.uint32)), i
        assert np.array_equal(o2.T.reshape(-1).view(np.uint32), v2[i].view(np.uint32)), i
        assert np.array_equal(ob2, (ob1 + 1) % 8)
    pt = kat_patchtrans(v1[:3], v2[:3], b1[:3])
    for i in range(3):
        assert np.array_equal(O.get_desc(patches[i], atan2_float=(variant == "pt_atan2f")), pt[i])


def test_get_desc_ipp_magnitude_variant_is_ulps_away():
    """With IPP (the cv2 wheel's default) cv::magnitude is an approximate kernel: the descriptor moves by
    a few ulp, the bins and the structure do not."""
    patches, v1, v2, b1, T = load_desc_kat("pt_ipp")
    _, w1, w2, wb1, wT = load_desc_kat("pt")
    assert np.array_equal(b1, wb1)
    np.testing.assert_allclose(T, wT, rtol=4e-7)
    for a, b in ((v1, w1), (v2, w2)):
        assert np.abs(a - b).max() <= 6e-7 * max(np.abs(b).max(), 1.0)
        np.testing.assert_allclose(a, b, rtol=1e-6, atol=1e-30)


def test_get_desc_fp64_truth_close_to_fixture():
    """Error budget of the reference's own fp32 arithmetic against exact arithmetic (fp64 mode)."""
    patches, v1, v2, b1, _ = load_desc_kat("pt")
    worst = 0.0
    for i in range(0, 48, 4):                                              # the smooth synthetic patches
        pt64 = O.get_desc(patches[i], dt=np.float64)
        pt32 = kat_patchtrans(v1[i:i + 1], v2[i:i + 1], b1[i:i + 1])[0]
        worst = max(worst, np.abs(pt32 - pt64).max() / pt64.max())
    assert worst <= 2e-5, worst


# ---- internal consistency of the oracle ----------------------------------------------------------
def test_hinge_pairsum_sorted_equals_bruteforce(oracle_lib):
    import ctypes
    rng = np.random.default_rng(0)
    pos = rng.normal(0, 1, 300).astype(np.float32)
    neg = rng.normal(0.3, 1, 411).astype(np.float32)
    a, b = O.hinge_pairsum(pos, neg), O.hinge_pairsum_bruteforce(pos, neg)
    c = oracle_lib.dlco_oracle_hinge_pairsum(pos.ctypes.data_as(ctypes.c_void_p), 300,
                                             neg.ctypes.data_as(ctypes.c_void_p), 411, 1.0)
    assert abs(a - b) <= 1e-9 * b and abs(c - b) <= 1e-9 * b


def test_get_desc_structure(synth):
    patches, _, _ = synth
    pt = O.get_desc(patches[0])
    assert pt.shape == (4096, 8) and pt.dtype == np.float32
    assert (np.count_nonzero(pt, axis=1) <= 2).all()
    nz = np.nonzero(pt[100])[0]
    if len(nz) == 2:                                                        # adjacent bins (circular)
        assert (nz[1] - nz[0]) in (1, 7)
    # transposed flatten: row p is pixel (r, c) = (p % 64, p // 64): transposing the patch permutes rows
    pt_t = O.get_desc(patches[0].T.copy())
    assert pt_t.shape == pt.shape


def test_full_distance_c_and_numpy_agree(oracle_lib, synth):
    import ctypes
    patches, _, idx = synth
    rings = O.ring_params()[[10, 3000]]
    filt, _ = make_bank(oracle_lib, rings)
    t1, t2 = O.get_desc(patches[idx[0, 0]]), O.get_desc(patches[idx[0, 2]])
    want = O.full_distance(filt, t1, t2)
    got = np.zeros(2)
    oracle_lib.dlco_oracle_pool_pair(16, 4096, filt.ctypes.data_as(ctypes.c_void_p),
                                     np.ascontiguousarray(t1).ctypes.data_as(ctypes.c_void_p),
                                     np.ascontiguousarray(t2).ctypes.data_as(ctypes.c_void_p),
                                     got.ctypes.data_as(ctypes.c_void_p))
    np.testing.assert_allclose(got, want, rtol=1e-12)


def test_c_pr_rda_chain_equals_numpy_restatement(oracle_lib):
    """oracle/dlco_oracle.c:dlco_oracle_pr_rda (the timed CPU baseline of row c1) follows the
    numpy restatement of src/pr-learn.cpp:303-329 bit for bit."""
    import ctypes
    rng = np.random.default_rng(5)
    n, F = 400, 96
    labels = (np.arange(n) % 2 == 0).astype(np.uint8)
    dists = (rng.random((n, F)) * (1.0 + 0.4 * (1 - labels[:, None]))).astype(np.float32)
    ip, ineg, npt, nnt = O.split_train_val(labels)
    ip, ineg = np.asarray(ip, np.int32), np.asarray(ineg, np.int32)
    w_ref, df_ref, r_ref, _ = O.pr_rda_steps(dists, ip, ineg, npt, nnt, 0.035, 0.25, 300)
    w = np.zeros(F, np.float32)
    df = np.zeros(F, np.float32)
    st = ctypes.c_ulonglong(2215)
    vp = ctypes.c_void_p
    oracle_lib.dlco_oracle_pr_rda.restype = None
    oracle_lib.dlco_oracle_pr_rda.argtypes = [vp, ctypes.c_long, ctypes.c_int, vp, ctypes.c_int, vp, ctypes.c_int,
                                              ctypes.c_float, ctypes.c_float, ctypes.c_ulonglong,
                                              ctypes.c_ulonglong, ctypes.POINTER(ctypes.c_ulonglong), vp, vp]
    oracle_lib.dlco_oracle_pr_rda(dists.ctypes.data, F, F, ip.ctypes.data, npt, ineg.ctypes.data, nnt,
                                  0.035, 0.25, 0, 300, ctypes.byref(st), w.ctypes.data, df.ctypes.data)
    assert st.value == r_ref.state
    np.testing.assert_array_equal(df, df_ref)
    np.testing.assert_array_equal(w, w_ref)
