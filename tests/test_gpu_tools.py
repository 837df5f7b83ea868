"""The drop-in command-line tools (tools/*.cpp -> bin/) end to end on a GPU box, through real HDF5
files: argv, stdout line formats and output datasets as the reference's workspace scripts expect
them, numbers against the oracle / the library's Python drivers."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from conftest import make_bank  # noqa: E402
from oracle import dlco_oracle as O  # noqa: E402
from oracle.h5lite import H5File  # noqa: E402

BIN = os.path.join(ROOT, "bin")
BAR = "0...10...20...30...40...50...60...70...80...90...100 - done."


def pack(path, tmp, sets):
    """sets: list of (name, array, chunk_rows, deflate) -> HDF5 file via bin/h5tool pack"""
    cmd = [os.path.join(BIN, "h5tool"), "pack", path]
    tn = {np.dtype(np.uint8): "u8", np.dtype(np.int32): "i32", np.dtype(np.float32): "f32", np.dtype(np.float64): "f64"}
    for name, arr, cr, lvl in sets:
        arr = np.ascontiguousarray(arr)
        b = os.path.join(tmp, name + ".bin")
        arr.tofile(b)
        cmd += [name, tn[arr.dtype], str(arr.ndim)] + [str(d) for d in arr.shape] + [b, str(cr), str(lvl)]
    subprocess.check_call(cmd)


def run(tool, *args):
    r = subprocess.run([os.path.join(BIN, tool)] + [str(a) for a in args], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    return r.returncode, r.stdout.decode(), r.stderr.decode()


@pytest.fixture(scope="module")
def workspace(tmp_path_factory, gpu_ctx, oracle_lib):
    """filters.h5 (24 rings = 192 filters), train / test image sets; built once for the module."""
    import __graft_entry__
    __graft_entry__.build()
    tmp = str(tmp_path_factory.mktemp("ws"))
    rings_all = O.ring_params()
    pick = np.r_[np.arange(0, 5120, 320), [3, 40, 900, 1500, 2600, 3300, 4100, 5000]]
    rings = rings_all[pick]
    filt, prp = make_bank(oracle_lib, rings)
    pack(os.path.join(tmp, "filters.h5"), tmp,
         [("RingParams", rings.astype(np.float64), 128, 9), ("PRParams", prp, 128, 9),
          ("PRFilters", filt.reshape(-1, 64, 64), 128, 1)])
    ws = {"tmp": tmp, "filters": os.path.join(tmp, "filters.h5"), "filt": filt, "prp": prp}
    for name, seed, n_points, n_pairs in [("train", 11, 150, 700), ("test", 12, 60, 300)]:
        patches, pid = O.synth_patches(n_points, per_point=3, seed=seed)
        idx = O.synth_pairs(pid, n_pairs, seed=seed + 100)
        pack(os.path.join(tmp, name + "-img.h5"), tmp, [("Indices", idx, 128, 0), ("Patches", patches, 128, 9)])
        ws[name] = (patches, idx)
    return ws


def test_usage_and_exit_codes():
    for tool in ["comp-fulldists", "comp-uprjdists", "pr-learn", "pj-learn", "pj-sweep", "pr-stats", "pj-stats"]:
        rc, out, _ = run(tool, "-help")
        assert rc == 1 and "Usage:" in out
    rc, out, _ = run("pr-learn", "-bogus", "a", "b", "c")
    assert rc == 1 and "ERROR: Invalid -bogus option." in out


def test_pipeline(workspace, gpu_ctx):
    ws, tmp = workspace, workspace["tmp"]
    flt = ws["filters"]
    dev = gpu_ctx.device
    # ---- comp-fulldists (workspace/03-compdist.sh:13) ------------------------------------------
    dist = {}
    for name in ("train", "test"):
        out = os.path.join(tmp, name + "-dist.h5")
        rc, so, se = run("comp-fulldists", flt, os.path.join(tmp, name + "-img.h5"), out)
        assert rc == 0, se
        patches, idx = ws[name]
        assert so.startswith("Bins: #8 Sigma: 1.4 bNorm: 1\nLoad Indices.\nLoad Patches.\n" + BAR + "\nLoad PRFilters.\n" + BAR)
        assert "Export Pair Labels: #%d\n" % len(idx) in so and "Start Compute L2 distances.\n" in so
        assert "Compute Capability: 10.0" in so
        assert re.search(r"\rStep: %d / %d\nDone\.\n\nTotal: \d+\.\d{9} sec\n\n$" % (len(idx), len(idx)), so)
        f = H5File(out)
        lab, d = f.read("Label"), f.read("Distance")
        assert lab.shape == (len(idx), 1) and lab.dtype == np.uint8 and d.shape == (len(idx), 24) and d.dtype == np.float32
        assert np.array_equal(lab.ravel(), O.pair_labels(idx))
        want = O.full_distances(ws["filt"], patches[:], idx[:64])
        # end to end against the oracle's own (cv2-pinned) fp32 get_desc at north_star's 1e-5
        rel = np.abs(d[:64] - want).max(axis=1) / want.max(axis=1)
        assert rel.max() <= 1e-5, rel.max()
        lib = gpu_ctx.fulldist_build(gpu_ctx.bank(ws["filt"]), torch.from_numpy(patches).to(dev), torch.from_numpy(idx).to(dev))
        assert np.array_equal(lib.cpu().numpy(), d)
        dist[name] = (d, lab.ravel())

    # ---- pr-learn (workspace/04-prlearn.sh:23-27) -----------------------------------------------
    prm = os.path.join(tmp, "train-0.035-0.250-pr.h5")
    rc, so, se = run("pr-learn", "-iters", 300000, "-mu", 0.035, "-gamma", 0.25, flt, os.path.join(tmp, "train-dist.h5"), prm)
    assert rc == 0, se
    assert so.startswith("mu: 0.035 gamma: 0.25 maxdim: 640 nIters: 300000\nLoad PRParams.\nLoad RingParams.\n"
                         "Load Labels: 700\nLoad Distances: 700 x 24\n" + BAR + "\nPositive samples #350\nNegative samples #350\n"
                         "Positive train #280\nNegative train #280\nPositive valid #70\nNegative valid #70\n\nFound GPU: ")
    pat = r"^(Best|Step): (\d+)  Loss: (\d+\.\d{6}) Regul: (\d+\.\d{6}) Obj: (\d+\.\d{6}) \((\d+\.\d{6})\)  NNZ: (\d+) \((\d+)\)  Ttime: \d+\.\d{4} Vtime: \d+\.\d{4}$"
    steps = [re.match(pat, ln) for ln in so.splitlines() if ln[:5] in ("Best:", "Step:")]
    assert all(steps) and [int(m.group(2)) for m in steps] == [100000, 200000, 300000]
    # the same run through the Python driver of the library: identical numbers
    import dlco_b200
    lines = []
    d_tr, lab_tr = dist["train"]
    L = dlco_b200.learn.PrLearner(gpu_ctx, torch.from_numpy(d_tr).to(dev), lab_tr, 0.035, 0.25, pr_params=ws["prp"],
                                  emit=lines.append)
    L.train(300000)
    strip = lambda s: re.sub(r"  Ttime: .*$", "", s)                                  # noqa: E731
    assert [strip(x) for x in so.splitlines() if x[:5] in ("Best:", "Step:", "Stat:")] == [strip(x) for x in lines]
    saved = [ln for ln in so.splitlines() if ln.endswith("[saved]")]
    w = H5File(prm).read("w")
    assert w.shape == (len(saved), 24) and H5File(prm).info("w")["maxshape"][0] == 0xFFFFFFFFFFFFFFFF
    assert np.array_equal(w, np.stack(L.saved[0]))
    assert len(saved) >= 1 and (w[-1] > 0).sum() >= 2

    # ---- pr-stats (workspace/05-prstats.sh:26-30) ------------------------------------------------
    rc, so, se = run("pr-stats", flt, "-dst", os.path.join(tmp, "test-dist.h5"), "-prj", prm)
    assert rc == 0, se
    assert so.startswith("MaxDim: 640 nChannels: 8\nLoad PRParams.\nLoad RingParams.\nSet: [%s] {full}\n    Load TestLabel: %s\n"
                         "    Validate Model: " % (os.path.join(tmp, "test-dist.h5"), BAR))
    ms = re.findall(r"^  ModelStat: AUC #(\S+)  FPR95: (\d+\.\d\d) Dim/MaxDim \[(\d+)/640\] nPR: (\d+) \(#(\d+)\) \[(.*)\]\(#(\d+)\)->\[(.*)\]$",
                    so, re.M)
    assert len(ms) == len(w)
    aucs = [float(m[0]) for m in ms]
    assert aucs == sorted(aucs)                                                       # multimap order
    d_te, lab_te = dist["test"]
    for m in ms:
        k = int(m[6])
        n_pr, dim, nz, fpr95, auc = O.pr_stats(ws["prp"], d_te, lab_te, w[k])
        assert (int(m[3]), int(m[2]), int(m[4])) == (n_pr, dim, nz)
        assert abs(float(m[1]) - float(fpr95) * 100) <= 0.1 + 0.005 and abs(float(m[0]) - auc) <= 1e-5
    best = re.search(r"\n  BestModel: \[(.*)\] #(\d+)\n$", so)
    assert best and best.group(1) == prm and int(best.group(2)) == int(ms[-1][6])

    # ---- comp-uprjdists (workspace/07-compunproj.sh:29) ------------------------------------------
    wid = len(w) - 1
    unproj = {}
    for name in ("train", "test"):
        out = os.path.join(tmp, name + "-unproj.h5")
        rc, so, se = run("comp-uprjdists", flt, os.path.join(tmp, name + "-img.h5"), "-prj", prm, "-id", wid, "-out", out)
        assert rc == 0, se
        sel = O.select_pr_filters(ws["filt"], w[wid])
        assert "Load Learnt Filters: [%s]#%d\nPRFilters: %d x 4096\nDescriptor size: %d\n" % (prm, wid, len(sel), 8 * len(sel)) in so
        assert "Start Compute L1 distances.\n" in so and so.endswith(" sec\n\n")
        patches, idx = ws[name]
        d = H5File(out).read("Distance")
        assert d.shape == (len(idx), 8 * len(sel)) and np.array_equal(H5File(out).read("Label").ravel(), O.pair_labels(idx))
        want = O.uprj_distances(sel, patches, idx[:48])
        # differences of values clamped to [0, 1]; end to end against the oracle's own (cv2-pinned) fp32
        # get_desc at north_star's 1e-5, relative to the row's largest entry
        assert (np.abs(d[:48] - want).max(axis=1) / np.abs(want).max(axis=1)).max() <= 1e-5
        unproj[name] = (d, O.pair_labels(idx))

    # ---- pj-learn (workspace/08-pjlearn.sh:24-27) ------------------------------------------------
    pjm = os.path.join(tmp, "train-pj.h5")
    rc, so, se = run("pj-learn", "-iters", 300, "-mu", 0.001, "-gamma", 0.1, os.path.join(tmp, "train-unproj.h5"), pjm)
    assert rc == 0, se
    D = unproj["train"][0].shape[1]
    assert so.startswith("mu: 0.001 gamma: 0.1 nIters: 300\nLoad Labels: 700\nLoad Distances: 700 x %d\n" % D)
    pat = r"^(Best|Step): (\d+)  Loss: \d+\.\d{6} Regul: \d+\.\d{6} Obj: \d+\.\d{6} \(\d+\.\d{6}\) Rank: (\d+) \((\d+)\) Ttime: \d+\.\d{4} Vtime: \d+\.\d{4}$"
    steps = [re.match(pat, ln) for ln in so.splitlines() if ln[:5] in ("Best:", "Step:")]
    assert all(steps) and [int(m.group(2)) for m in steps] == [100, 200, 300]
    assert re.search(r"^Stat: Dim \[\d+\] AUC: \d\.\d{6} \(\d\.\d{6}\) FPR95: \d+\.\d\d \(\d+\.\d\d\) \[saved\]$", so, re.M)
    lines = []
    x_tr, lab_tr = unproj["train"]
    P = dlco_b200.learn.PjLearner(gpu_ctx, torch.from_numpy(x_tr).to(dev), lab_tr, 0.001, 0.1, emit=lines.append)
    P.train(300)
    strip = lambda s: re.sub(r" Ttime: .*$", "", s)                                   # noqa: E731
    la = [strip(x).split() for x in so.splitlines() if x[:5] in ("Best:", "Step:", "Stat:")]
    lb = [strip(x).split() for x in lines]
    assert len(la) == len(lb)
    for ta, tb in zip(la, lb):                  # same lines; printed floats may differ in the last digit (library eig
        assert len(ta) == len(tb)               # is not bitwise reproducible from run to run)
        for xa, xb in zip(ta, tb):
            fa, fb = xa.strip("()"), xb.strip("()")
            if re.fullmatch(r"-?\d+\.\d+", fa):
                assert abs(float(fa) - float(fb)) <= 2e-6 + 0.0101 * (len(fa.split(".")[1]) == 2), (ta, tb)
            else:
                assert xa == xb, (ta, tb)
    f = H5File(pjm)
    W, A = f.read("W"), f.read("A")
    assert A.shape == (D, D) and W.shape[1] == D and f.info("W")["layout"][0] == "contiguous"
    # eigenvectors are defined up to sign (and cuSOLVER is not run-to-run bitwise reproducible):
    # compare A, and W through W^T W
    A_py, W_py = P.A_save.cpu().numpy(), P.W_save.cpu().numpy()
    assert W.shape == W_py.shape
    assert np.abs(A - A_py).max() <= 1e-4 * np.abs(A_py).max()
    assert np.abs(W.T @ W - W_py.T @ W_py).max() <= 1e-4 * np.abs(A_py).max()
    assert np.abs(W.T.astype(np.float64) @ W - A).max() <= 1e-5 * max(np.abs(A).max(), 1e-30)

    # ---- pj-sweep: the 08-pjlearn.sh loop in lock step; files and logs as the loop would leave them ----
    sw = os.path.join(tmp, "sweep")
    rc, so2, se = run("pj-sweep", os.path.join(tmp, "train-unproj.h5"), sw, "-mu", "0.001,0.0020", "-gamma", "0.1",
                      "-iters", 300)
    assert rc == 0, se
    assert "2 grid points" in so2 and "config-iterations/s" in so2
    log = open(os.path.join(sw, "logging", "train-0.001-0.1-pj.log")).read()
    assert log.startswith("mu: 0.001 gamma: 0.1 nIters: 300\nLoad Labels: 700\nLoad Distances: 700 x %d\n" % D)
    assert os.path.exists(os.path.join(sw, "logging", "train-0.0020-0.1-pj.log"))

    def toks(text):
        out = []
        for ln in text.splitlines():
            if ln[:5] in ("Best:", "Step:", "Stat:"):
                out.append(re.sub(r" Ttime: .*$", "", ln).replace("(", " ").replace(")", " ").replace("[", " ").replace("]", " ").split())
        return out
    a, b = toks(so), toks(log)
    assert len(a) == len(b) and len(a) >= 3
    for la, lb in zip(a, b):
        assert len(la) == len(lb)
        for x, y in zip(la, lb):
            if re.fullmatch(r"-?\d+\.\d+", x):
                assert abs(float(x) - float(y)) <= 3e-6 + 0.0101 * ("." in x and len(x.split(".")[1]) == 2), (la, lb)
            else:
                assert x == y, (la, lb)
    fs = H5File(os.path.join(sw, "train-0.001-0.1-pj.h5"))
    A_s, W_s = fs.read("A"), fs.read("W")
    assert W_s.shape == W.shape and np.abs(A_s - A).max() <= 1e-4 * np.abs(A).max()
    assert np.abs(W_s.T.astype(np.float64) @ W_s - A_s).max() <= 1e-5 * max(np.abs(A_s).max(), 1e-30)
    f2 = H5File(os.path.join(sw, "train-0.0020-0.1-pj.h5"))
    assert f2.read("A").shape == (D, D)
    # -prox sign: the projection by the tensor-core sign iteration; same files, model within the bar
    rc, so4, se = run("pj-sweep", os.path.join(tmp, "train-unproj.h5"), os.path.join(tmp, "sweep_sign"), "-mu", "0.001,0.0020",
                      "-gamma", "0.1", "-iters", 300, "-prox", "sign")
    assert rc == 0, se
    fg = H5File(os.path.join(tmp, "sweep_sign", "train-0.001-0.1-pj.h5"))
    A_g, W_g = fg.read("A"), fg.read("W")
    assert W_g.shape == W_s.shape and np.abs(A_g - A_s).max() <= 1e-4 * np.abs(A_s).max()
    assert np.abs(W_g.T.astype(np.float64) @ W_g - A_g).max() <= 3e-5 * max(np.abs(A_g).max(), 1e-30)
    c = toks(open(os.path.join(tmp, "sweep_sign", "logging", "train-0.001-0.1-pj.log")).read())
    assert len(c) == len(b) and [ln[0] for ln in c] == [ln[0] for ln in b]
    # pj-learn itself with the sign prox: through the flag and through the environment (unchanged scripts)
    pjs = os.path.join(tmp, "train-pj-sign.h5")
    rc, so5, se = run("pj-learn", "-iters", 300, "-mu", 0.001, "-gamma", 0.1, "-prox", "sign",
                      os.path.join(tmp, "train-unproj.h5"), pjs)
    assert rc == 0, se
    fl = H5File(pjs)
    A_l, W_l = fl.read("A"), fl.read("W")
    assert W_l.shape == W_g.shape and np.array_equal(A_l, A_g)                       # same kernels as pj-sweep
    assert np.abs(W_l.T.astype(np.float64) @ W_l - A_l).max() <= 3e-5 * np.abs(A_l).max()   # W: batch-1 library eig
    assert [ln[0] for ln in toks(so5)] == [ln[0] for ln in b]
    env = dict(os.environ, DLCO_PJ_PROX="sign")
    pje = os.path.join(tmp, "train-pj-env.h5")
    r = subprocess.run([os.path.join(BIN, "pj-learn"), "-iters", "300", "-mu", "0.001", "-gamma", "0.1",
                        os.path.join(tmp, "train-unproj.h5"), pje], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr
    assert np.array_equal(H5File(pje).read("A"), A_l)
    # -shard r/n partitions the grid
    rc, so3, se = run("pj-sweep", os.path.join(tmp, "train-unproj.h5"), os.path.join(tmp, "sweep1"), "-mu", "0.001,0.0020",
                      "-gamma", "0.1", "-iters", 20, "-shard", "1/2")
    assert rc == 0 and "1 grid points" in so3 and os.listdir(os.path.join(tmp, "sweep1", "logging")) == ["train-0.0020-0.1-pj.log"]

    # ---- pj-stats (defined by us, see tools/pj-stats.cpp) ----------------------------------------
    rc, so, se = run("pj-stats", flt, "-maxdim", D, "-dst", os.path.join(tmp, "test-unproj.h5"), "-prg", prm, "-prj", pjm)
    assert rc == 0, se
    m = re.search(r"^  ModelStat: AUC #(\S+)  FPR95: (\d+\.\d\d) Dim/MaxDim \[(\d+)/%d\] " % D, so, re.M)
    assert m, so
    dim, fpr95, auc = O.pj_stats(W, unproj["test"][0], unproj["test"][1])
    assert int(m.group(3)) == dim and abs(float(m.group(2)) - float(fpr95) * 100) <= 0.1 + 0.005
    assert abs(float(m.group(1)) - auc) <= 1e-5
    assert "\n  BestModel: [%s] #0\n" % pjm in so
    # MaxDim filter
    rc, so, _ = run("pj-stats", "-maxdim", 1, "-dst", os.path.join(tmp, "test-unproj.h5"), "-prj", pjm)
    assert rc == 0 and "ModelStat" not in so and "BestModel: [] #-1" in so


def test_fulldists_rejects_nan_free_but_bad_inputs(workspace):
    """missing datasets fail loudly with a non-zero exit code (no partial success)"""
    tmp = workspace["tmp"]
    rc, _, se = run("comp-fulldists", workspace["filters"], workspace["filters"], os.path.join(tmp, "x.h5"))
    assert rc == 2 and "Indices" in se
