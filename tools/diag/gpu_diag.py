"""Step-by-step GPU diagnostic (not a pytest file): prints error norms for every kernel."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import _oracle_lib, make_bank  # noqa: E402
from oracle import dlco_oracle as O  # noqa: E402
import dlco_b200  # noqa: E402


def main():
    lib = _oracle_lib()
    ctx = dlco_b200.Context(0)
    print("device", ctx.device_name())
    dev = ctx.device
    patches, pid = O.synth_patches(24, per_point=3, seed=7)
    idx = O.synth_pairs(pid, 40, seed=8)
    d_patches = torch.from_numpy(patches).to(dev)
    d_idx = torch.from_numpy(idx).to(dev)

    # ---- get_desc
    ids = torch.arange(patches.shape[0], dtype=torch.int32, device=dev)
    pt = ctx.get_desc(d_patches, ids).cpu().numpy()
    ref32 = np.stack([O.get_desc(p) for p in patches])
    ref64 = np.stack([O.get_desc(p, dt=np.float64) for p in patches])
    print("get_desc: max|gpu-o32| %.3e  max|gpu-o64| %.3e  max|o32-o64| %.3e  max val %.3f" % (
        np.abs(pt - ref32).max(), np.abs(pt - ref64).max(), np.abs(ref32 - ref64).max(), ref64.max()))

    # ---- hinge
    rng = np.random.default_rng(0)
    pos = rng.normal(0, 1, 5000).astype(np.float32)
    neg = rng.normal(1, 1, 7001).astype(np.float32)
    got = ctx.hinge_pairsum(torch.from_numpy(pos).to(dev), torch.from_numpy(neg).to(dev))
    want = O.hinge_pairsum(pos, neg)
    print("hinge: got %.9e want %.9e rel %.2e" % (got, want, abs(got - want) / want))

    # ---- pooling, small bank (256 filters = 2 M tiles)
    rings = O.ring_params()
    sel = np.r_[0:8, 700:708, 2500:2508, 5112:5120]
    filt, _ = make_bank(lib, rings[sel])
    bank = ctx.bank(filt)
    want = np.zeros((len(idx), len(sel)))
    for k, row in enumerate(idx):
        want[k] = O.full_distance(filt, pt[row[0]], pt[row[2]])
    for impl in ("simt", "tc"):
        os.environ["DLCO_POOL_IMPL"] = impl
        t = time.time()
        got = ctx.fulldist_build(bank, d_patches, d_idx)
        torch.cuda.synchronize()
        got = got.cpu().numpy()
        err = np.abs(got - want)
        rel = err.max(axis=1) / want.max(axis=1)
        print("fulldist[%s] small bank: max abs %.3e  max row-rel %.3e  (%.2fs) finite=%s" % (
            impl, err.max(), rel.max(), time.time() - t, np.isfinite(got).all()))
        if rel.max() > 1e-3:
            print(" got", got[0, :6], "\n want", want[0, :6])
    os.environ["DLCO_POOL_IMPL"] = "tc"
    # host variant
    h_out = ctx.fulldist_build_host(bank, torch.from_numpy(patches).pin_memory(), torch.from_numpy(idx).pin_memory())
    print("fulldist host vs want: max abs %.3e" % np.abs(h_out.numpy() - want).max())
    # ragged sizes
    for n in (1, 15, 17):
        got = ctx.fulldist_build(bank, d_patches, d_idx, 0, n).cpu().numpy()
        print("ragged n=%d max abs %.3e" % (n, np.abs(got - want[:n]).max()))
    bank.close()

    # ---- full bank timing
    filt, _ = make_bank(lib, rings)
    t = time.time()
    bank = ctx.bank(filt)
    print("full bank upload+split %.2fs" % (time.time() - t))
    npairs = int(os.environ.get('DIAG_PAIRS', '2048'))
    P2, pid2 = O.synth_patches(64, per_point=3, seed=11)
    idx2 = O.synth_pairs(pid2, npairs, seed=12)
    dP2 = torch.from_numpy(P2).to(dev)
    dI2 = torch.from_numpy(idx2).to(dev)
    out = ctx.fulldist_build(bank, dP2, dI2)
    torch.cuda.synchronize()
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = ctx.fulldist_build(bank, dP2, dI2)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print("full bank %d pairs: %.2f ms -> %.0f pairs/s, %.1f TFLOP/s dense-equiv" % (
            npairs, ms, npairs / ms * 1e3, npairs * 5.369e9 / ms / 1e9))
    # check a few pairs of the full bank against the oracle (desc from GPU)
    pt2 = ctx.get_desc(dP2, torch.arange(P2.shape[0], dtype=torch.int32, device=dev)).cpu().numpy()
    o = out.cpu().numpy()
    worst = 0
    for k in (0, 100, npairs - 1):
        w = O.full_distance(filt, pt2[idx2[k, 0]], pt2[idx2[k, 2]])
        worst = max(worst, np.abs(o[k] - w).max() / w.max())
    print("full bank sample row-rel err %.3e" % worst)
    print("launches", ctx.launch_count())


if __name__ == "__main__":
    main()
