"""Host-side multi-GPU logic (opencv-dlco_b200/shard.py) with world_size 2 over gloo on CPU.

The numeric backend is a numpy stand-in built from the oracle (tests may use oracle/), so what is
checked here is the partitioning and the collectives: a pair-sharded pj-learn gradient,
validation objective and FPR95 evaluation must equal the unsharded oracle result."""
import os
import socket
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class NumpyBackend:
    """The subset of dlco_b200.Context that shard.py calls, on CPU tensors, through the oracle."""

    def __init__(self):
        from oracle import dlco_oracle as O
        self.O = O

    @staticmethod
    def _rows(dists, rows):
        d = dists.numpy()
        return d if rows is None else d[rows.numpy().astype(np.int64)]

    def pj_margins(self, dists, rows, W, rank):
        x = self._rows(dists, rows)
        if rank == 0:
            return torch.zeros(x.shape[0], dtype=torch.float32)
        return torch.from_numpy(self.O.pj_margins(W.numpy()[:rank], x).astype(np.float32))

    def pj_hinge_weights(self, m):
        m = m.numpy()
        b = len(m) // 2
        viol = (m[:b, None] + np.float32(1.0)) > m[None, b:]
        c = np.concatenate([viol.sum(1), -viol.sum(0)]).astype(np.float32)
        return torch.from_numpy(c)

    def pj_wsyrk(self, dists, rows, c):
        x = self._rows(dists, rows).astype(np.float64)
        return torch.from_numpy(((x.T * c.numpy().astype(np.float64)) @ x).astype(np.float32))

    def pr_margins(self, dists, rows, w):
        return torch.from_numpy((self._rows(dists, rows).astype(np.float64) @ w.numpy().astype(np.float64)).astype(np.float32))

    def hinge_pairsum(self, pos, neg, margin=1.0):
        return self.O.hinge_pairsum(pos.numpy(), neg.numpy(), margin)

    def roc_stats(self, pd, labels):
        f, a = self.O.roc_stats(pd.numpy(), labels.numpy())
        return float(f), float(a)


def _problem():
    rng = np.random.default_rng(5)
    n, D, F = 1000, 24, 40
    labels = (np.arange(n) % 2 == 0).astype(np.uint8)
    rng.shuffle(labels)
    dj = (rng.normal(0, 0.2, (n, D)) * (1.0 + (1 - labels)[:, None])).astype(np.float32)
    dr = (rng.random((n, F)) * (1.0 + 0.4 * (1 - labels)[:, None])).astype(np.float32)
    W = (rng.normal(0, 1.5, (7, D))).astype(np.float32)
    w = np.where(rng.random(F) < 0.3, rng.random(F), 0).astype(np.float32)
    return labels, dj, dr, W, w


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    import torch.distributed as dist
    import dlco_b200
    from dlco_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        labels, dj, dr, W, w = _problem()
        n = len(labels)
        coll = shard.Collective()
        assert (coll.rank, coll.world) == (rank, world)
        lo, hi = shard.pair_range(n, rank, world, align=128)
        be = NumpyBackend()
        ip, ineg, npt, nnt = dlco_b200.learn.split_train_val(labels)
        rng = dlco_b200.learn.CvRNG(2215)
        b = 50
        rp = np.array([ip[rng.uniform(0, npt)] for _ in range(b)])
        rn = np.array([ineg[rng.uniform(0, nnt)] for _ in range(b)])
        spj = shard.ShardedPj(be, coll, torch.from_numpy(dj[lo:hi]), torch.from_numpy(labels[lo:hi]), lo, n)
        g = spj.grad(rp, rn, torch.from_numpy(W), W.shape[0]).numpy()
        A = torch.from_numpy((W.T @ W).astype(np.float32))
        vl = spj.validate(ip[npt:], ineg[nnt:], torch.from_numpy(W), W.shape[0], A, 0.001)
        st = spj.stats(torch.from_numpy(W), W.shape[0])
        spr = shard.ShardedPr(be, coll, torch.from_numpy(dr[lo:hi]), torch.from_numpy(labels[lo:hi]), lo, n)
        pv = spr.validate(ip[npt:], ineg[nnt:], torch.from_numpy(w), 0.035)
        ps = spr.stats(torch.from_numpy(w))
        np.savez(out % rank, g=g, vl=vl, st=st, pv=pv, ps=ps, lohi=[lo, hi])
    finally:
        dist.destroy_process_group()


def test_pair_range_covers_everything():
    from dlco_b200 import shard
    for n in (0, 1, 127, 128, 1000, 500000):
        for world in (1, 2, 3, 8):
            r = [shard.pair_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            for a, b in zip(r, r[1:]):
                assert a[1] == b[0]
            assert all(lo % 128 == 0 for lo, _ in r if lo < n)


def test_config_shard_is_a_partition():
    from dlco_b200 import shard
    grid = shard.sweep_grid([0.0001, 0.0005, 0.001, 0.002, 0.003],
                            [0.025, 0.05, 0.1, 0.15, 0.2, 0.25, 0.5, 0.75, 1.0])     # workspace/08-pjlearn.sh:7-13
    assert len(grid) == 45 and grid[1] == (0.0001, 0.05)
    for world in (1, 2, 8):
        parts = [shard.config_shard(len(grid), r, world) for r in range(world)]
        assert sorted(sum(parts, [])) == list(range(45))
        assert max(map(len, parts)) - min(map(len, parts)) <= 1


def test_sharded_pj_and_pr_match_unsharded(tmp_path):
    import torch.multiprocessing as mp
    from oracle import dlco_oracle as O
    import dlco_b200
    world = 2
    out = str(tmp_path / "rank%d.npz")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    res = [np.load(out % r) for r in range(world)]
    assert res[0]["lohi"][1] == res[1]["lohi"][0]

    labels, dj, dr, W, w = _problem()
    ip, ineg, npt, nnt = dlco_b200.learn.split_train_val(labels)
    rng = dlco_b200.learn.CvRNG(2215)
    b = 50
    rp = np.array([ip[rng.uniform(0, npt)] for _ in range(b)])
    rn = np.array([ineg[rng.uniform(0, nnt)] for _ in range(b)])
    pos, neg = dj[rp], dj[rn]
    want_g, _, _ = O.pj_grad(pos, neg, O.pj_margins(W, pos).astype(np.float32), O.pj_margins(W, neg).astype(np.float32))
    want_vl = O.pj_validate(dj[ip[npt:]], dj[ineg[nnt:]], W, W.T @ W, 0.001)
    want_st = O.roc_stats(O.pj_patch_dist(W, dj)[1].astype(np.float32), labels)
    want_pv = O.pr_validate(dr[ip[npt:]], dr[ineg[nnt:]], w, 0.035)
    want_ps = O.roc_stats((dr.astype(np.float64) @ w.astype(np.float64)).astype(np.float32), labels)
    scale = np.abs(want_g).max()
    for r in res:
        assert np.abs(r["g"] - want_g).max() <= 2e-6 * scale
        np.testing.assert_allclose(r["vl"], want_vl, rtol=2e-6)
        np.testing.assert_allclose(r["st"], [float(want_st[0]), want_st[1]], rtol=1e-6)
        np.testing.assert_allclose(r["pv"], want_pv, rtol=2e-6)
        np.testing.assert_allclose(r["ps"], [float(want_ps[0]), want_ps[1]], rtol=1e-6)
    # every rank ends with the same gradient bit for bit (all-reduce result)
    assert np.array_equal(res[0]["g"], res[1]["g"])
