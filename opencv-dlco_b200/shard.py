"""Multi-GPU partitioning of the DLCO hot path: one process per GPU,
torch.distributed for the plumbing (NCCL on the GPUs, gloo in the CPU tests).

The reference is single-GPU (cuda::setDevice(0), src/pj-learn.cpp:267); everything here is new
design.  What shards how:

  * comp-fulldists / comp-uprjdists rows, and the mu x gamma sweeps: independent units, no
    data-path collective -- `pair_range`, `config_shard`.
  * pj-learn gradient with a pair-sharded "Distance" matrix (`ShardedPj.grad`): every rank draws the
    same cv::RNG batch, computes the margins of the sampled rows it owns, the margins are exchanged
    (one all-reduce of 2*szBatch floats; each slot has exactly one owner so the sum is exact), the
    hinge weights are derived redundantly, and the partial sum_k c_k x_k x_k^T of the owned rows is
    all-reduced (D x D floats, src/pj-learn.cpp:367-418 is the single-GPU loop it replaces).
  * validation objective (`ShardedPr/ShardedPj.validate`): margins of the owned hold-out rows,
    all-gather of the negative margins, local positives x all negatives hinge sum, all-reduce of
    one f64 scalar (src/pr-learn.cpp:340-358, src/pj-learn.cpp:501-527).
  * FPR95 evaluation (`stats`): all-gather of the per-pair distances and labels, then the ROC of
    src/misc.cpp:227-263 on every rank.

`backend` is the object that runs the local numeric steps: a dlco_b200.Context on a GPU box; the
CPU tests pass a numpy stand-in so that the partitioning and the collectives are exercised with
world_size 2 over gloo.  Nothing here imports oracle/.
"""
import numpy as np


def pair_range(n, rank, world, align=128):
    """Contiguous row range [lo, hi) of rank `rank`; boundaries are multiples of `align` (the
    reference's sChunk = 128 HDF5 row chunk, src/comp-fulldists.cpp:286) except the last."""
    if world <= 1:
        return 0, n
    blocks = (n + align - 1) // align
    per, extra = divmod(blocks, world)
    lo_b = rank * per + min(rank, extra)
    hi_b = lo_b + per + (1 if rank < extra else 0)
    return min(n, lo_b * align), min(n, hi_b * align)


def config_shard(n_cfg, rank, world):
    """Round-robin assignment of the mu x gamma grid (workspace/04-prlearn.sh, 08-pjlearn.sh)."""
    return list(range(rank, n_cfg, max(world, 1)))


def sweep_grid(mus, gammas):
    """The scripts' nested loops: mu outer, gamma inner -> list of (mu, gamma)."""
    return [(float(m), float(g)) for m in mus for g in gammas]


class Collective:
    """Thin wrapper over torch.distributed; a world of 1 is a no-op (no process group needed)."""

    def __init__(self, group=None, peer=None):
        """peer: a dlco_b200.PeerComm -- float32 CUDA sums then go through the library's own all-reduce (one kernel per
        GPU over NVLink peer memory, bit-identical on every rank) instead of the process group's collective."""
        import torch
        self.torch = torch
        self.dist = None
        self.group = group
        self.peer = peer
        self.rank, self.world = 0, 1
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                self.dist = dist
                self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        except ImportError:
            pass

    def allreduce_sum(self, t):
        if self.world > 1:
            if self.peer is not None and t.is_cuda and t.dtype == self.torch.float32 and t.is_contiguous():
                return self.peer.allreduce(t)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t

    def allreduce_max(self, t):
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        return t

    def allgather_cat(self, t):
        """Concatenate 1-D (or row-stacked) tensors of possibly different lengths, rank order."""
        if self.world == 1:
            return t
        torch = self.torch
        n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
        sizes = [torch.zeros_like(n) for _ in range(self.world)]
        self.dist.all_gather(sizes, n, group=self.group)
        sizes = [int(s.item()) for s in sizes]
        m = max(sizes)
        pad = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[:t.shape[0]] = t
        parts = [torch.empty_like(pad) for _ in range(self.world)]
        self.dist.all_gather(parts, pad, group=self.group)
        return torch.cat([p[:s] for p, s in zip(parts, sizes)], 0)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier(group=self.group)


class _Sharded:
    """Rows [lo, hi) of the N-row "Distance" matrix and of "Label" live on this rank."""

    def __init__(self, backend, coll, local_dists, local_labels, lo, n_total):
        self.be, self.coll, self.torch = backend, coll, coll.torch
        self.dists, self.lo, self.hi, self.n = local_dists, lo, lo + local_dists.shape[0], n_total
        self.labels = local_labels
        self.dev = local_dists.device

    def _own(self, rows):
        """global row ids (numpy) -> (positions in `rows` owned here, local row ids tensor)"""
        rows = np.asarray(rows, np.int64)
        pos = np.nonzero((rows >= self.lo) & (rows < self.hi))[0]
        loc = self.torch.from_numpy((rows[pos] - self.lo).astype(np.int32)).to(self.dev)
        return pos, loc

    def _scatter_sum(self, n, pos, vals):
        """n-vector with `vals` at `pos`, zero elsewhere, summed over ranks (each slot has one owner)"""
        full = self.torch.zeros((n,), dtype=self.torch.float32, device=self.dev)
        if len(pos):
            full[self.torch.from_numpy(pos).to(self.dev)] = vals
        return self.coll.allreduce_sum(full)

    def _hinge_mean(self, m_pos_local, m_neg_local, n_pos, n_neg):
        torch = self.torch
        neg_all = self.coll.allgather_cat(m_neg_local)
        s = self.be.hinge_pairsum(m_pos_local, neg_all) if m_pos_local.numel() and neg_all.numel() else 0.0
        tot = self.coll.allreduce_sum(torch.tensor([s], dtype=torch.float64, device=self.dev))
        # Loss / nPosVal / nNegVal in float, as the reference (src/pj-learn.cpp:524)
        return float(np.float32(np.float32(np.float32(tot.item()) / np.float32(n_pos)) / np.float32(n_neg)))

    def _roc(self, pd_local):
        pd = self.coll.allgather_cat(pd_local)
        lab = self.coll.allgather_cat(self.labels)
        return self.be.roc_stats(pd, lab)


class ShardedPj(_Sharded):
    """pj-learn steps over a pair-sharded unprojected-distance matrix."""

    def grad(self, rows_pos, rows_neg, W, rank):
        """rows_pos / rows_neg: the szBatch sampled GLOBAL row ids (numpy, identical on every
        rank: same RNG(2215) stream).  Returns dLoss (D x D), identical on every rank."""
        b = len(rows_pos)
        rows = np.concatenate([np.asarray(rows_pos, np.int64), np.asarray(rows_neg, np.int64)])
        pos, loc = self._own(rows)
        m_loc = self.be.pj_margins(self.dists, loc, W, rank) if len(pos) else None
        m = self._scatter_sum(2 * b, pos, m_loc)
        c = self.be.pj_hinge_weights(m)
        c_loc = c[self.torch.from_numpy(pos).to(self.dev)] if len(pos) else c[:0]
        part = self.be.pj_wsyrk(self.dists, loc, c_loc)
        return self.coll.allreduce_sum(part)

    def validate(self, pos_val, neg_val, W, rank, A, mu):
        """pos_val / neg_val: GLOBAL hold-out row ids -> (LossVal, Regul = mu * trace(A))"""
        _, lp = self._own(pos_val)
        _, ln = self._own(neg_val)
        mp = self.be.pj_margins(self.dists, lp, W, rank)
        mn = self.be.pj_margins(self.dists, ln, W, rank)
        loss = self._hinge_mean(mp, mn, len(pos_val), len(neg_val))
        regul = float(np.float32(float(mu) * float(self.torch.trace(A.double()).item())))
        return loss, regul

    def stats(self, W, rank):
        return self._roc(self.be.pj_margins(self.dists, None, W, rank))


class ShardedPr(_Sharded):
    """pr-learn evaluation passes over a pair-sharded full-distance matrix.  The RDA chain itself
    is serial (batch 1, src/pr-learn.cpp:303-329): "replicas only" -- configurations are sharded
    with `config_shard`, not the chain."""

    def validate(self, pos_val, neg_val, w, mu):
        _, lp = self._own(pos_val)
        _, ln = self._own(neg_val)
        mp = self.be.pr_margins(self.dists, lp, w)
        mn = self.be.pr_margins(self.dists, ln, w)
        loss = self._hinge_mean(mp, mn, len(pos_val), len(neg_val))
        regul = float(np.float32(float(mu) * float(w.double().abs().sum().item())))
        return loss, regul

    def stats(self, w):
        return self._roc(self.be.pr_margins(self.dists, None, w))
