"""Host-side drivers of the pr-learn / pj-learn loops over the C ABI.

These mirror the control flow of the reference tools' main() bodies (src/pr-learn.cpp:230-433,
src/pj-learn.cpp:214-594) -- index shuffling with cv::RNG, the 80/20 split, LogStep
validation, "Best:/Step:/Stat:" log lines in the reference's printf formats -- while every
numeric step runs in libdlco_b200.so.  The C++ tools in tools/ do the same from C++; this
module is what the tests and bench.py drive.  Nothing here imports oracle/.
"""
import time

import numpy as np

_COEFF = 4164903690
_M32 = 0xFFFFFFFF
_M64 = 0xFFFFFFFFFFFFFFFF


class CvRNG:
    """cv::RNG: multiply-with-carry, state = (u32)state * 4164903690 + (state >> 32);
    RNG() starts at 0xffffffff; uniform(a, b) = a + next() % (b - a)."""

    def __init__(self, seed=_M32):
        self.state = seed if seed else _M32

    def next(self):
        s = ((self.state & _M32) * _COEFF + (self.state >> 32)) & _M64
        self.state = s
        return s & _M32

    def uniform(self, a, b):
        return a if a == b else a + self.next() % (b - a)


def rand_shuffle(a, rng):
    """cv::randShuffle(vector<int>), iterFactor 1: for i: swap(a[i], a[next() % n])."""
    a = list(a)
    n = len(a)
    s = rng.state
    for i in range(n):
        s = ((s & _M32) * _COEFF + (s >> 32)) & _M64
        j = (s & _M32) % n
        a[i], a[j] = a[j], a[i]
    rng.state = s
    return a


def split_train_val(labels, n_div=0.80):
    """src/pr-learn.cpp:230-254 (== src/pj-learn.cpp:214-237): IdxPos / IdxNeg shuffled with the
    process-global theRNG() (default state), positives first; train = first 80 %."""
    labels = np.asarray(labels).reshape(-1)
    g = CvRNG()
    idx_pos = rand_shuffle(np.nonzero(labels == 1)[0].tolist(), g)
    idx_neg = rand_shuffle(np.nonzero(labels == 0)[0].tolist(), g)
    nd = float(np.float32(n_div))
    return (np.asarray(idx_pos, np.int32), np.asarray(idx_neg, np.int32),
            int(len(idx_pos) * nd), int(len(idx_neg) * nd))


class PrLearner:
    """pr-learn for one or many (mu, gamma) configurations on one GPU.
    dists: (N, F) float32 cuda tensor ("Distance"), labels: (N,) uint8 numpy ("Label")."""

    def __init__(self, ctx, dists, labels, mu, gamma, log_step=100000, max_dim=640, pr_params=None,
                 emit=None):
        import torch
        self.ctx, self.dists, self.torch = ctx, dists, torch
        self.labels = np.asarray(labels).reshape(-1).astype(np.uint8)
        self.mu = np.atleast_1d(np.asarray(mu, np.float32))
        self.gamma = np.atleast_1d(np.asarray(gamma, np.float32))
        self.n_cfg = len(self.mu)
        self.log_step, self.max_dim, self.pr_params = log_step, max_dim, pr_params
        self.emit = emit or (lambda s: None)
        ip, ineg, self.n_pos_trn, self.n_neg_trn = split_train_val(self.labels)
        self.idx_pos = torch.from_numpy(ip).to(ctx.device)
        self.idx_neg = torch.from_numpy(ineg).to(ctx.device)
        self.pos_val = self.idx_pos[self.n_pos_trn:].contiguous()
        self.neg_val = self.idx_neg[self.n_neg_trn:].contiguous()
        F = dists.shape[1]
        self.w = torch.zeros((self.n_cfg, F), dtype=torch.float32, device=ctx.device)
        self.dfavg = torch.zeros_like(self.w)
        self.rng_state = 2215                                  # RNG rng(2215), src/pr-learn.cpp:241
        self.t = 0
        self.obj_best = np.full(self.n_cfg, np.finfo(np.float32).max, np.float32)
        self.w_best = torch.zeros_like(self.w)
        self.d_labels = torch.from_numpy(self.labels).to(ctx.device)
        self.saved = [[] for _ in range(self.n_cfg)]           # accepted w rows per config

    def run_steps(self, n):
        """steps t .. t+n-1 (no validation in between)."""
        self.rng_state = self.ctx.pr_rda_run(self.dists, self.idx_pos, self.n_pos_trn, self.idx_neg,
                                             self.n_neg_trn, self.mu, self.gamma, self.t, n, self.rng_state,
                                             self.w, self.dfavg)
        self.t += n

    def validate(self, c):
        return self.ctx.pr_validate(self.dists, self.pos_val, self.neg_val, self.w[c], float(self.mu[c]))

    def stats(self, c):
        pd = self.ctx.pr_patchdist(self.dists, self.w[c])
        return self.ctx.roc_stats(pd, self.d_labels)

    def train(self, n_iter):
        """for t in 0..n_iter inclusive, validation whenever t is a positive multiple of LogStep
        (the reference's `step == LogStep` test fires after the update of step t)."""
        t_train = time.perf_counter()
        ls = self.log_step
        while self.t <= n_iter:
            m = (self.t + ls - 1) // ls * ls or ls               # next validation step (> 0)
            t_end = min(n_iter, m)
            self.run_steps(t_end - self.t + 1)                   # self.t becomes t_end + 1
            if t_end == m:
                self.ctx.sync()
                ttime = time.perf_counter() - t_train
                for c in range(self.n_cfg):
                    self._log(c, t_end, ttime)
                t_train = time.perf_counter()

    def _log(self, c, t, ttime):
        tv = time.perf_counter()
        loss, regul = self.validate(c)
        vtime = time.perf_counter() - tv
        obj = np.float32(loss) + np.float32(regul)
        nnz = int((self.w[c] != 0).sum().item())
        tag = "" if self.n_cfg == 1 else "[%g/%g] " % (self.mu[c], self.gamma[c])
        if obj < self.obj_best[c]:
            self.obj_best[c] = obj
            self.w_best[c].copy_(self.w[c])
            self.emit("%sBest: %i  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f)  NNZ: %i (%i)  Ttime: %.4f Vtime: %.4f"
                      % (tag, t, loss, regul, obj, self.obj_best[c], nnz, nnz, ttime, vtime))
            fpr95, auc = self.stats(c)
            w_host = self.w[c].cpu().numpy()
            n_pr, dim, nz_dim = pr_dims(self.pr_params, w_host) if self.pr_params is not None else (nnz, nnz * 8, nnz * 8)
            if dim <= self.max_dim:
                self.saved[c].append(w_host)
                self.emit("%sStat: nPR #%i (#%i) Dim/MaxDim [%i/%i] AUC: %.6f FPR95: %.2f [saved]"
                          % (tag, n_pr, nz_dim, dim, self.max_dim, auc, fpr95 * 100))
            else:
                self.emit("%sStat: nPR #%i (#%i) Dim/MaxDim [%i/%i] AUC: %.6f FPR95: %.2f"
                          % (tag, n_pr, nz_dim, dim, self.max_dim, auc, fpr95 * 100))
        else:
            nnz_best = int((self.w_best[c] != 0).sum().item())
            self.emit("%sStep: %i  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f)  NNZ: %i (%i)  Ttime: %.4f Vtime: %.4f"
                      % (tag, t, loss, regul, obj, self.obj_best[c], nnz, nnz_best, ttime, vtime))


def pr_dims(pr_params, w, n_channels=8):
    """nPR / Dim / nzDim bookkeeping of ComputePRStats (src/misc.cpp:182-217), including its
    quirk nPR = nzDim - dup_rows / 2 (dup_rows counts every row that has a twin)."""
    w = np.asarray(w).reshape(-1)
    rows = []
    for i in np.nonzero(w > 0)[0]:
        blk = pr_params[8 * i:8 * i + 8]
        rows.extend(r for r in blk if np.count_nonzero(r))
    rows = np.asarray(rows, np.float32).reshape(-1, 3)
    dup = 0
    for i in range(len(rows)):
        same = np.all(rows == rows[i], axis=1)
        same[i] = False
        dup += bool(same.any())
    nz_dim = len(rows)
    n_pr = nz_dim - dup // 2
    return n_pr, n_pr * n_channels, nz_dim


class PjLearner:
    """pj-learn (nuclear-norm RDA, src/pj-learn.cpp:286-594) for one (mu, gamma) on one GPU.
    dists: (N, D) float32 cuda tensor of unprojected distances, labels: (N,) uint8 numpy."""

    def __init__(self, ctx, dists, labels, mu, gamma, sz_batch=200, log_step=100, emit=None, split=None):
        import torch
        self.ctx, self.dists, self.torch = ctx, dists, torch
        self.labels = np.asarray(labels).reshape(-1).astype(np.uint8)
        self.mu, self.gamma = float(np.float32(mu)), float(np.float32(gamma))
        self.sz_batch, self.log_step = sz_batch, log_step
        self.emit = emit or (lambda s: None)
        ip, ineg, self.n_pos_trn, self.n_neg_trn = split or split_train_val(self.labels)
        self.h_idx_pos, self.h_idx_neg = ip, ineg
        dev = ctx.device
        self.pos_val = torch.from_numpy(ip[self.n_pos_trn:].copy()).to(dev)
        self.neg_val = torch.from_numpy(ineg[self.n_neg_trn:].copy()).to(dev)
        D = dists.shape[1]
        z = lambda: torch.zeros((D, D), dtype=torch.float32, device=dev)     # noqa: E731
        self.W, self.A, self.dLoss, self.dfavg, self.evec = z(), z(), z(), z(), z()
        self.evals = torch.zeros((D,), dtype=torch.float32, device=dev)
        self.rank = 0                       # reference: W = zeros(D, D) at t = 0 -> all margins 0
        self.rng = CvRNG(2215)              # src/pj-learn.cpp:225
        self.t = 0
        self.d_labels = torch.from_numpy(self.labels).to(dev)
        self.obj_best = np.finfo(np.float32).max
        self.auc_best, self.fpr95_best = 0.0, 1.0
        self.W_save = self.A_save = None
        self.rank_best = 0

    def sample_batch(self):
        """interleaved pos, neg draws (src/pj-learn.cpp:311-316) -> row indices into dists"""
        b = self.sz_batch
        rp, rn = np.empty(b, np.int32), np.empty(b, np.int32)
        for k in range(b):
            rp[k] = self.h_idx_pos[self.rng.uniform(0, self.n_pos_trn)]
            rn[k] = self.h_idx_neg[self.rng.uniform(0, self.n_neg_trn)]
        return rp, rn

    def step(self):
        torch, ctx = self.torch, self.ctx
        rp, rn = self.sample_batch()
        d_rp = torch.from_numpy(rp).to(ctx.device)
        d_rn = torch.from_numpy(rn).to(ctx.device)
        ctx.pj_rda_grad(self.dists, d_rp, d_rn, self.W, self.rank, out=self.dLoss)
        ctx.pj_rda_update(self.t, self.sz_batch, self.mu, self.gamma, self.dLoss, self.dfavg, self.A)
        ctx.pj_eig(self.A, self.evals, self.evec)
        self.rank = ctx.pj_prox_apply(self.evals, self.evec, self.A, self.W)
        self.t += 1

    def validate(self):
        return self.ctx.pj_validate(self.dists, self.pos_val, self.neg_val, self.W, self.rank, self.A, self.mu)

    def stats(self):
        pd = self.ctx.pj_patchdist(self.dists, self.W, self.rank)
        return self.ctx.roc_stats(pd, self.d_labels)

    def train(self, n_iter):
        t_train = time.perf_counter()
        step = 0
        while self.t <= n_iter:
            self.step()
            if step == self.log_step:
                self.ctx.sync()
                ttime = time.perf_counter() - t_train
                self._log(self.t - 1, ttime)
                step = 0
                t_train = time.perf_counter()
            step += 1

    def _log(self, t, ttime):
        tv = time.perf_counter()
        loss, regul = self.validate()
        vtime = time.perf_counter() - tv
        obj = float(np.float32(loss) + np.float32(regul))
        # the reference prints W.rows; an empty W is replaced by a D x D zero matrix (:487-490)
        rows = self.rank if self.rank > 0 else self.dists.shape[1]
        if obj < self.obj_best:
            self.obj_best = obj
            self.rank_best = rows                                              # W_Best = W.clone()
            self.emit("Best: %i  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f"
                      % (t, loss, regul, obj, self.obj_best, rows, rows, ttime, vtime))
            fpr95, auc = self.stats()
            keep = self.auc_best <= auc and self.fpr95_best >= fpr95           # :560-566
            if keep:
                self.auc_best, self.fpr95_best = auc, fpr95
                self.W_save = self.W[:max(self.rank, 0)].clone() if self.rank > 0 else self.W.clone()
                self.A_save = self.A.clone()
            self.emit("Stat: Dim [%i] AUC: %.6f (%.6f) FPR95: %.2f (%.2f)%s"
                      % (self.rank, auc, self.auc_best, fpr95 * 100, self.fpr95_best * 100,
                         " [saved]" if keep else ""))
        else:
            self.emit("Step: %i  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f"
                      % (t, loss, regul, obj, self.obj_best, rows, self.rank_best, ttime, vtime))


class PjSweep:
    """The (mu, gamma) grid of workspace/08-pjlearn.sh on one data set, in lock step on one GPU.

    Every run of the reference's grid shuffles with the default theRNG() and samples with
    RNG(2215) (src/pj-learn.cpp:214-227), so all configurations draw the same batches: the sweep
    gathers each batch once, evaluates all gradients with one grouped SYRK launch and all
    eigen-decompositions with one batched call (dlco_pj_rda_step_grid).  State is device resident;
    an iteration never synchronises with the host.  Logs are the reference's lines, prefixed
    with "[mu/gamma] " as PrLearner does for its grid."""

    def __init__(self, ctx, dists, labels, mus, gammas, sz_batch=200, log_step=100, emit=None, split=None,
                 prox="eig"):
        """prox="eig": the reference's route (eigen-decomposition, clip, V L V^T; batched cuSOLVER).
        prox="sign": the projection by the tensor-core matrix-sign iteration (dlco_pj_psd_project); W is
        rebuilt from the projected A only where it is needed (ranks are exact, validation / statistics /
        saved models go through one eigen-decomposition per log step instead of one per iteration)."""
        import torch
        assert prox in ("eig", "sign")
        self.prox, self._w_stale = prox, False
        self.ctx, self.dists, self.torch = ctx, dists, torch
        self.labels = np.asarray(labels).reshape(-1).astype(np.uint8)
        self.mu = np.atleast_1d(np.asarray(mus, np.float32))
        self.gamma = np.atleast_1d(np.asarray(gammas, np.float32))
        assert len(self.mu) == len(self.gamma)
        self.n_cfg = len(self.mu)
        self.sz_batch, self.log_step = sz_batch, log_step
        self.emit = emit or (lambda s: None)
        ip, ineg, self.n_pos_trn, self.n_neg_trn = split or split_train_val(self.labels)
        dev = ctx.device
        self.idx_pos = torch.from_numpy(np.ascontiguousarray(ip, np.int32)).to(dev)
        self.idx_neg = torch.from_numpy(np.ascontiguousarray(ineg, np.int32)).to(dev)
        self.pos_val = self.idx_pos[self.n_pos_trn:].contiguous()
        self.neg_val = self.idx_neg[self.n_neg_trn:].contiguous()
        D, n = dists.shape[1], self.n_cfg
        z = lambda: torch.zeros((n, D, D), dtype=torch.float32, device=dev)     # noqa: E731
        self.W, self.A, self.dfavg = z(), z(), z()
        self.work = torch.empty((n * (2 * D * D + D),), dtype=torch.float32, device=dev)
        self.d_rank = torch.zeros((n,), dtype=torch.int32, device=dev)
        self.d_mu = torch.from_numpy(self.mu).to(dev)
        self.d_gamma = torch.from_numpy(self.gamma).to(dev)
        self.rng_state = torch.tensor([2215], dtype=torch.int64, device=dev)    # RNG rng(2215)
        self.t = 0
        self.d_labels = torch.from_numpy(self.labels).to(dev)
        big = float(np.finfo(np.float32).max)
        self.obj_best = [big] * n
        self.auc_best, self.fpr95_best = [0.0] * n, [1.0] * n
        self.rank_best = [0] * n
        self.W_save, self.A_save = [None] * n, [None] * n

    def step(self):
        if self.prox == "sign":
            self.ctx.pj_rda_step_grid_sign(self.dists, self.idx_pos, self.n_pos_trn, self.idx_neg, self.n_neg_trn,
                                           self.sz_batch, self.rng_state, self.d_mu, self.d_gamma, self.t, self.A,
                                           self.dfavg, self.d_rank, self.work)
            self._w_stale = True
        else:
            self.ctx.pj_rda_step_grid(self.dists, self.idx_pos, self.n_pos_trn, self.idx_neg, self.n_neg_trn,
                                      self.sz_batch, self.rng_state, self.d_mu, self.d_gamma, self.t, self.W, self.A,
                                      self.dfavg, self.d_rank, self.work)
        self.t += 1

    def ranks(self):
        """Ranks of the current A (prox="sign": call refresh_w() first for the eigenvalue-based count that the
        log lines and W follow; between refreshes d_rank holds the sign iteration's trace-based estimate)."""
        if self.prox == "eig":
            self.ctx.pj_grid_check(self.n_cfg)                   # synchronises; raises on a failed eig
        return self.d_rank.cpu().numpy().astype(int)

    def refresh_w(self):
        """prox="sign": W from the current (projected) A -- one batched eigen-decomposition."""
        if self._w_stale:
            self.ctx.pj_factor_grid(self.A, self.d_rank, self.W, self.work)
            self._w_stale = False

    def validate(self, c, rank):
        self.refresh_w()
        return self.ctx.pj_validate(self.dists, self.pos_val, self.neg_val, self.W[c], int(rank), self.A[c],
                                    float(self.mu[c]))

    def stats(self, c, rank):
        self.refresh_w()
        pd = self.ctx.pj_patchdist(self.dists, self.W[c], int(rank))
        return self.ctx.roc_stats(pd, self.d_labels)

    def train(self, n_iter):
        t_train = time.perf_counter()
        step = 0
        while self.t <= n_iter:
            self.step()
            if step == self.log_step:
                self.refresh_w()                 # prox="sign": W and the eigenvalue-based ranks of the projected A
                ranks = self.ranks()
                ttime = time.perf_counter() - t_train
                for c in range(self.n_cfg):
                    self._log(c, self.t - 1, ttime, ranks[c])
                step = 0
                t_train = time.perf_counter()
            step += 1

    def _log(self, c, t, ttime, rank):
        tag = "[%g/%g] " % (self.mu[c], self.gamma[c])
        tv = time.perf_counter()
        loss, regul = self.validate(c, rank)
        vtime = time.perf_counter() - tv
        obj = float(np.float32(loss) + np.float32(regul))
        rows = rank if rank > 0 else self.dists.shape[1]
        if obj < self.obj_best[c]:
            self.obj_best[c] = obj
            self.rank_best[c] = rows
            self.emit("%sBest: %i  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f"
                      % (tag, t, loss, regul, obj, self.obj_best[c], rows, rows, ttime, vtime))
            fpr95, auc = self.stats(c, rank)
            keep = self.auc_best[c] <= auc and self.fpr95_best[c] >= fpr95
            if keep:
                self.auc_best[c], self.fpr95_best[c] = auc, fpr95
                self.W_save[c] = self.W[c, :rank].clone() if rank > 0 else self.W[c].clone()
                self.A_save[c] = self.A[c].clone()
            self.emit("%sStat: Dim [%i] AUC: %.6f (%.6f) FPR95: %.2f (%.2f)%s"
                      % (tag, rank, auc, self.auc_best[c], fpr95 * 100, self.fpr95_best[c] * 100,
                         " [saved]" if keep else ""))
        else:
            self.emit("%sStep: %i  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f"
                      % (tag, t, loss, regul, obj, self.obj_best[c], rows, self.rank_best[c], ttime, vtime))
