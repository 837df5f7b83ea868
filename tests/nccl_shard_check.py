"""2-GPU check of the pair-sharded pj-learn step over NCCL (run by hand on a multi-GPU box):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tests/nccl_shard_check.py

Each rank holds half of the rows; the all-reduced gradient, validation objective and FPR95 must
equal what rank 0 computes alone from the whole matrix.  Prints one JSON line on rank 0."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dlco_b200  # noqa: E402
from dlco_b200 import shard  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = dlco_b200.Context(local)
    dev = ctx.device
    n, D, r = 200000, 544, 64
    g = torch.Generator(device="cpu").manual_seed(1)
    lab = (torch.arange(n) % 2 == 0).to(torch.uint8)
    x = torch.randn((n, D), generator=g) * 0.1 * (1.0 + (1 - lab.float())[:, None])
    W = torch.randn((r, D), generator=g) * 0.03
    ip, ineg, npt, nnt = dlco_b200.learn.split_train_val(lab.numpy())
    rng = dlco_b200.learn.CvRNG(2215)
    b = 200
    rp = np.array([ip[rng.uniform(0, npt)] for _ in range(b)], np.int32)
    rn = np.array([ineg[rng.uniform(0, nnt)] for _ in range(b)], np.int32)
    coll = shard.Collective()
    lo, hi = shard.pair_range(n, rank, world)
    d_W = W.to(dev)
    A = (d_W.T @ d_W).contiguous()
    sp = shard.ShardedPj(ctx, coll, x[lo:hi].to(dev), lab[lo:hi].to(dev), lo, n)
    grad = sp.grad(rp, rn, d_W, r)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier()
    ev0.record()
    for _ in range(20):
        grad = sp.grad(rp, rn, d_W, r)
    ev1.record(); torch.cuda.synchronize()
    ms = torch.tensor([ev0.elapsed_time(ev1) / 20], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    loss, regul = sp.validate(ip[npt:], ineg[nnt:], d_W, r, A, 0.001)
    f, a = sp.stats(d_W, r)
    if rank == 0:
        xa, la = x.to(dev), lab.to(dev)
        g1 = ctx.pj_rda_grad(xa, torch.from_numpy(rp).to(dev), torch.from_numpy(rn).to(dev), d_W, r)
        l1, r1 = ctx.pj_validate(xa, torch.from_numpy(ip[npt:]).to(dev), torch.from_numpy(ineg[nnt:]).to(dev), d_W, r, A, 0.001)
        f1, a1 = ctx.roc_stats(ctx.pj_patchdist(xa, d_W, r), la)
        err = float((grad - g1).abs().max() / g1.abs().max().clamp_min(1e-30))
        assert float(g1.abs().max()) > 0, "hinge inactive: test data too separable"
        ok = err <= 2e-6 and abs(loss - l1) <= 1e-6 * abs(l1) and abs(f - f1) <= 1e-6 and abs(a - a1) <= 1e-9
        print(json.dumps({"world": world, "grad_rel_err": err, "loss": [loss, l1], "fpr95": [f, f1], "auc": [a, a1],
                          "sharded_grad_ms": float(ms.item()), "ok": bool(ok)}))
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
