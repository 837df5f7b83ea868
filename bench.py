#!/usr/bin/env python
"""bench.py -- headline benchmark of the DLCO hot path on B200.

Workload (BASELINE.json configs[1]): comp-fulldists full build, 500k synthetic 64x64 patch
pairs x the full gen-poolregion candidate set (5120 rings / 40960 filters) per GPU.
One "step" = one pass of the whole 500k-pair build.  N>1: every rank builds its own 500k-pair
shard of an N x 500k job (pairs are independent: no data-path collective, weak scaling).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--pairs P] [--impl b200|reference]

Prints ONE JSON line (rank 0).  `value` = pairs/s with inputs resident in HBM; `e2e` = the same
through dlco_fulldist_build_host (pinned host patches/indices in, distances out to pinned host
memory, copies inside the timed region); `roofline` = the pooling GEMM kernel against the
measured tensor peak; `cpu_baseline` = the oracle's CPU port on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import time

# The CPU legs (cpu_baseline, --impl reference) use every host thread the box has: torch.distributed.run exports
# OMP_NUM_THREADS=1 to its workers, which would leave OpenBLAS single-threaded (round-1 SCALE record).  Set the
# thread counts before numpy / OpenBLAS load; `threads_in_use()` reports what the BLAS pool really has.
HOST_THREADS = os.cpu_count() or 1
for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
    os.environ[_v] = str(HOST_THREADS)

import numpy as np  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

RINGS, NFILT, K = 5120, 40960, 4096
FLOP_PER_PAIR = 2.0 * 2 * NFILT * K * 8          # dense-equivalent fp32 flops
GTX970_PAIRS_PER_S = 35.5                        # reference's own GPU path, derived from mtimes


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def threads_in_use():
    """Threads of the BLAS pool numpy will use (threadpoolctl), and OpenMP's view."""
    try:
        from threadpoolctl import threadpool_info
        info = threadpool_info()
        blas = [i["num_threads"] for i in info if i.get("user_api") == "blas"]
        return max(blas) if blas else HOST_THREADS
    except Exception:
        return HOST_THREADS


def ncu_traffic(kernel_substr, profiles=None):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the first kernel whose name contains
    `kernel_substr`, from the newest `ncu --set full --page raw --csv` export under profiles/ that has it.
    -> (bytes, file name) or (None, None)."""
    import csv
    import glob
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    files = sorted(glob.glob(os.path.join(profiles or os.path.join(ROOT, "profiles"), "r*_ncu_full.csv")), reverse=True)
    for path in files:
        try:
            rows = list(csv.reader(open(path, newline="")))
            head = rows[0]
            ki = head.index("Kernel Name")
            cols = {}
            for want in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                cols[want] = [i for i, h in enumerate(head) if h == want or h.endswith("." + want)]
            if not all(cols.values()):
                continue
            for row in rows[2:]:
                if len(row) > ki and kernel_substr in row[ki]:
                    tot = 0.0
                    for want, idxs in cols.items():
                        i = idxs[0]
                        tot += float(row[i].replace(",", "")) * unit.get(rows[1][i], 1.0)
                    return tot, os.path.basename(path)
        except (OSError, ValueError, IndexError):
            continue
    return None, None


# --------------------------------------------------------------------------------------------
# synthetic inputs (torch, on device): same shapes / layouts as conv-impatches + gen-poolregion
# --------------------------------------------------------------------------------------------
def synth_bank(torch, dev):
    """gen-poolregion-shaped bank (src/gen-poolregion.cpp:51-80,189-213,247-257), float64 math,
    box-truncated Gaussians normalised to unit sum; an input generator, not the oracle."""
    import math
    sig = torch.arange(1, 33, dtype=torch.float64, device=dev) * 0.5
    phi = torch.arange(0, 5, dtype=torch.float64, device=dev) * (math.pi / 16)
    r = torch.arange(0, 32, dtype=torch.float64, device=dev)
    S, Ph, R = torch.meshgrid(sig, phi, r, indexing="ij")
    S, Ph, R = S.reshape(-1), Ph.reshape(-1), R.reshape(-1)
    xc, yc = R * torch.cos(Ph), R * torch.sin(Ph)
    o0 = torch.stack([yc, yc, -yc, -yc, xc, xc, -xc, -xc], 1).reshape(-1)
    o1 = torch.stack([xc, -xc, xc, -xc, yc, -yc, -yc, yc], 1).reshape(-1)
    sg = S.repeat_interleave(8)
    x0, y0 = o1, o0                                   # get_PR_filter(ps, Offsets[:,1], Offsets[:,0], s)
    grid = torch.arange(64, dtype=torch.float64, device=dev) - 31.5
    out = torch.empty((NFILT, K), dtype=torch.float32, device=dev)
    for a in range(0, NFILT, 4096):
        b = a + 4096
        dx = grid[None, :] - x0[a:b, None]
        dy = grid[None, :] - y0[a:b, None]
        s = sg[a:b, None]
        mx = dx.abs() <= 3 * s + 1
        my = dy.abs() <= 3 * s + 1
        gx = torch.exp(-dx * dx / (2 * s * s)) * mx
        gy = torch.exp(-dy * dy / (2 * s * s)) * my
        f = gx[:, :, None] * gy[:, None, :]
        f = f / f.sum(dim=(1, 2), keepdim=True)
        out[a:b] = f.reshape(-1, K).float()
    return out


def synth_patches(torch, dev, n_points, per_point=3, seed=20240601):
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    P = n_points * per_point
    out = torch.empty((P, 64, 64), dtype=torch.uint8, device=dev)
    chunk = 8192
    for a in range(0, n_points, chunk):
        n = min(chunk, n_points - a)
        low = torch.randn((n, 1, 8, 8), generator=g, device=dev)
        base = torch.nn.functional.interpolate(low, size=(64, 64), mode="bicubic", align_corners=False)
        base = base.repeat_interleave(per_point, dim=0)
        gain = 0.8 + 0.4 * torch.rand((n * per_point, 1, 1, 1), generator=g, device=dev)
        img = 128 + 55 * gain * base + 4 * torch.randn(base.shape, generator=g, device=dev)
        out[a * per_point:(a + n) * per_point] = img.clamp(0, 255).round().to(torch.uint8)[:, 0]
    return out


def synth_pairs(torch, dev, n_points, per_point, n_pairs, seed=20240602):
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    P = n_points * per_point
    a = torch.randint(0, P, (n_pairs,), generator=g, device=dev)
    pos = torch.arange(n_pairs, device=dev) % 2 == 0
    sib = (a // per_point) * per_point + (a % per_point + torch.randint(1, per_point, (n_pairs,), generator=g, device=dev)) % per_point
    rnd = torch.randint(0, P, (n_pairs,), generator=g, device=dev)
    rnd = torch.where(rnd // per_point == a // per_point, (rnd + per_point) % P, rnd)
    b = torch.where(pos, sib, rnd)
    idx = torch.stack([a, a // per_point, b, b // per_point], 1).to(torch.int32)
    return idx.contiguous()


# --------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi) during the timed region
# --------------------------------------------------------------------------------------------
class Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.p = index, None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
            out, _ = self.p.communicate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------
# CPU arm: the oracle's port of src/comp-fulldists.cpp:320-342 (get_desc on CPU, SGEMM pooling
# through numpy/OpenBLAS with every host thread, clamp, diff^2, two reductions)
# --------------------------------------------------------------------------------------------
def _desc_worker(p):
    from oracle import dlco_oracle as O
    return O.get_desc(p)


def cpu_fulldist(filters, patches, idx, pool=None, chunk=128):
    """Per pair: get_desc x2, SGEMM (40960x4096)(4096x8) x2, min, diff^2, two reductions --
    the reference's per-pair work (no de-duplication of repeated patches), batched over
    `chunk` pairs per SGEMM call so OpenBLAS can use every core."""
    from oracle import dlco_oracle as O
    out = np.empty((len(idx), filters.shape[0] // 8), np.float32)
    for c0 in range(0, len(idx), chunk):
        sub = idx[c0:c0 + chunk]
        slots = [patches[i] for row in sub for i in (row[0], row[2])]
        if pool is not None:
            descs = pool.map(_desc_worker, slots, chunksize=8)
        else:
            descs = [O.get_desc(p) for p in slots]
        T = np.concatenate(descs, axis=1)                        # 4096 x (16 * n)
        Pm = np.minimum(filters @ T, np.float32(1.0))             # SGEMM + clamp
        Pm = Pm.reshape(filters.shape[0], len(sub), 2, 8)
        d = Pm[:, :, 1, :] - Pm[:, :, 0, :]
        l2 = (d * d).sum(axis=2)                                  # 40960 x n
        out[c0:c0 + len(sub)] = l2.reshape(-1, 8, len(sub)).sum(axis=1).T
    return out


def cpu_arm_setup():
    import ctypes
    from oracle import dlco_oracle as O
    lib_path = os.path.join(ROOT, "oracle", "libdlco_oracle.so")
    if not os.path.exists(lib_path):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    lib = ctypes.CDLL(lib_path)
    rings = O.ring_params()
    filt = np.zeros((NFILT, K), np.float32)
    lib.dlco_oracle_pr_bank(64, RINGS, rings.ctypes.data_as(ctypes.c_void_p),
                            filt.ctypes.data_as(ctypes.c_void_p), None)
    patches, pid = O.synth_patches(48, per_point=3, seed=5)
    return filt, patches, pid


def cpu_time_sample(filt, patches, pid, n_pairs, pool, seed):
    from oracle import dlco_oracle as O
    idx = O.synth_pairs(pid, n_pairs, seed=seed)
    t = time.perf_counter()
    cpu_fulldist(filt, patches, idx, pool)
    return time.perf_counter() - t


def run_cpu_baseline(target_s=12.0):
    import multiprocessing as mp
    cores = HOST_THREADS
    filt, patches, pid = cpu_arm_setup()
    pool = mp.get_context("fork").Pool(min(cores, 32))
    try:
        t0 = cpu_time_sample(filt, patches, pid, 16, pool, 1)            # calibrate (+ warm BLAS)
        n = int(max(16, min(2048, 16 * target_s / max(t0, 1e-3))))
        dt = cpu_time_sample(filt, patches, pid, n, pool, 2)
    finally:
        pool.close()
    return {"value": n / dt, "unit": "pairs/s", "cores": threads_in_use(), "kind": "port",
            "sample": "%d pairs x 5120 rings (get_desc on %d procs + OpenBLAS SGEMM 40960x4096x%d on %d threads), %.1f s"
                      % (n, min(cores, 32), 16 * n, threads_in_use(), dt)}


def run_reference_arm(args, rank):
    """--impl reference: the CPU implementation timed alone (rank 0 only), all host threads whatever the launcher
    exported (see HOST_THREADS above)."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = HOST_THREADS
    filt, patches, pid = cpu_arm_setup()
    pool = mp.get_context("fork").Pool(min(cores, 32))
    try:
        t0 = cpu_time_sample(filt, patches, pid, 16, pool, 1)
        n = int(max(16, min(1024, 16 * 6.0 / max(t0, 1e-3))))          # ~6 s per step
        for i in range(args.warmup):
            cpu_time_sample(filt, patches, pid, n, pool, 10 + i)
        tt = 0.0
        for i in range(args.steps):
            tt += cpu_time_sample(filt, patches, pid, n, pool, 100 + i)
    finally:
        pool.close()
    v = n * args.steps / tt
    thr = threads_in_use()
    log("[bench] reference arm: %d BLAS threads, %d get_desc processes, %d host cores" % (thr, min(cores, 32), cores))
    line = {
        "impl": "reference", "metric": "full-dist pairs/s", "value": v, "unit": "pairs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": tt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "comp-fulldists full build, 500k pairs x 5120 rings (40960 filters) per GPU; "
                               "CPU arm timed on a bounded sample of %d pairs per step" % n},
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": thr, "kind": "port",
                         "sample": "%d pairs/step: oracle get_desc (%d processes) + OpenBLAS SGEMM pooling (%d threads)"
                                   % (n, min(cores, 32), thr)},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def strong_scaling_leg(torch, dist, ctx, bank, dev, rank, world, n_pairs, steps, barrier):
    """ONE n_pairs job split over the ranks by shard.pair_range (weak scaling above gives every rank its own job):
    every rank builds its row range through the host entry point straight into ONE host buffer shared by all ranks
    (POSIX shared memory, each rank page-locks the slice it writes) -- the gather a single-process multi-GPU tool does
    with threads (tools/comp-fulldists.cpp).  Inputs: the same patches / pairs on every rank (seeded alike)."""
    from dlco_b200 import shard
    import shutil
    nbytes = n_pairs * RINGS * 4
    path = "/dev/shm/dlco_bench_strong_%s.f32" % os.environ.get("MASTER_PORT", "0")
    free = shutil.disk_usage("/dev/shm").free if os.path.isdir("/dev/shm") else 0
    if free < nbytes + (1 << 30):
        return {"skipped": "/dev/shm has %.1f GB free, the gathered matrix needs %.1f GB" % (free / 1e9, nbytes / 1e9)}
    per_point = 3
    n_points = max(64, int(n_pairs * 0.9) // per_point)
    patches = synth_patches(torch, dev, n_points, per_point, seed=20240601)
    idx = synth_pairs(torch, dev, n_points, per_point, n_pairs, seed=20240602)
    h_patches, h_idx = patches.cpu().pin_memory(), idx.cpu().pin_memory()
    del patches, idx
    if rank == 0:
        with open(path, "wb") as f:
            f.truncate(nbytes)
    barrier()
    whole = torch.from_file(path, shared=True, size=n_pairs * RINGS, dtype=torch.float32).view(n_pairs, RINGS)
    lo, hi = shard.pair_range(n_pairs, rank, world)
    mine = whole[lo:hi]
    cudart = torch.cuda.cudart()
    reg = cudart.cudaHostRegister(mine.data_ptr(), mine.numel() * 4, 0)
    try:
        ctx.fulldist_build_host(bank, h_patches, h_idx, lo, min(hi, lo + 2048), out=mine[:min(hi - lo, 2048)])   # warm-up
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            ctx.fulldist_build_host(bank, h_patches, h_idx, lo, hi, out=mine)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        te = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        barrier()
        res = {"pairs_total": n_pairs, "pairs_per_rank": hi - lo, "steps": steps, "pairs_per_s": n_pairs * steps / float(te.item()),
               "ms_per_step": float(te.item()) / steps * 1e3, "gather": "ranks write their row range of one /dev/shm host matrix (page-locked slice)",
               "host_register_rc": str(reg)}
        if rank == 0:        # the gathered matrix is complete and finite: sample rows of every rank's range
            probe = whole[:: max(1, n_pairs // 997)]
            res["gathered_checksum"] = float(probe.double().sum().item())
            res["gathered_finite"] = bool(torch.isfinite(probe).all().item())
    finally:
        cudart.cudaHostUnregister(mine.data_ptr())
        del mine, whole
        barrier()
        if rank == 0:
            try:
                os.remove(path)
            except OSError:
                pass
    return res


def collective_leg(torch, dist, ctx, dev, rank, world):
    """The exchange steps of a pair-sharded pj-learn iteration (opencv-dlco_b200/shard.py: every rank owns a row range
    of "Distance"; margins all-reduce of 2 x 200 floats, then the D x D gradient all-reduce) over NCCL / NVLink, checked
    against the single-GPU gradient, plus the bare all-reduce of the D x D matrix."""
    from dlco_b200 import learn, shard
    n, D, r, b = 200000, 544, 64, 200
    g = torch.Generator(device="cpu").manual_seed(1)
    lab = (torch.arange(n) % 2 == 0).to(torch.uint8)
    x = torch.randn((n, D), generator=g) * 0.1 * (1.0 + (1 - lab.float())[:, None])
    W = (torch.randn((r, D), generator=g) * 0.03).to(dev)
    ip, ineg, npt, nnt = learn.split_train_val(lab.numpy())
    rng = learn.CvRNG(2215)
    rp = np.array([ip[rng.uniform(0, npt)] for _ in range(b)], np.int32)
    rn = np.array([ineg[rng.uniform(0, nnt)] for _ in range(b)], np.int32)
    coll = shard.Collective()
    lo, hi = shard.pair_range(n, rank, world)
    sp = shard.ShardedPj(ctx, coll, x[lo:hi].to(dev), lab[lo:hi].to(dev), lo, n)
    grad = sp.grad(rp, rn, W, r)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    dist.barrier()
    e0.record()
    for _ in range(20):
        grad = sp.grad(rp, rn, W, r)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1) / 20], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    buf = torch.zeros((D, D), dtype=torch.float32, device=dev)
    for _ in range(5):
        dist.all_reduce(buf)
    torch.cuda.synchronize(dev)
    dist.barrier()
    e0.record()
    for _ in range(100):
        dist.all_reduce(buf)
    e1.record()
    torch.cuda.synchronize(dev)
    us = torch.tensor([e0.elapsed_time(e1) * 10.0], dtype=torch.float64, device=dev)
    dist.all_reduce(us, op=dist.ReduceOp.MAX)
    res = {"op": "ncclAllReduce sum f32", "bytes": D * D * 4, "allreduce_us": float(us.item()),
           "allreduce_busbw_gbs": 2.0 * (world - 1) / world * D * D * 4 / (float(us.item()) * 1e-6) / 1e9,
           "sharded_pj_grad_ms": float(ms.item()), "rows_per_rank": hi - lo, "D": D, "batch": 2 * b}
    # the same exchange through the library's own all-reduce: one kernel per GPU reading the peers' partial sums over
    # NVLink peer memory (csrc/comm.cu; CUDA IPC between the ranks' processes), summed in rank order
    try:
        import dlco_b200
        peer = dlco_b200.PeerComm(ctx, D * D)
        for _ in range(5):
            peer.allreduce(buf)
        torch.cuda.synchronize(dev)
        dist.barrier()
        e0.record()
        for _ in range(100):
            peer.allreduce(buf)
        e1.record()
        torch.cuda.synchronize(dev)
        peer.check()
        pus = torch.tensor([e0.elapsed_time(e1) * 10.0], dtype=torch.float64, device=dev)
        dist.all_reduce(pus, op=dist.ReduceOp.MAX)
        sp2 = shard.ShardedPj(ctx, shard.Collective(peer=peer), x[lo:hi].to(dev), lab[lo:hi].to(dev), lo, n)
        grad2 = sp2.grad(rp, rn, W, r)
        torch.cuda.synchronize(dev)
        dist.barrier()
        e0.record()
        for _ in range(20):
            grad2 = sp2.grad(rp, rn, W, r)
        e1.record()
        torch.cuda.synchronize(dev)
        peer.check()
        pms = torch.tensor([e0.elapsed_time(e1) / 20], dtype=torch.float64, device=dev)
        dist.all_reduce(pms, op=dist.ReduceOp.MAX)
        # bit-identical on every rank: compare rank 0's result with the others'
        ref = grad2.clone()
        dist.broadcast(ref, 0)
        same = torch.tensor([1.0 if torch.equal(ref, grad2) else 0.0], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        res["peer_memory"] = {"op": "dlco_comm_allreduce_f32 (one kernel per GPU, NVLink peer loads, rank-order sum)",
                              "allreduce_us": float(pus.item()), "sharded_pj_grad_ms": float(pms.item()),
                              "bit_identical_on_all_ranks": bool(same.item() == 1.0),
                              "rel_diff_vs_nccl_path": float((grad2 - grad).abs().max() / grad.abs().max().clamp_min(1e-30))}
        grad = grad2
        peer.close()
    except Exception as ex:
        res["peer_memory"] = {"failed": repr(ex)}
    if rank == 0:
        g1 = ctx.pj_rda_grad(x.to(dev), torch.from_numpy(rp).to(dev), torch.from_numpy(rn).to(dev), W, r)
        res["grad_rel_err_vs_single_gpu"] = float((grad - g1).abs().max() / g1.abs().max().clamp_min(1e-30))
    return res


# --------------------------------------------------------------------------------------------
def main():
    import faulthandler
    faulthandler.enable()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--pairs", type=int, default=500000)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-rows", action="store_true", help="skip the rows b/c/d block (bench_paths quick run)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the DLCO hot path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import dlco_b200
    ctx = dlco_b200.Context(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    n_pairs = args.pairs
    per_point = 3
    n_points = max(64, int(n_pairs * 0.9) // per_point)      # ~Liberty: 450k patches for 500k pairs
    t_setup = time.time()
    # the reference's own candidate set, bit-exact (gen-poolregion restated on the device: no filters.h5, no upload)
    bank = ctx.bank_analytic(dlco_b200.ring_params())
    patches = synth_patches(torch, dev, n_points, per_point, seed=20240601 + rank)
    idx = synth_pairs(torch, dev, n_points, per_point, n_pairs, seed=20240602 + rank)
    out = torch.empty((n_pairs, RINGS), dtype=torch.float32, device=dev)
    torch.cuda.synchronize(dev)
    log("[bench] rank %d setup %.1fs: %d patches, %d pairs" % (rank, time.time() - t_setup, patches.shape[0], n_pairs))

    def step():
        ctx.fulldist_build(bank, patches, idx, out=out)

    for _ in range(args.warmup):
        step()
    barrier()
    clocks = Clocks(local)
    if rank == 0:
        clocks.start()
    l0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launch_count() - l0
    clk = clocks.stop() if rank == 0 else None
    checksum = float(out[:: max(1, n_pairs // 1000)].double().sum().item())
    if not np.isfinite(checksum):
        raise SystemExit("bench.py: non-finite distances")
    # per-kernel split and the dominant kernel's average launch time: a separate pass over the same workload with
    # an event pair around every launch (the timed region above runs without them)
    prof_steps = 1 if n_pairs >= 100000 else min(args.steps, 3)
    ctx.profile(True)
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for _ in range(prof_steps):
        step()
    p1.record()
    torch.cuda.synchronize(dev)
    prof_ms = p0.elapsed_time(p1)
    pool_ms, pool_n = ctx.profile_read(1)
    desc_ms, desc_n = ctx.profile_read(0)
    red_ms, red_n = ctx.profile_read(6)
    ctx.profile(False)

    tmax = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_max = float(tmax.item())
    value = world * n_pairs * args.steps / (ms_max * 1e-3)

    # ---- end to end through the host-buffer entry point -------------------------------------
    e2e = None
    if not args.no_e2e:
        h_patches = patches.cpu().pin_memory()
        h_idx = idx.cpu().pin_memory()
        del out
        torch.cuda.empty_cache()
        h_out = torch.empty((n_pairs, RINGS), dtype=torch.float32).pin_memory()
        e2e_steps = max(1, args.steps)
        ctx.fulldist_build_host(bank, h_patches, h_idx, 0, min(n_pairs, 4096), out=h_out)   # warm-up
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            ctx.fulldist_build_host(bank, h_patches, h_idx, out=h_out)
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        te = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": world * n_pairs * e2e_steps / float(te.item()), "unit": "pairs/s",
               "h2d_bytes_per_step": n_pairs * (2 * 4096 + 16), "d2h_bytes_per_step": n_pairs * RINGS * 4,
               "steps": e2e_steps, "host_checksum": float(h_out[:: max(1, n_pairs // 1000)].double().sum().item())}

    # ---- N > 1: one n_pairs job split over the ranks (strong scaling, gathered on the host), and the collective steps
    strong, collective = None, None
    if world > 1 and not args.no_e2e:
        log("[bench] rank %d: strong-scaling leg" % rank)
        try:
            del h_out
            strong = strong_scaling_leg(torch, dist, ctx, bank, dev, rank, world, n_pairs, max(1, min(args.steps, 3)), barrier)
        except Exception as ex:
            strong = {"failed": repr(ex)}
        h_out = None
    if world > 1 and not args.no_rows:
        log("[bench] rank %d: collective leg (strong: %s)" % (rank, "ok" if strong and "pairs_per_s" in strong else strong))
        try:
            collective = collective_leg(torch, dist, ctx, dev, rank, world)
        except Exception as ex:
            collective = {"failed": repr(ex)}

    # ---- N > 1: the 5 x 9 (mu, gamma) pj-learn sweep of one data set sharded over the ranks by configuration
    # (c % world == rank, as `pj-sweep -shard r/n`), unprojected distances replicated, no collective
    sweep_sharded = None
    if world > 1 and not args.no_rows:
        log("[bench] rank %d: sharded sweep leg (collective: %s)" % (rank, collective if collective and "failed" in collective else "ok"))
        try:
            import bench_paths
            from dlco_b200 import learn
            D, n_u = 544, 100000
            ud, ulabels = bench_paths.synth_uprj(torch, dev, n_u, D, 12)
            mus = np.repeat(np.array([0.0001, 0.0005, 0.001, 0.002, 0.003], np.float32), 9)
            gms = np.tile(np.array([0.025, 0.05, 0.1, 0.15, 0.2, 0.25, 0.5, 0.75, 1.0], np.float32), 5)
            mine = [c for c in range(45) if c % world == rank]
            S = learn.PjSweep(ctx, ud, ulabels, mus[mine], gms[mine], prox="sign")
            for _ in range(5):
                S.step()
            n_it = 20
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            for _ in range(n_it):
                S.step()
            s1.record()
            barrier()
            tm = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            sweep_sharded = {"n_cfg_total": 45, "cfg_per_rank_max": -(-45 // world), "D": D, "prox": "sign",
                             "ms_per_iter": float(tm.item()) / n_it,
                             "cfg_iters_per_s": 45 * n_it / (float(tm.item()) * 1e-3)}
            del S
            # the 9 x 45 sweep of workspace/08-pjlearn.sh shards better by data set: every rank advances all 45
            # configurations of its own unprojected-distance matrix (9 data sets over `world` ranks)
            S = learn.PjSweep(ctx, ud, ulabels, mus, gms, prox="sign")
            for _ in range(3):
                S.step()
            n_it = 10
            barrier()
            s0.record()
            for _ in range(n_it):
                S.step()
            s1.record()
            barrier()
            tm = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            sweep_sharded["by_data_set"] = {"data_sets_in_flight": world, "cfg_per_rank": 45,
                                            "ms_per_iter": float(tm.item()) / n_it,
                                            "cfg_iters_per_s": world * 45 * n_it / (float(tm.item()) * 1e-3)}
            del S, ud
        except Exception as ex:
            sweep_sharded = {"failed": repr(ex)}

    log("[bench] rank %d: legs done (stdout closed: %s, isatty: %s)" % (rank, sys.stdout.closed, sys.stdout.isatty()))
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    bf16 = peaks.get("bf16_tflops_sustained")
    peak_src = "of measured: MEASURED_PEAKS.json bf16_tflops_sustained (kernel timed inside a long step)"
    if not bf16:
        bf16, peak_src = 1400.0, "of fallback: 1.4 PFLOP/s sustained bf16 (B200_PROFILING.md)"
    # what one launch of k_pool_fulldist does: every non-empty (128-row tile, 64-wide K chunk) block of
    # the bank against every 256-column pair tile, 3 fp16 MMA passes (lo*hi + hi*lo + hi*hi)
    info = bank.info()
    batch = int(os.environ.get("DLCO_FULLDIST_BATCH", "256"))
    launches_per_step = -(-n_pairs // batch)
    pairs_issued = 0
    for b0 in range(0, n_pairs, batch):
        pairs_issued += -(-min(batch, n_pairs - b0) // 16) * 16
    issued_flop_per_pair_slot = 2.0 * 3 * info["k_chunks"] * 128 * 64 * 16
    issued_per_step = issued_flop_per_pair_slot * pairs_issued
    avg_ms = pool_ms / max(pool_n, 1)
    issued_tflops = issued_per_step * prof_steps / (pool_ms * 1e-3) / 1e12 if pool_n else None
    dense_equiv_tflops = FLOP_PER_PAIR * n_pairs * prof_steps / (pool_ms * 1e-3) / 1e12 if pool_n else None
    traffic, traffic_src = ncu_traffic("k_pool_fulldist<0")
    roofline = {
        "bound": "tensor", "kernel": "k_pool_fulldist (tcgen05 kind::f16, 3 fp16 passes = fp32-equivalent)",
        "achieved": issued_tflops, "peak": bf16, "unit": "TFLOP/s",
        "frac": (issued_tflops / bf16) if issued_tflops else None,
        # dram__bytes_read.sum + dram__bytes_write.sum of one 256-pair launch, read from the ncu export under profiles/
        "traffic": traffic if batch == 256 else None, "traffic_source": traffic_src,
        "algorithmic_bytes_per_launch": batch * (8192 + RINGS * 4),
        "timed_in": "separate pass of %d step(s) with an event pair around every launch, right after the timed region "
                    "(which runs without them); %.1f ms per profiled step vs %.1f ms timed" % (prof_steps, prof_ms / prof_steps, ms / args.steps),
        "peak_source": peak_src, "launches": pool_n, "avg_launch_ms": avg_ms,
        "achieved_is": "fp16 tensor flops the kernel issues (non-empty tile x K-chunk blocks x 3 passes)",
        "issued_gflop_per_pair": issued_flop_per_pair_slot / 1e9,
        "algorithmic_gflop_per_pair_dense_fp32": FLOP_PER_PAIR / 1e9,
        "algorithmic_tflops_dense_equiv": dense_equiv_tflops,
        "bank": info, "launches_per_step": launches_per_step,
        "kernel_share_of_step": pool_ms / prof_ms if prof_ms else None,
        "desc_kernel_ms_total": desc_ms, "desc_launches": desc_n,
        "reduce_kernel_ms_total": red_ms, "reduce_launches": red_n,
        "out_stream_gbs": n_pairs * args.steps * (RINGS * 4 + 8192) / (ms * 1e-3) / 1e9,
    }
    # ---- the other rows of the path (SURVEY 8 b/c/d): pj-learn RDA iters/s, pr-learn steps/s, GEMV / FPR95
    # evaluation bandwidth -- BASELINE.json's metric names them next to full-dist pairs/s
    rows = None
    if world == 1 and not args.no_rows:
        try:
            import bench_paths
            del patches, idx
            if e2e is not None:
                h_out = h_patches = h_idx = None
            bank.close()
            torch.cuda.empty_cache()
            r = bench_paths.run(ctx, quick=True)
            rows = {
                "pj_learn_rda_iters_per_s_1cfg": r["pj_step"]["iters_per_s"],
                "pj_learn_sweep_cfg_iters_per_s_45cfg": r["pj_sweep_step"]["cfg_iters_per_s"],
                "pj_learn_rda_iters_per_s_1cfg_sign_prox": r["pj_step_sign_1cfg"]["cfg_iters_per_s"],
                "pj_learn_sweep_cfg_iters_per_s_45cfg_sign_prox": r["pj_sweep_step_sign"]["cfg_iters_per_s"],
                "pj_sign_prox_gemm_frac_of_3xtf32_peak": r["pj_sweep_step_sign"]["gemm_frac_of_3xtf32_peak"],
                "uprjdist_build_pairs_per_s": r["uprjdist_build"]["pairs_per_s"],
                "pj_syrk_grouped_frac_of_3xtf32_peak": r["pj_grad_grid"]["syrk_frac_of_3xtf32_peak"],
                "pr_learn_rda_steps_per_s_1cfg": r["pr_rda_1cfg"]["steps_per_s_per_cfg"],
                "pr_learn_rda_cfg_steps_per_s_256cfg": r["pr_rda_256cfg"]["cfg_steps_per_s"],
                "pr_stats_gemv_gbs": r["pr_patchdist_gemv"]["achieved_gbs"],
                "pr_stats_gemv_frac_hbm": r["pr_patchdist_gemv"]["frac_hbm"],
                "pr_validate_frac_hbm": r["pr_validate"]["frac_hbm"],
                "pj_stats_dist_rank80_gbs": r["pj_patchdist_rank80"]["achieved_gbs"],
                "pj_stats_dist_rank80_frac_hbm": r["pj_patchdist_rank80"]["frac_hbm"],
                "hinge_pairsum_gterms_per_s": r["hinge_pairsum_50k_x_50k"]["gterms_per_s"],
                "roc_stats_100k_ms": r["roc_stats"]["ms"],
                "source": "bench_paths.run(quick=True): 500k x 5120 full distances, 500k x 544 unprojected distances, synthetic",
            }
        except Exception as ex:   # the headline line must still be printed
            rows = {"failed": repr(ex)}
    cpu, rows_cpu = None, None
    if world == 1 and not args.no_cpu and not args.no_rows:
        try:
            import bench_paths
            rows_cpu = bench_paths.cpu_leg(quick=True)
            rows_cpu["threads"] = threads_in_use()
        except Exception as ex:
            rows_cpu = {"failed": repr(ex)}
    if world == 1 and not args.no_cpu:
        try:
            cpu = run_cpu_baseline()
        except Exception as ex:  # the GPU line must still be printed
            cpu = {"value": None, "unit": "pairs/s", "cores": os.cpu_count(), "kind": "port",
                   "sample": "failed: %r" % (ex,)}
    line = {
        "metric": "full-dist pairs/s", "value": value, "unit": "pairs/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "comp-fulldists full build: %d synthetic 64x64 patch pairs x 5120 rings "
                               "(40960 filters) per GPU, pair-sharded" % n_pairs,
                   "pairs_per_gpu": n_pairs, "filters": NFILT, "arith": "3 fp16 products per fp32 product on tcgen05 (22 mantissa bits), fp32 accumulate",
                   "bank": "gen-poolregion defaults (nr 32, nphi 5, nsigma 32), bit-exact, generated on the device",
                   "l2": "inputs larger than L2 (1.04 GB of filter planes streamed per 256-pair batch, 10 GB of output per step)",
                   "reference_gtx970_pairs_per_s": GTX970_PAIRS_PER_S},
        "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
        "cpu_baseline": cpu, "rows": rows, "rows_cpu_baseline": rows_cpu, "strong_scaling": strong,
        "collective": collective, "pj_sweep_sharded": sweep_sharded, "checksum": checksum,
    }
    log("[bench] rank 0: printing the result line (%d keys)" % len(line))
    sys.stdout.write(json.dumps(line) + "\n")
    sys.stdout.flush()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
