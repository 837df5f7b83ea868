"""Two or more GPUs in one box: the peer-memory all-reduce (csrc/comm.cu) and the pair-sharded pj-learn built on it
(`pj-learn -gpus N`).  Skipped on a single-GPU box (run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py`;
results of the last run are kept in profiles/)."""
import ctypes
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
from oracle.h5lite import H5File  # noqa: E402
from test_workspace_scripts import h5pack  # noqa: E402

needs2 = pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")


@needs2
def test_peer_allreduce_in_one_process():
    import __graft_entry__
    __graft_entry__.build()
    import dlco_b200
    n = min(torch.cuda.device_count(), 4)
    lib = dlco_b200.load_library()
    ctxs = [dlco_b200.Context(g) for g in range(n)]
    arr = (ctypes.c_void_p * n)(*[c.handle for c in ctxs])
    comms = (ctypes.c_void_p * n)()
    count_max = 544 * 544
    rc = lib.dlco_comm_create_local(arr, n, count_max, comms)
    assert rc == 0, lib.dlco_last_error()
    g = torch.Generator().manual_seed(3)
    for count in (400, 544 * 544, 1001):                       # margins, gradient, an odd length (scalar path)
        for rep in range(7):                                   # consecutive calls alternate the staging parity
            xs = [torch.randn(count, generator=g) for _ in range(n)]
            d_in = [x.to("cuda:%d" % k) for k, x in enumerate(xs)]
            d_out = [torch.empty_like(t) for t in d_in]
            for k in range(n):
                with torch.cuda.device(k):
                    ctxs[k].use_torch_stream()
                    assert lib.dlco_comm_allreduce_f32(comms[k], ctypes.c_void_p(d_in[k].data_ptr()),
                                                       ctypes.c_void_p(d_out[k].data_ptr()), count) == 0, lib.dlco_last_error()
            for k in range(n):
                assert lib.dlco_comm_check(comms[k]) == 0, lib.dlco_last_error()
            want = xs[0].clone()
            for x in xs[1:]:
                want = want + x                                  # rank order, fp32: the kernel's own order
            outs = [t.cpu() for t in d_out]
            for o in outs:
                assert torch.equal(o, want)                      # bit-identical on every rank
    # latency of the D x D all-reduce (all GPUs timed from GPU 0's events)
    d_in = [torch.randn(count_max, device="cuda:%d" % k) for k in range(n)]
    d_out = [torch.empty_like(t) for t in d_in]

    def once():
        for k in range(n):
            with torch.cuda.device(k):
                lib.dlco_comm_allreduce_f32(comms[k], ctypes.c_void_p(d_in[k].data_ptr()), ctypes.c_void_p(d_out[k].data_ptr()), count_max)
    for _ in range(5):
        once()
    for k in range(n):
        torch.cuda.synchronize(k)
    with torch.cuda.device(0):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    for _ in range(50):
        once()
    with torch.cuda.device(0):
        e1.record()
    for k in range(n):
        torch.cuda.synchronize(k)
    print("peer-memory all-reduce of %d floats over %d GPUs: %.1f us per call" % (count_max, n, e0.elapsed_time(e1) * 1000 / 50))
    for k in range(n):
        lib.dlco_comm_destroy(comms[k])
    for c in ctxs:
        c.close()


@needs2
def test_pj_learn_pair_sharded_matches_single_gpu(tmp_path):
    """`pj-learn -gpus 2` (pairs sharded, margins + gradient all-reduced over peer memory) against `pj-learn` on one GPU:
    the same validation points, ranks and model up to the drift of two free-running fp32 recursions whose gradient sums
    associate differently (a flipped near-tie hinge decision moves a run by ~1e-4 .. 1e-3, see
    tests/test_gpu_baseline_configs.py: the fp32 oracle drifts as much from its fp64 twin)."""
    import __graft_entry__
    __graft_entry__.build()
    from test_gpu_baseline_configs import synth_uprj
    x, lab = synth_uprj(30000, 160, 21)
    src = str(tmp_path / "train-unproj.h5")
    h5pack(src, ("Label", lab.reshape(-1, 1), 128, 1), ("Distance", x, 128, 0))
    outs = {}
    for tag, extra in (("one", []), ("two", ["-gpus", "2"])):
        dst = str(tmp_path / (tag + "-pj.h5"))
        r = subprocess.run([os.path.join(ROOT, "bin", "pj-learn"), "-iters", "400", "-mu", "0.001", "-gamma", "0.1", src, dst] + extra,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r.returncode == 0, r.stderr.decode()
        outs[tag] = (r.stdout.decode(), H5File(dst).read("W"), H5File(dst).read("A"))
    pat = re.compile(r"^(Best|Step): (\d+)  Loss: (\S+) Regul: (\S+) Obj: (\S+) \((\S+)\) Rank: (\d+) \((\d+)\)", re.M)
    a, b = pat.findall(outs["one"][0]), pat.findall(outs["two"][0])
    assert len(a) == len(b) == 4
    for la, lb in zip(a, b):
        assert la[1] == lb[1] and abs(int(la[6]) - int(lb[6])) <= 1
        for i in (2, 3, 4):
            assert abs(float(la[i]) - float(lb[i])) <= 1e-2 * abs(float(la[i])) + 1.5e-6, (la, lb)
    A1, A2 = outs["one"][2], outs["two"][2]
    assert abs(outs["one"][1].shape[0] - outs["two"][1].shape[0]) <= 1
    assert np.abs(A1 - A2).max() <= 2e-2 * np.abs(A1).max()
    sa = re.findall(r"^Stat: .*$", outs["one"][0], re.M)
    sb = re.findall(r"^Stat: .*$", outs["two"][0], re.M)
    assert len(sa) == len(sb) >= 1
    print("one GPU :", sa[-1], "\ntwo GPUs:", sb[-1])
