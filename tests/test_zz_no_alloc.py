"""dlco_roc_stats_static (include/dlco_b200.h): the FPR95 / AUC evaluation with its sort scratch in a fixed __device__
array, for callers that cannot allocate.  Not in the reference; checked against dlco_roc_stats, the existing entry
point, through the C-ABI, on two calls in a row that share the array.  Collected last on purpose (pytest -x)."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _case(seed, n):
    rng = np.random.default_rng(seed)
    lab = (rng.random(n) < 0.4).astype(np.uint8)
    pd = (rng.gamma(2.0, 1.0, n) * np.where(lab == 1, 0.5, 1.0)).astype(np.float32)
    pd[::7] = pd[3]                                           # ties: the stable order decides
    return pd, lab


def _call(ctx, fn, d_pd, d_lab, n):
    f, a = ctypes.c_float(), ctypes.c_double()
    rc = fn(ctx.handle, ctypes.c_void_p(d_pd.data_ptr()), ctypes.c_void_p(d_lab.data_ptr()), n, ctypes.byref(f), ctypes.byref(a))
    return rc, f.value, a.value


def test_roc_stats_static_matches_roc_stats_on_two_calls(gpu_ctx):
    lib = gpu_ctx.lib
    max_n = lib.dlco_roc_stats_static_max_n()
    assert max_n == 1 << 20
    # the second call is smaller than the first, so the array still holds the first call's keys beyond its own
    cases = [_case(1, 300001), _case(2, 4097), _case(3, max_n)]
    dev = [(torch.from_numpy(pd).to(gpu_ctx.device), torch.from_numpy(lab).to(gpu_ctx.device)) for pd, lab in cases]
    want = []
    for d_pd, d_lab in dev:
        rc, f, a = _call(gpu_ctx, lib.dlco_roc_stats, d_pd, d_lab, d_pd.numel())
        assert rc == 0
        want.append((f, a))
    gpu_ctx.sync()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(gpu_ctx.device)[0]
    got = []
    for d_pd, d_lab in dev:
        rc, f, a = _call(gpu_ctx, lib.dlco_roc_stats_static, d_pd, d_lab, d_pd.numel())
        assert rc == 0
        got.append((f, a))
    assert torch.cuda.mem_get_info(gpu_ctx.device)[0] == free0          # nothing was allocated on the device
    for (f, a), (fw, aw) in zip(got, want):
        assert f == fw                                        # the bar of tests/test_gpu_learn.py::test_roc_stats
        assert abs(a - aw) <= 1e-12
    assert 0.0 < got[0][0] < 1.0 and 0.5 < got[0][1] <= 1.0


def test_roc_stats_static_rejects_sizes_beyond_its_scratch(gpu_ctx):
    lib = gpu_ctx.lib
    pd, lab = _case(4, 1000)
    d_pd, d_lab = torch.from_numpy(pd).to(gpu_ctx.device), torch.from_numpy(lab).to(gpu_ctx.device)
    rc, _, _ = _call(gpu_ctx, lib.dlco_roc_stats_static, d_pd, d_lab, lib.dlco_roc_stats_static_max_n() + 1)
    assert rc != 0
    assert b"exceeds the static scratch" in lib.dlco_last_error()
    rc, f, a = _call(gpu_ctx, lib.dlco_roc_stats_static, d_pd, d_lab, 1000)      # and the entry point still works after it
    rc2, fw, aw = _call(gpu_ctx, lib.dlco_roc_stats, d_pd, d_lab, 1000)
    assert rc == 0 and rc2 == 0 and f == fw and abs(a - aw) <= 1e-12
