import sys, json
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import dlco_b200
from bench_paths import timeit
ctx = dlco_b200.Context(0); dev = ctx.device
res = {}
for n_cfg, D in ((45, 544), (1, 544), (45, 480), (8, 544)):
    g = torch.Generator(device=dev); g.manual_seed(1)
    xb = torch.randn((n_cfg, 400, D), generator=g, device=dev) * 0.1
    c = torch.randint(-200, 200, (n_cfg, 400, 1), generator=g, device=dev).float()
    A0 = -(10.0) * (torch.matmul((xb * c).transpose(1, 2), xb) / 40000 + 0.001 * torch.eye(D, device=dev))
    A0 = (0.5 * (A0 + A0.transpose(1, 2))).contiguous()
    A = A0.clone()
    def f():
        A.copy_(A0); ctx.pj_psd_project(A)
    ms, best = timeit(torch, f, warm=2, rep=5)
    cp = timeit(torch, lambda: A.copy_(A0), warm=2, rep=5)[0]
    ev = timeit(torch, lambda: ctx.pj_eig_batched(A0), warm=1, rep=3)[0]
    w = torch.linalg.eigvalsh(A0.double())
    ref = None
    ctx.pj_psd_project(A.copy_(A0))
    wv, V = torch.linalg.eigh(A0.double())
    ref = (V * wv.clamp(min=0)[:, None, :]) @ V.transpose(1, 2)
    err = ((A.double() - ref).abs().amax(dim=(1, 2)) / ref.abs().amax(dim=(1, 2))).max().item()
    res["%dx%d" % (n_cfg, D)] = {"psd_sign_ms": ms - cp, "cusolver_eig_ms": ev, "relerr": err,
                                 "gemm_tflops_fp32_equiv": 28 * 2.0 * D**3 * n_cfg / ((ms - cp) * 1e-3) / 1e12}
print(json.dumps(res))
