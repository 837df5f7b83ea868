"""time of the matrix-sign projection of 45 (and 1) D = 544 matrices under the current environment (experiments)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import dlco_b200
from bench_paths import timeit
ctx = dlco_b200.Context(0); dev = ctx.device
for n_cfg, D in ((45, 544), (1, 544)):
    g = torch.Generator(device=dev); g.manual_seed(1)
    xb = torch.randn((n_cfg, 400, D), generator=g, device=dev) * 0.1
    c = torch.randint(-200, 200, (n_cfg, 400, 1), generator=g, device=dev).float()
    A0 = -(10.0) * (torch.matmul((xb * c).transpose(1, 2), xb) / 40000 + 0.001 * torch.eye(D, device=dev))
    A0 = (0.5 * (A0 + A0.transpose(1, 2))).contiguous()
    A = A0.clone()
    def f():
        A.copy_(A0); ctx.pj_psd_project(A)
    ms, best = timeit(torch, f, warm=2, rep=8)
    cp = timeit(torch, lambda: A.copy_(A0), warm=2, rep=8)[0]
    print("n_cfg=%d psd_sign_ms=%.4f" % (n_cfg, ms - cp))
