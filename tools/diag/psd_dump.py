"""dlco_pj_psd_project on seeded matrices, results saved: run once per environment setting and compare the files
(tests/test_gpu_learn.py::test_sign_sym_kernel_matches_the_generic_prep).    python tools/diag/psd_dump.py out.npz"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import dlco_b200

ctx = dlco_b200.Context(0)
dev = ctx.device
out = {}
for n_cfg, D in ((5, 544), (1, 544), (4, 300), (2, 100)):
    g = torch.Generator(device=dev)
    g.manual_seed(7 + D + n_cfg)
    xb = torch.randn((n_cfg, 400, D), generator=g, device=dev) * 0.1
    c = torch.randint(-200, 200, (n_cfg, 400, 1), generator=g, device=dev).float()
    A = -(10.0) * (torch.matmul((xb * c).transpose(1, 2), xb) / 40000 + 0.001 * torch.eye(D, device=dev))
    A = (0.5 * (A + A.transpose(1, 2))).contiguous()
    rank, S = ctx.pj_psd_project(A, want_sign=True)
    out["A_%d_%d" % (n_cfg, D)] = A.cpu().numpy()
    out["S_%d_%d" % (n_cfg, D)] = S.cpu().numpy()
    out["rank_%d_%d" % (n_cfg, D)] = rank.cpu().numpy()
np.savez(sys.argv[1], **out)
