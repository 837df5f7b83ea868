"""one configuration of the pr-learn RDA chain, 100 000 steps at F = 5120 (ncu target / timing)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
import dlco_b200
import bench_paths as bp
ctx = dlco_b200.Context(0); dev = ctx.device
n, F = (int(sys.argv[2]) if len(sys.argv) > 2 else 100000), 5120
d, lab = bp.synth_dists(torch, dev, n, F, 3)
lab_h = np.asarray(lab)
pos = torch.from_numpy(np.nonzero(lab_h == 1)[0].astype(np.int32)).to(dev)
neg = torch.from_numpy(np.nonzero(lab_h == 0)[0].astype(np.int32)).to(dev)
w = torch.zeros((F,), device=dev); df = torch.zeros((F,), device=dev)
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
st = ctx.pr_rda_run(d, pos, pos.numel(), neg, neg.numel(), 0.035, 0.25, 0, 2000, 2215, w, df)
torch.cuda.synchronize(); t0 = time.perf_counter()
ctx.pr_rda_run(d, pos, pos.numel(), neg, neg.numel(), 0.035, 0.25, 2000, steps, st, w, df)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("%.3f us per step, nnz %d" % (dt / steps * 1e6, int((w != 0).sum())))
