import sys, json
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
import dlco_b200, bench as B
ctx = dlco_b200.Context(0); dev = ctx.device
full = B.synth_bank(torch, dev)
sel = full[(torch.arange(64, device=dev) * 601 + 16000) % full.shape[0]].contiguous(); del full
bank = ctx.bank(sel)
n = 100000
npnt = n * 9 // 30
patches = B.synth_patches(torch, dev, npnt, 3); idx = B.synth_pairs(torch, dev, npnt, 3, n)
out = torch.empty((n, 8), dtype=torch.float32, device=dev)
for _ in range(2): ctx.fulldist_build(bank, patches, idx, out=out)
torch.cuda.synchronize(); ctx.profile(True)
for _ in range(3): ctx.fulldist_build(bank, patches, idx, out=out)
torch.cuda.synchronize()
d, dn = ctx.profile_read(0); p, pn = ctx.profile_read(1)
print(json.dumps({"desc_ns_per_patch": d / 3 / (2 * n) * 1e6, "pool_ms_per_pass": p / 3, "desc_ms_per_pass": d / 3, "checksum": float(out.double().sum().item())}))
