#!/usr/bin/env python
"""comp-fulldists AS A PROCESS: HDF5 read of the patches, the full build, chunked + deflated
"Distance" write -- the numbers the device-resident bench leaves out.  Writes one JSON line.

  python tools/measure_comp_fulldists.py [--pairs 500000] [--gzip 1] [--gpus 1] [--dir /tmp/dlco_tool]

Inputs are synthetic (bench.synth_patches / synth_pairs) written through bin/h5tool pack; filters.h5 comes from
bin/gen-poolregion (the default bank)."""
import argparse
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BIN = os.path.join(ROOT, "bin")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=500000)
    ap.add_argument("--gzip", type=int, default=1)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--dir", default="/tmp/dlco_tool")
    args = ap.parse_args()
    import torch
    import bench as B
    os.makedirs(args.dir, exist_ok=True)
    flt = os.path.join(args.dir, "filters.h5")
    env = dict(os.environ, DLCO_GZIP=str(args.gzip))
    res = {"pairs": args.pairs, "gzip": args.gzip, "gpus": args.gpus}
    if not os.path.exists(flt):
        t = time.time()
        subprocess.check_call([os.path.join(BIN, "gen-poolregion"), flt], stdout=subprocess.DEVNULL, env=dict(os.environ, DLCO_GZIP="1"))
        res["gen_poolregion_s"] = time.time() - t
        res["filters_h5_bytes"] = os.path.getsize(flt)
    img = os.path.join(args.dir, "img-%d.h5" % args.pairs)
    if not os.path.exists(img):
        dev = torch.device("cuda", 0)
        n_points = max(64, int(args.pairs * 0.9) // 3)
        patches = B.synth_patches(torch, dev, n_points, 3).cpu().numpy()
        idx = B.synth_pairs(torch, dev, n_points, 3, args.pairs).cpu().numpy()
        patches.tofile(img + ".p.bin")
        idx.tofile(img + ".i.bin")
        subprocess.check_call([os.path.join(BIN, "h5tool"), "pack", img,
                               "Indices", "i32", "2", str(idx.shape[0]), "4", img + ".i.bin", "256", "0",
                               "Patches", "u8", "3", str(patches.shape[0]), "64", "64", img + ".p.bin", "256", "1"])
        os.remove(img + ".p.bin")
        os.remove(img + ".i.bin")
        del patches, idx
        torch.cuda.empty_cache()
    res["img_h5_bytes"] = os.path.getsize(img)
    dst = os.path.join(args.dir, "dist.h5")
    if os.path.exists(dst):
        os.remove(dst)
    cmd = [os.path.join(BIN, "comp-fulldists"), flt, img, dst]
    if args.gpus > 1:
        cmd += ["-gpus", str(args.gpus)]
    t = time.time()
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env)
    res["wall_s"] = time.time() - t
    res["rc"] = r.returncode
    so, se = r.stdout.decode(), r.stderr.decode()
    m = re.search(r"Total: ([0-9.]+) sec", so)
    res["tool_total_s"] = float(m.group(1)) if m else None          # compute + Distance write (the reference's own timer)
    m = re.search(r"([0-9.]+) pairs/s including HDF5 output", se)
    res["pairs_per_s_incl_hdf5_write"] = float(m.group(1)) if m else None
    res["pairs_per_s_whole_process"] = args.pairs / res["wall_s"]
    res["dist_h5_bytes"] = os.path.getsize(dst) if os.path.exists(dst) else None
    res["host_threads"] = os.cpu_count()
    res["stderr_tail"] = se[-300:]
    os.remove(dst)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
