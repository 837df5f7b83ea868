#!/usr/bin/env python
"""The whole (mu, gamma) sweep of workspace/08-pjlearn.sh as one job on the GPUs of this box: 9 unprojected-distance
data sets (3 per name, as the reference's top-3 pr models give) x 45 configurations at D = 544, `bin/pj-sweep -prox sign`
one data set per GPU, then the reference's UNCHANGED 09-pjstats.sh / 10-pj-top.sh (tests/golden/workspace_scripts.json)
on the logs.  One data set is also run with `-prox eig` (on the GPU that is free in the second round) and the two
routes' per-bucket selections are compared.  Prints one JSON line: wall time, config-iterations/s, the selections.

    python tools/sweep_9x45.py [--iters 5000] [--pairs 100000] [--gpus 8] [--out gpurun_out/r02_sweep_9x45.json]

Synthetic data (there is no data set on the box): low-rank informative subspace + noise, seeded per data set."""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

MUS = "0.0001,0.0005,0.0010,0.0020,0.0030"
GAMMAS = "0.025,0.050,0.100,0.150,0.200,0.250,0.500,0.750,1.000"
NAMES = ("yosemite", "notredame", "liberty")
MODELS = ("0.025-0.075-pr#3", "0.035-0.250-pr#7", "0.050-0.100-pr#1")


def synth_set(seed, n, d):
    rng = np.random.default_rng(seed)
    lab = (np.arange(n) % 2 == 0).astype(np.uint8)
    k = 96
    basis = (rng.standard_normal((k, d)) / np.sqrt(d)).astype(np.float32)
    sc = np.ones((n, k), np.float32)
    sc[:, :32] = np.where(lab[:, None] == 1, 0.45, 1.0)
    x = (rng.standard_normal((n, k)).astype(np.float32) * sc) @ basis + rng.standard_normal((n, d)).astype(np.float32) * 0.04
    return np.clip(x, -1, 1).astype(np.float32), lab


def parse_top(path):
    out, cur = {}, None
    for ln in open(path):
        m = re.match(r"TOP maxdim #(\d+) minimisers:", ln)
        if m:
            cur = int(m.group(1))
            out[cur] = []
            continue
        m = re.match(r" mean: (\S+)  YO:(\S+) ND:(\S+) LY:(\S+) dim #\[(\d+)/(\d+)/(\d+)\] AUC\[\S+\] (\S+)$", ln)
        if m and cur is not None:
            out[cur].append((float(m.group(1)), m.group(8)))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=5000)
    ap.add_argument("--pairs", type=int, default=60000)
    ap.add_argument("--dim", type=int, default=544)
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--eig-mus", default="0.0005,0.0020", help="the mu rows the eig route repeats for the comparison")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r02_sweep_9x45.json"))
    args = ap.parse_args()
    import __graft_entry__
    __graft_entry__.build()
    from test_workspace_scripts import h5pack, make_workspace
    tmp = tempfile.mkdtemp(prefix="sweep9x45_")
    ws, env = make_workspace(tmp)
    ws_eig = os.path.join(tmp, "workspace_eig")
    shutil.copytree(ws, ws_eig)
    for w in (ws, ws_eig):
        os.makedirs(os.path.join(w, "distances"), exist_ok=True)
        os.makedirs(os.path.join(w, "pj-learn", "logging"), exist_ok=True)
    t_gen = time.time()
    sets = []
    for mi, model in enumerate(MODELS):
        for ni, name in enumerate(NAMES):
            ds = "%s-%s-%s" % (name, NAMES[mi], model)             # evaluated on `name`, pr model learnt on NAMES[mi]
            x, lab = synth_set(9000 + 10 * mi + ni, args.pairs, args.dim)
            path = os.path.join(ws, "distances", ds + "-unproj.h5")
            h5pack(path, ("Label", lab.reshape(-1, 1), 128, 1), ("Distance", x, 128, 0))
            sets.append((ds, path))
    t_gen = time.time() - t_gen
    sweep = os.path.join(ROOT, "bin", "pj-sweep")
    jobs = [("sign", ds, path, ws, MUS) for ds, path in sets]
    jobs += [("eig", ds, path, ws_eig, args.eig_mus) for ds, path in sets[:3]]          # the three data sets of model 0
    running, done, free = [], [], list(range(args.gpus))
    t0 = time.time()
    t_sign_done = None
    queue = list(jobs)
    while queue or running:
        while queue and free:
            route, ds, path, w, mus = queue.pop(0)
            gpu = free.pop(0)
            log = open(os.path.join(tmp, "%s-%s.out" % (route, ds.replace("#", "_"))), "w")
            p = subprocess.Popen([sweep, path, os.path.join(w, "pj-learn"), "-iters", str(args.iters), "-mu", mus, "-gamma", GAMMAS,
                                  "-prox", route, "-gpu", str(gpu)], stdout=log, stderr=subprocess.STDOUT)
            running.append((p, route, ds, gpu, time.time(), log))
        time.sleep(0.2)
        for job in list(running):
            p, route, ds, gpu, ts, log = job
            if p.poll() is None:
                continue
            running.remove(job)
            log.close()
            if p.returncode != 0:
                print(open(log.name).read()[-2000:], file=sys.stderr)
                raise SystemExit("pj-sweep %s on %s failed (rc %d)" % (route, ds, p.returncode))
            done.append({"route": route, "ds": ds, "gpu": gpu, "seconds": time.time() - ts})
            free.append(gpu)
            if route == "sign" and sum(1 for d in done if d["route"] == "sign") == len(sets):
                t_sign_done = time.time() - t0
    t_all = time.time() - t0
    tops = {}
    for route, w in (("sign", ws), ("eig", ws_eig)):
        for script in ("09-pjstats.sh", "10-pj-top.sh"):                      # unchanged, byte for byte
            r = subprocess.run(["bash", script], cwd=w, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            if r.returncode != 0:
                raise SystemExit("%s failed: %s" % (script, r.stderr.decode()[-1000:]))
        tops[route] = parse_top(os.path.join(w, "pj-top10.log"))
    # selection per dimension bucket: over everything (sign), and sign vs eig over the configurations both ran
    eig_labels = {lab for rows in tops["eig"].values() for _, lab in rows}
    sel, agree = {}, {}
    for maxdim in (48, 64, 80, 120):
        rows = tops["sign"].get(maxdim, [])
        sel[str(maxdim)] = {"n_fit": len(rows), "best": rows[0][1] if rows else None, "mean_fpr95": rows[0][0] if rows else None}
        common = [r for r in rows if r[1] in eig_labels]
        e_rows = tops["eig"].get(maxdim, [])
        tab_e = {lab: m for m, lab in e_rows}
        agree[str(maxdim)] = {
            "sign_best": common[0][1] if common else None, "sign_mean": common[0][0] if common else None,
            "eig_best": e_rows[0][1] if e_rows else None, "eig_mean": e_rows[0][0] if e_rows else None,
            "same_set_fits": sorted(lab for _, lab in common) == sorted(tab_e),
            "max_abs_diff_mean_fpr95_pp": max([abs(m - tab_e[lab]) for m, lab in common if lab in tab_e], default=None),
        }
    n_cfg = 45
    res = {
        "what": "9 data sets x 45 (mu, gamma) configurations, D = %d, %d pairs each, %d iterations, bin/pj-sweep -prox sign, one "
                "data set per GPU; then the unchanged 09-pjstats.sh / 10-pj-top.sh" % (args.dim, args.pairs, args.iters),
        "gpus": args.gpus, "iters": args.iters, "data_gen_seconds": t_gen,
        "sign_sweep_wall_seconds": t_sign_done, "all_jobs_wall_seconds": t_all,
        "sign_config_iterations_per_s": len(sets) * n_cfg * args.iters / t_sign_done if t_sign_done else None,
        "jobs": done, "selection_sign": sel, "sign_vs_eig_on_model0": agree,
        "eig_repeats": "model 0's three data sets, mu in {%s} x 9 gamma, same iterations" % args.eig_mus,
    }
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(res, f, indent=1)
    shutil.copy(os.path.join(ws, "pj-top10.log"), args.out.replace(".json", "_pj-top10.log"))
    print(json.dumps(res))
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
