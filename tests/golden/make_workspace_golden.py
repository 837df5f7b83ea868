#!/usr/bin/env python
"""Fixtures for running the reference's workspace scripts UNCHANGED against bin/ (tests/test_workspace_scripts.py,
tests/test_gpu_scripts.py).  Run HERE (needs /root/reference); the tests and the GPU box only read what this wrote.

    python tests/golden/make_workspace_golden.py

Written (tests/golden/):
  workspace_scripts.json         the reference's workspace/NN-*.sh, byte for byte with their sha256 (the compatibility
                                 contract the tools are tested against -- test infrastructure, not product; the
                                 tests write them into a temporary workspace and run them unchanged)
  export_kat.npz                 inputs / expected outputs of workspace/11-opencv-export.sh: W of the four exported
                                 pj-learn models, sha256 + size of workspace/opencv/vgg_generated_{48,64,80,120}.i
  yosemite-0.025-0.075-pr.h5     the second pr-learn model those exports use (5 KB, written by libhdf5)
  pj_select_kat.json             the reference's own pj-select-*.log / pj-top10.log next to the "[saved]" log lines they
                                 were derived from (one data set's 45 runs): pins 09-pjstats.sh / 10-pj-top.sh + the bc
                                 stand-in (tests/golden/bc) to the reference's committed results
"""
import hashlib
import json
import os
import re
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
WS = "/root/reference/workspace"

from oracle import h5lite  # noqa: E402


def sha(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def main():
    scripts = {}
    for name in sorted(os.listdir(WS)):
        if re.match(r"\d\d-.*\.sh$", name):
            scripts[name] = {"text": open(os.path.join(WS, name)).read(), "sha256": sha(os.path.join(WS, name))}
    json.dump(scripts, open(os.path.join(HERE, "workspace_scripts.json"), "w"), indent=1, sort_keys=True)
    manifest = scripts
    txt = open(os.path.join(WS, "11-opencv-export.sh")).read()
    out = {}
    rows = []
    for m in re.finditer(r"-prg (\S+) -id (\d+) \\\s*-prj (\S+) \\\s*(\S+\.i)", txt):
        prg, wid, prj, ifile = m.group(1), int(m.group(2)), m.group(3), m.group(4)
        tag = re.search(r"_(\d+)\.i$", ifile).group(1)
        out["W" + tag] = h5lite.read(os.path.join(WS, prj), "W")
        rows.append({"i_file": ifile, "prg": prg, "id": wid, "prj": prj, "tag": tag,
                     "sha256": sha(os.path.join(WS, ifile)), "size": os.path.getsize(os.path.join(WS, ifile))})
    out["table"] = np.array(json.dumps(rows))
    np.savez_compressed(os.path.join(HERE, "export_kat.npz"), **out)
    src = os.path.join(WS, "pr-learn/olderbest/yosemite-0.025-0.075-pr.h5")
    shutil.copyfile(src, os.path.join(HERE, "yosemite-0.025-0.075-pr.h5"))
    os.chmod(os.path.join(HERE, "yosemite-0.025-0.075-pr.h5"), 0o644)

    # 09-pjstats.sh / 10-pj-top.sh on the reference's own logs: last "[saved]" line of every pj-learn log of the three
    # runs-on-"liberty-0.035-0.250-pr#7" data sets, and the select / top10 logs the reference derived from all 405
    logs = {}
    logdir = os.path.join(WS, "pj-learn", "logging")
    for name in sorted(os.listdir(logdir)):
        lines = [ln.rstrip("\n") for ln in open(os.path.join(logdir, name)) if "saved" in ln]
        logs[name] = lines[-1] if lines else ""
    sel = {n: open(os.path.join(WS, n)).read() for n in
           ("pj-select-liberty.log", "pj-select-notredame.log", "pj-select-yosemite.log", "pj-top10.log")}
    json.dump({"last_saved": logs, "outputs": sel}, open(os.path.join(HERE, "pj_select_kat.json"), "w"), indent=0)
    print("written:", sorted(manifest), "export_kat.npz", os.path.getsize(os.path.join(HERE, "export_kat.npz")),
          "pj_select_kat.json", os.path.getsize(os.path.join(HERE, "pj_select_kat.json")))


if __name__ == "__main__":
    main()
