#!/usr/bin/env python
"""Builds the committed golden fixtures from the reference tree's own run outputs.

Run HERE (the container that has /root/reference); the GPU box and the test suite only read
the fixtures this script wrote:

    python tests/golden/make_golden.py

Sources (all read-only data artefacts of the reference, no source code is copied):
  workspace/opencv/vgg_generated_{48,64,80,120}.i   exported PR filters (hex float bits) + PJ
  workspace/11-opencv-export.sh                      which w row / pj.h5 each .i came from
  workspace/pr-learn/*.h5 + logging/*.log            learnt w rows <-> "[saved]" log lines
  workspace/pj-learn/*.h5 + logging/*.log            learnt W, A <-> last "[saved]" log line
  cv2 (OpenCV 4.13 in this image)                    the OpenCV primitives the reference calls

Fixtures written (tests/golden/):
  pr_select_kat.json     per .i file: rings with w > 0, expected row count, sha256 of the dense
                         selected-filter matrix rebuilt from PRidx[] / PR[]
  pr_w_rows.json         liberty-0.035-0.250-pr.h5: sparse w rows + the log lines they pair with
  liberty-0.035-0.250-pr.h5   the 5 KB HDF5 file itself (h5lite reader KAT: chunked, gzip-9,
                         unlimited rows)
  pj_model_kat.npz       W (rank x D) of the 48-D model, every 8th row of A, trace(A), log line
  cv2_kat.npz            GaussianBlur / randShuffle / contourArea / sort outputs of cv2 on seeded inputs
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
REF = "/root/reference"
WS = os.path.join(REF, "workspace")

from oracle import h5lite  # noqa: E402


def parse_i_file(path):
    txt = open(path).read()

    def arr(name):
        m = re.search(r"static const unsigned int %s\[\] =\s*\{(.*?)\};" % name, txt, re.S)
        return [int(t, 16) for t in re.findall(r"0x[0-9a-fA-F]+", m.group(1))]

    def scalar(name):
        return int(re.search(r"static const int %s = (\d+);" % name, txt).group(1))

    out = {}
    for pre in ("PR", "PJ"):
        rows, cols = scalar(pre + "rows"), scalar(pre + "cols")
        idx, vals = arr(pre + "idx"), arr(pre)
        dense = np.zeros(rows * cols, np.uint32)
        p = 0
        for start, count in zip(idx[0::2], idx[1::2]):
            dense[start:start + count] = vals[p:p + count]
            p += count
        assert p == len(vals)
        out[pre] = dense.view(np.float32).reshape(rows, cols)
    return out


def export_table():
    """(i file, prg h5, id, prj h5) from workspace/11-opencv-export.sh"""
    txt = open(os.path.join(WS, "11-opencv-export.sh")).read()
    out = []
    for m in re.finditer(r"-prg (\S+) -id (\d+) \\\s*-prj (\S+) \\\s*(\S+\.i)", txt):
        out.append((m.group(4), m.group(1), int(m.group(2)), m.group(3)))
    return out


def main():
    kat = []
    pj_done = False
    for ifile, prg, wid, prj in export_table():
        parsed = parse_i_file(os.path.join(WS, ifile))
        w = h5lite.read(os.path.join(WS, prg), "w")[wid]
        rings = np.nonzero(w > 0)[0]
        kat.append({
            "i_file": ifile, "prg": prg, "id": wid, "prj": prj,
            "rings": rings.tolist(), "w": [float(v) for v in w[rings]],
            "rows": int(parsed["PR"].shape[0]),
            "sha256": hashlib.sha256(parsed["PR"].tobytes()).hexdigest(),
            "pj_shape": list(parsed["PJ"].shape),
        })
        if not pj_done:                      # the 48-D model: smallest W / A
            W = h5lite.read(os.path.join(WS, prj), "W")
            A = h5lite.read(os.path.join(WS, prj), "A")
            assert np.array_equal(parsed["PJ"][parsed["PJ"] != 0], W[W != 0])
            log = os.path.join(WS, "pj-learn", "logging", os.path.basename(prj).replace(".h5", ".log"))
            saved = [ln.strip() for ln in open(log) if "saved" in ln]
            best = [ln.strip() for ln in open(log) if ln.startswith("Best:")]
            np.savez_compressed(os.path.join(HERE, "pj_model_kat.npz"), W=W, A_rows=A[::8].copy(),
                                A_trace=np.float64(np.trace(A.astype(np.float64))), A_shape=np.array(A.shape),
                                A_sym_err=np.float64(np.abs(A - A.T).max()),
                                last_saved=np.array(saved[-1]), best_lines=np.array(best[-3:]),
                                source=np.array(prj))
            pj_done = True
    json.dump(kat, open(os.path.join(HERE, "pr_select_kat.json"), "w"), indent=1)

    # pr-learn: w rows <-> log lines
    name = "liberty-0.035-0.250-pr"
    src = os.path.join(WS, "pr-learn", name + ".h5")
    shutil.copyfile(src, os.path.join(HERE, name + ".h5"))
    os.chmod(os.path.join(HERE, name + ".h5"), 0o644)
    w = h5lite.read(src, "w")
    log = [ln.rstrip("\n") for ln in open(os.path.join(WS, "pr-learn", "logging", name + ".log"))]
    pairs = []
    for i, ln in enumerate(log):
        if ln.startswith("Stat:") and "[saved]" in ln:
            pairs.append({"best": log[i - 1], "stat": ln})
    rows = [{"idx": np.nonzero(r)[0].tolist(), "val": [float(v) for v in r[np.nonzero(r)[0]]]} for r in w]
    json.dump({"file": name + ".h5", "mu": 0.035, "gamma": 0.25, "shape": list(w.shape), "rows": rows,
               "saved_lines": pairs[:len(rows)], "header": log[:16]},
              open(os.path.join(HERE, "pr_w_rows.json"), "w"), indent=1)

    # OpenCV primitives the oracle restates
    import cv2
    rng = np.random.default_rng(12345)
    img = rng.integers(0, 256, (64, 64)).astype(np.float32)
    blur = cv2.GaussianBlur(img, (0, 0), 1.4, borderType=cv2.BORDER_REPLICATE)
    gk = cv2.getGaussianKernel(13, 1.4, cv2.CV_32F).reshape(-1)
    ix = cv2.filter2D(blur, cv2.CV_32F, np.array([[-1, 0, 1]], np.float32), borderType=cv2.BORDER_REPLICATE)
    iy = cv2.filter2D(blur, cv2.CV_32F, np.array([[-1], [0], [1]], np.float32), borderType=cv2.BORDER_REPLICATE)
    mag = cv2.magnitude(ix, iy)
    # randShuffle on theRNG() in its default state (0xffffffff): only a FRESH process has it
    import subprocess
    import tempfile
    tmp = os.path.join(tempfile.mkdtemp(), "shuf.npz")
    subprocess.check_call([sys.executable, "-c",
                           "import cv2, numpy as np\n"
                           "a = np.arange(1000, dtype=np.int32).reshape(1, -1)\n"
                           "b = np.arange(777, dtype=np.int32).reshape(1, -1)\n"
                           "cv2.randShuffle(a); cv2.randShuffle(b)\n"
                           "np.savez(%r, a=a, b=b)\n" % tmp])
    sh = np.load(tmp)
    a, b = sh["a"], sh["b"]
    pts = np.stack([np.sort(rng.random(500)), np.sort(rng.random(500))], 1).astype(np.float32)
    pts = np.vstack([pts, [[1.0, 0.0]]]).astype(np.float32)
    area = cv2.contourArea(pts)
    np.savez_compressed(os.path.join(HERE, "cv2_kat.npz"), img=img, blur=blur, gk=gk, ix=ix, iy=iy, mag=mag,
                        shuf_a=a.reshape(-1), shuf_b=b.reshape(-1), pts=pts, area=np.float64(area),
                        cv2_version=np.array(cv2.__version__))
    print("fixtures written to", HERE)


if __name__ == "__main__":
    main()
