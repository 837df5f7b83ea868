"""The reference's workspace scripts, UNCHANGED (tests/golden/workspace_scripts.json holds them byte for byte with
their sha256), run against bin/ in a temporary workspace -- the part that needs no GPU:

  * 11-opencv-export.sh -> export-opencv must regenerate workspace/opencv/vgg_generated_{48,64,80,120}.i from the
    reference's own pr-learn / pj-learn models (fixtures) byte for byte: an end-to-end bit-exact check of h5mini's
    reader on libhdf5-written files, SelectPRFilters, the W convention and the exporter;
  * 09-pjstats.sh + 10-pj-top.sh on the reference's own pj-learn logs reproduce its committed pj-select-*.log and
    pj-top10.log (pins the awk-column contract of the "[saved]" lines and the bc stand-in this image needs);
  * conv-impatches on a synthetic Brown-layout directory.

The GPU half (02 .. 10 on synthetic data sets) is tests/test_gpu_scripts.py."""
import hashlib
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import ROOT, make_bank
from oracle import dlco_oracle as O
from oracle.h5lite import H5File

G = os.path.join(ROOT, "tests", "golden")
BIN = os.path.join(ROOT, "bin")


def materialise_scripts(ws):
    """Write the reference's scripts into `ws` byte for byte (sha256 checked)."""
    scripts = json.load(open(os.path.join(G, "workspace_scripts.json")))
    for name, rec in scripts.items():
        data = rec["text"].encode()
        assert hashlib.sha256(data).hexdigest() == rec["sha256"], name
        with open(os.path.join(ws, name), "wb") as f:
            f.write(data)
    return sorted(scripts)


def make_workspace(tmp):
    """tmp/bin -> the repo's bin (the scripts call ../bin/<tool>), tmp/shim/bc, tmp/workspace/."""
    ws = os.path.join(tmp, "workspace")
    os.makedirs(ws)
    os.symlink(BIN, os.path.join(tmp, "bin"))
    os.makedirs(os.path.join(tmp, "shim"))
    shutil.copy(os.path.join(G, "bc"), os.path.join(tmp, "shim", "bc"))
    materialise_scripts(ws)
    env = dict(os.environ)
    if shutil.which("bc") is None:                      # this image has no bc(1): tests/golden/bc stands in
        env["PATH"] = os.path.join(tmp, "shim") + os.pathsep + env["PATH"]
    env["LC_ALL"] = "C"
    return ws, env


def h5pack(path, *sets):
    """sets: (name, array, chunk_rows, deflate) through bin/h5tool pack (h5mini::Writer)."""
    args = [os.path.join(BIN, "h5tool"), "pack", path]
    tmpfiles = []
    for name, arr, chunk, level in sets:
        arr = np.ascontiguousarray(arr)
        raw = path + "." + name + ".bin"
        arr.tofile(raw)
        tmpfiles.append(raw)
        dt = {np.dtype(np.uint8): "u8", np.dtype(np.int32): "i32", np.dtype(np.float32): "f32",
              np.dtype(np.float64): "f64"}[arr.dtype]
        args += [name, dt, str(arr.ndim)] + [str(d) for d in arr.shape] + [raw, str(chunk), str(level)]
    subprocess.check_call(args)
    for f in tmpfiles:
        os.remove(f)


@pytest.fixture(scope="module")
def tools_built():
    import __graft_entry__
    __graft_entry__.build()
    for t in ("export-opencv", "conv-impatches", "h5tool", "gen-poolregion"):
        assert os.path.exists(os.path.join(BIN, t)), t


def test_scripts_fixture_is_complete():
    names = sorted(json.load(open(os.path.join(G, "workspace_scripts.json"))))
    assert names == ["01-export.sh", "02-genpoolregs.sh", "03-compdist.sh", "04-prlearn.sh", "05-prstats.sh",
                     "06-pr-top.sh", "07-compunproj.sh", "08-pjlearn.sh", "09-pjstats.sh", "10-pj-top.sh",
                     "11-opencv-export.sh"]


def test_export_opencv_script_reproduces_reference_headers(tools_built, oracle_lib, tmp_path):
    ws, env = make_workspace(str(tmp_path))
    z = np.load(os.path.join(G, "export_kat.npz"))
    table = json.loads(str(z["table"]))
    # filters.h5: the rings any exported model selects, bit-exact from the oracle; all other rows zero
    kat = json.load(open(os.path.join(G, "pr_select_kat.json")))
    rings = O.ring_params()
    used = sorted(set(sum([k["rings"] for k in kat], [])))
    filt, _ = make_bank(oracle_lib, rings[used])
    full = np.zeros((40960, 64, 64), np.float32)
    for j, r in enumerate(used):
        full[8 * r:8 * r + 8] = filt[8 * j:8 * j + 8].reshape(8, 64, 64)
    h5pack(os.path.join(ws, "filters.h5"), ("PRFilters", full, 128, 1))
    del full
    os.makedirs(os.path.join(ws, "pr-learn", "olderbest"))
    os.makedirs(os.path.join(ws, "pj-learn"))
    os.makedirs(os.path.join(ws, "opencv"))
    shutil.copy(os.path.join(G, "yosemite-0.025-0.075-pr.h5"), os.path.join(ws, "pr-learn", "olderbest"))
    shutil.copy(os.path.join(G, "liberty-0.035-0.250-pr.h5"), os.path.join(ws, "pr-learn"))
    for row in table:                                                     # the pj models: W as pj-learn writes it
        h5pack(os.path.join(ws, row["prj"]), ("W", z["W" + row["tag"]], 0, 0))
    r = subprocess.run(["bash", "11-opencv-export.sh"], cwd=ws, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()
    out = r.stdout.decode()
    assert out.count("Load PRFilters:\n0...10...20...30...40...50...60...70...80...90...100 - done.\n") == 4
    assert "PRFilters: 60 x 4096 [480]\nPJFilters: 48 x [480]\n" in out and "PJFilters: 120 x [544]\n" in out
    for row in table:
        data = open(os.path.join(ws, row["i_file"]), "rb").read()
        # the committed .i files end in one more newline than src/export-opencv.cpp:383-386 writes ("\n};\n" then
        # fclose): everything else is identical byte for byte
        assert len(data) + 1 == row["size"], row["i_file"]
        assert hashlib.sha256(data + b"\n").hexdigest() == row["sha256"], row["i_file"]


def test_pjstats_and_pjtop_scripts_on_reference_logs(tmp_path):
    """09-pjstats.sh takes the LAST "[saved]" line of every pj-learn log and reads awk fields $6 (AUC), $9 (FPR95),
    $3 ([Dim]); 10-pj-top.sh averages FPR95 over the three data sets with bc.  On the reference's own 405 logs the
    unchanged scripts must give the reference's own pj-select-*.log / pj-top10.log."""
    ws, env = make_workspace(str(tmp_path))
    kat = json.load(open(os.path.join(G, "pj_select_kat.json")))
    os.makedirs(os.path.join(ws, "pj-learn", "logging"))
    for name, last in kat["last_saved"].items():
        with open(os.path.join(ws, "pj-learn", "logging", name), "w") as f:
            f.write("Step: 100  Loss: 0.5 Regul: 0.1 Obj: 0.6 (0.6) Rank: 99 (99) Ttime: 0.1 Vtime: 0.1\n")
            f.write("Stat: Dim [99] AUC: 0.500000 (0.500000) FPR95: 99.00 (99.00) [saved]\n")   # an earlier save
            if last:
                f.write(last + "\n")
    for script in ("09-pjstats.sh", "10-pj-top.sh"):
        r = subprocess.run(["bash", script], cwd=ws, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert r.returncode == 0, r.stderr.decode()
    for ds in ("liberty", "notredame", "yosemite"):
        got = open(os.path.join(ws, "pj-select-%s.log" % ds)).read().splitlines()
        want = kat["outputs"]["pj-select-%s.log" % ds].splitlines()
        # `find` lists the logs in directory order: compare the ModelStat lines as sets, the rest in place
        assert sorted(ln for ln in got if "ModelStat" in ln) == sorted(ln for ln in want if "ModelStat" in ln)
        assert [ln for ln in got if "ModelStat" not in ln] == [ln for ln in want if "ModelStat" not in ln]
    got = open(os.path.join(ws, "pj-top10.log")).read()
    assert got == kat["outputs"]["pj-top10.log"]


def _write_bmp8(path, img):
    """8-bit palettised bottom-up BMP with a grey palette (the Brown data set's format)."""
    h, w = img.shape
    stride = (w + 3) & ~3
    pal = b"".join(bytes([i, i, i, 0]) for i in range(256))
    off = 14 + 40 + len(pal)
    hdr = b"BM" + (off + stride * h).to_bytes(4, "little") + b"\0\0\0\0" + off.to_bytes(4, "little")
    info = (40).to_bytes(4, "little") + w.to_bytes(4, "little") + h.to_bytes(4, "little") + (1).to_bytes(2, "little") + \
        (8).to_bytes(2, "little") + (0).to_bytes(4, "little") + (stride * h).to_bytes(4, "little") + \
        (2835).to_bytes(4, "little") * 2 + (256).to_bytes(4, "little") + (0).to_bytes(4, "little")
    rows = b"".join(img[y].tobytes() + b"\0" * (stride - w) for y in range(h - 1, -1, -1))
    open(path, "wb").write(hdr + info + pal + rows)


def test_conv_impatches(tools_built, tmp_path):
    """src/conv-impatches.cpp: m50_500000_500000_0.txt -> Indices, patchesNNNN.bmp (16 x 16 patches each) -> Patches."""
    rng = np.random.default_rng(3)
    d = os.path.join(str(tmp_path), "liberty")
    os.makedirs(d)
    n_patches = 300                                                       # 2 bitmaps, the second partly used
    pats = rng.integers(0, 256, (512, 64, 64)).astype(np.uint8)
    for b in range(2):
        img = pats[256 * b:256 * b + 256].reshape(16, 16, 64, 64).transpose(0, 2, 1, 3).reshape(1024, 1024)
        _write_bmp8(os.path.join(d, "patches%04d.bmp" % b), img)
    idx = np.stack([rng.integers(0, n_patches, 700), rng.integers(0, 90, 700), rng.integers(0, n_patches, 700),
                    rng.integers(0, 90, 700)], 1).astype(np.int32)
    idx[17, 0] = n_patches - 1                                            # the largest id decides NumPatches
    with open(os.path.join(d, "m50_500000_500000_0.txt"), "w") as f:
        for r in idx:
            f.write("%d %d 0 %d %d 0\n" % tuple(r))
    out = os.path.join(str(tmp_path), "liberty.h5")
    r = subprocess.run([os.path.join(BIN, "conv-impatches"), d + "/", out], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr.decode()
    so = r.stdout.decode()
    assert so.startswith("Open index file: %s//m50_500000_500000_0.txt\nExport Indices: #700\n" % d)
    assert "Export Patches: #300\n" in so and so.endswith("Done.\n\n")
    f = H5File(out)
    assert np.array_equal(f.read("Indices"), idx)
    p = f.read("Patches")
    assert p.shape == (300, 64, 64) and p.dtype == np.uint8 and np.array_equal(p, pats[:300])
    assert f.info("Indices")["chunk"] == (256, 4) and f.info("Patches")["chunk"] == (256, 64, 64)
    r = subprocess.run([os.path.join(BIN, "conv-impatches"), str(tmp_path) + "/nowhere", out], stdout=subprocess.PIPE)
    assert r.returncode == 1 and b"ERROR: File m50_500000_500000_0.txt not found." in r.stdout
