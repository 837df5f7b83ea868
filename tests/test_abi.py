"""The C-ABI shared library loads without a GPU and exports every symbol include/dlco_b200.h
declares; the Python binding table mirrors the header one-to-one; compute entry points fail
loudly (no CPU fallback) when no CUDA device is present."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__
    __graft_entry__.build()
    import dlco_b200
    return dlco_b200.load_library()


def header_functions():
    txt = open(os.path.join(ROOT, "include", "dlco_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(dlco_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_the_documented_entry_points():
    names = header_functions()
    for must in ("dlco_init", "dlco_fulldist_build", "dlco_fulldist_build_host", "dlco_uprjdist_build",
                 "dlco_hinge_pairsum", "dlco_pr_rda_run", "dlco_pr_validate", "dlco_pj_rda_grad",
                 "dlco_pj_rda_update", "dlco_pj_prox_apply", "dlco_pj_validate", "dlco_pr_patchdist",
                 "dlco_pj_patchdist", "dlco_roc_stats", "dlco_get_desc"):
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    import dlco_b200
    assert dlco_b200.MISSING == []
    for name in header_functions():
        assert hasattr(lib, name), "libdlco_b200.so does not export %s" % name


def test_binding_table_matches_header():
    import dlco_b200
    assert sorted(dlco_b200.SIGNATURES) == header_functions()


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = ctypes.c_void_p()
    rc = lib.dlco_init(0, ctypes.byref(h))
    assert rc == -2 and not h.value                      # DLCO_ENODEV
    assert b"no CPU fallback" in lib.dlco_last_error()
    import dlco_b200
    with pytest.raises(dlco_b200.DlcoError):
        dlco_b200.Context(0)


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "opencv-dlco_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "dlco_oracle" not in src, f


def test_tools_fail_loudly_without_a_gpu(tmp_path):
    """The C++ tools over the ABI have no CPU path either: on a machine without a CUDA device each one exits non-zero,
    says why on stderr and leaves no output file behind.  (On a GPU box this test has nothing to check.)"""
    import subprocess
    import numpy as np
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import __graft_entry__
    __graft_entry__.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    bin_ = os.path.join(root, "bin")
    # a small valid unprojected-distance file and a pr model for the tools that read one
    dist = str(tmp_path / "yosemite-m-unproj.h5")
    x = np.random.default_rng(0).random((64, 32), dtype=np.float32)
    lab = (np.arange(64) % 2).astype(np.uint8).reshape(-1, 1)
    for name, arr, dt in (("Label", lab, "u8"), ("Distance", x, "f32")):
        arr.tofile(str(tmp_path / (name + ".bin")))
    subprocess.check_call([os.path.join(bin_, "h5tool"), "pack", dist,
                           "Label", "u8", "2", "64", "1", str(tmp_path / "Label.bin"), "64", "1",
                           "Distance", "f32", "2", "64", "32", str(tmp_path / "Distance.bin"), "64", "0"])
    os.makedirs(str(tmp_path / "out" / "logging"))
    runs = [
        ([os.path.join(bin_, "gen-poolregion"), str(tmp_path / "filters.h5")], str(tmp_path / "filters.h5")),
        ([os.path.join(bin_, "pj-learn"), "-iters", "10", dist, str(tmp_path / "pj.h5")], str(tmp_path / "pj.h5")),
        ([os.path.join(bin_, "pj-sweep"), dist, str(tmp_path / "out"), "-iters", "10", "-mu", "0.001", "-gamma", "0.1"],
         str(tmp_path / "out" / "yosemite-m-0.001-0.1-pj.h5")),
    ]
    for cmd, product in runs:
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=120)
        assert r.returncode != 0, cmd
        assert b"no CUDA device" in r.stderr and b"no CPU fallback" in r.stderr, (cmd, r.stderr)
        assert not os.path.exists(product), product
