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
