import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


def _oracle_lib():
    path = os.path.join(ROOT, "oracle", "libdlco_oracle.so")
    src = os.path.join(ROOT, "oracle", "dlco_oracle.c")
    if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    lib = ctypes.CDLL(path)
    lib.dlco_oracle_hinge_pairsum.restype = ctypes.c_double
    lib.dlco_oracle_hinge_pairsum.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                              ctypes.c_int, ctypes.c_double]
    return lib


@pytest.fixture(scope="session")
def oracle_lib():
    return _oracle_lib()


def make_bank(lib, rings):
    """(filters nF x 4096, PRParams nF x 3) for the given RingParams rows, via the C oracle."""
    rings = np.ascontiguousarray(rings, np.float64)
    n = len(rings)
    filt = np.zeros((n * 8, 4096), np.float32)
    prp = np.zeros((n * 8, 3), np.float32)
    lib.dlco_oracle_pr_bank(64, n, rings.ctypes.data_as(ctypes.c_void_p),
                            filt.ctypes.data_as(ctypes.c_void_p), prp.ctypes.data_as(ctypes.c_void_p))
    return filt, prp


@pytest.fixture(scope="session")
def full_bank(oracle_lib):
    from oracle import dlco_oracle as O
    rings = O.ring_params()
    filt, prp = make_bank(oracle_lib, rings)
    return rings, filt, prp


@pytest.fixture(scope="session")
def synth():
    from oracle import dlco_oracle as O
    patches, pid = O.synth_patches(24, per_point=3, seed=7)
    idx = O.synth_pairs(pid, 40, seed=8)
    return patches, pid, idx


@pytest.fixture(scope="session")
def gpu_ctx():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import __graft_entry__
    __graft_entry__.build()            # no-op when libdlco_b200.so is newer than its sources
    import dlco_b200
    ctx = dlco_b200.Context(0)
    yield ctx
    ctx.close()
