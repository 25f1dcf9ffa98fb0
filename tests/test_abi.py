"""The C-ABI library builds for sm_100a without a GPU and exports every symbol include/qsb200.h declares."""
import ctypes
import os
import re

import pytest

from qsearch_b200 import build, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "qsb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(qsb_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_boundary():
    syms = declared_symbols()
    for must in ["qsb_init", "qsb_program_create", "qsb_eval_matrix", "qsb_eval_mat_jac", "qsb_eval_hs_grad",
                 "qsb_eval_residuals_jac", "qsb_solve_lm", "qsb_solve_lbfgs", "qsb_batch_run"]:
        assert must in syms


def test_cuda_library_builds_and_exports_every_symbol():
    path = build.build_cuda()
    dll = ctypes.CDLL(path)  # loading needs libcudart only, not a GPU
    for name in declared_symbols():
        assert hasattr(dll, name), name
    dll.qsb_version.restype = ctypes.c_char_p
    assert b"sm_100a" in dll.qsb_version()


def test_no_gpu_means_loud_failure_not_fallback():
    import numpy as np
    from qsearch_b200 import gates
    lib = _lib.Library(build.build_cuda())
    n = ctypes.c_int(0)
    rc = lib.dll.qsb_device_count(ctypes.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a GPU is visible here")
    from qsearch_b200.backends import CudaBackend
    prev = _lib.use_library(lib)
    try:
        circ = CudaBackend().prepare_circuit(gates.U3Gate())
        with pytest.raises(_lib.QsbError):
            circ.matrix(np.zeros(3))
    finally:
        _lib.use_library(prev)


def test_sass_is_sm100a_fp64():
    import subprocess
    path = build.build_cuda()
    out = subprocess.run(["cuobjdump", "-lelf", path], capture_output=True, text=True).stdout
    assert "sm_100a" in out
