"""Build libqsb200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m qsearch_b200.build            # product library  -> qsearch_b200/libqsb200.so
    python -m qsearch_b200.build --emu      # tests only: host emulation of the same sources -> tests/emu/_build/
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libqsb200.so")
EMU_DIR = os.path.join(ROOT, "tests", "emu", "_build")
EMU_LIB = os.path.join(EMU_DIR, "libqsb200_emu.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-O2", "-shared", "--threads", "0"]


def _sources():
    out = [os.path.join(ROOT, "include", "qsb200.h")]
    for f in sorted(os.listdir(CSRC)):
        out.append(os.path.join(CSRC, f))
    return out


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_cuda(force=False, verbose=False):
    if not force and not _stale(LIB, _sources()):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB, os.path.join(CSRC, "qsb_api.cu")]
    subprocess.run(cmd, check=True)
    return LIB


def build_exp(verbose=False):
    """DEVELOPER TOOL: evaluation kernels only (no solver kernels), for quick A/B runs of a kernel change through
    QSB_LIB_PATH (tools/prof_case.py).  Never loaded by the product."""
    out = os.path.join(HERE, "libqsb200_exp.so")
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-DQSB_EXPERIMENT"] + (["-Xptxas", "-v"] if verbose else []) + ["-o", out, os.path.join(CSRC, "qsb_api.cu")]
    subprocess.run(cmd, check=True)
    return out


def build_emu(force=False):
    """TEST INFRASTRUCTURE: compile the same .cu sources for the host with tests/emu/cuda_emu.h."""
    hdr = os.path.join(ROOT, "tests", "emu", "cuda_emu.h")
    if not force and not _stale(EMU_LIB, _sources() + [hdr]):
        return EMU_LIB
    os.makedirs(EMU_DIR, exist_ok=True)
    cmd = ["g++", "-x", "c++", "-std=c++20", "-O2", "-fPIC", "-shared", "-pthread", "-DQSB_EMU", "-DQSB_EMU_DEFINE",
           "-include", hdr, "-o", EMU_LIB, os.path.join(CSRC, "qsb_api.cu")]
    subprocess.run(cmd, check=True)
    return EMU_LIB


if __name__ == "__main__":
    if "--exp" in sys.argv:
        print(build_exp(verbose="-v" in sys.argv))
    elif "--emu" in sys.argv:
        print(build_emu(force="--force" in sys.argv))
    else:
        print(build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv))
