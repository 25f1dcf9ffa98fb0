"""Build libqsb200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m qsearch_b200.build            # product library  -> qsearch_b200/libqsb200.so
    python -m qsearch_b200.build --emu      # tests only: host emulation of the same sources -> tests/emu/_build/

The library is several translation units compiled in parallel (the solver kernels dominate the compile time) and then
linked; objects are cached under qsearch_b200/_obj/ (git-ignored) keyed on the modification times of the sources.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libqsb200.so")
OBJ_DIR = os.path.join(HERE, "_obj")
EMU_DIR = os.path.join(ROOT, "tests", "emu", "_build")
EMU_LIB = os.path.join(EMU_DIR, "libqsb200_emu.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-O2"]

# (object name, source, extra defines)
UNITS = [
    ("api", "qsb_api.cu", []),
    ("lmsys", "qsb_lmsys.cu", []),
    ("solve_lm", "qsb_solve.cu", ["-DQSB_SOLVE_TU=0"]),
    ("solve_lbfgs", "qsb_solve.cu", ["-DQSB_SOLVE_TU=1"]),
    ("solve_lm5", "qsb_solve.cu", ["-DQSB_SOLVE_TU=2"]),
    ("solve_lm5t", "qsb_solve.cu", ["-DQSB_SOLVE_TU=3"]),
    ("solve_lm4w", "qsb_solve.cu", ["-DQSB_SOLVE_TU=4"]),
    ("peak", "qsb_peak.cu", []),
    ("noalloc", "qsb_noalloc.cu", []),
]


def _unit_deps(src):
    """Files a translation unit depends on: the transitive closure of its quoted #include lines (+ this script)."""
    seen, todo = set(), [os.path.join(CSRC, src)]
    while todo:
        f = os.path.normpath(todo.pop())
        if f in seen or not os.path.exists(f):
            continue
        seen.add(f)
        with open(f) as fh:
            for line in fh:
                line = line.strip()
                if line.startswith('#include "'):
                    todo.append(os.path.join(os.path.dirname(f), line.split('"')[1]))
    return sorted(seen) + [os.path.abspath(__file__)]


def _sources():
    out = [os.path.join(ROOT, "include", "qsb200.h"), os.path.abspath(__file__)]
    for f in sorted(os.listdir(CSRC)):
        out.append(os.path.join(CSRC, f))
    return out


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return r.returncode, r.stdout


def _compile_all(jobs, verbose):
    """jobs: list of (object path, command).  Compiles the stale ones in parallel; raises on the first failure."""
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as ex:
        results = list(ex.map(lambda j: _run(j[1]), jobs))
    for (obj, cmd), (rc, out) in zip(jobs, results):
        if verbose or rc:
            sys.stdout.write(out)
        if rc:
            raise subprocess.CalledProcessError(rc, cmd)


def build_cuda(force=False, verbose=False, only=None):
    deps = _sources()
    if not force and not _stale(LIB, deps):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ_DIR, exist_ok=True)
    jobs, objs = [], []
    for name, src, defs in UNITS:
        obj = os.path.join(OBJ_DIR, name + ".o")
        objs.append(obj)
        if force or _stale(obj, _unit_deps(src)) or (only and name in only):
            jobs.append((obj, [nvcc] + NVCC_FLAGS + defs + (["-Xptxas", "-v"] if verbose else []) +
                         ["-c", "-o", obj, os.path.join(CSRC, src)]))
    _compile_all(jobs, verbose)
    subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB] + objs, check=True)
    return LIB


def build_emu(force=False, sanitize=None):
    """TEST INFRASTRUCTURE: compile the same .cu sources for the host with tests/emu/cuda_emu.h.
    sanitize = "address" | "undefined": an instrumented copy (own directory) for tests/emu/run_sanitized.py."""
    hdr = os.path.join(ROOT, "tests", "emu", "cuda_emu.h")
    deps = _sources() + [hdr]
    out_dir = EMU_DIR if not sanitize else os.path.join(EMU_DIR, sanitize)
    lib = EMU_LIB if not sanitize else os.path.join(out_dir, "libqsb200_emu_{}.so".format(sanitize))
    if not force and not _stale(lib, deps):
        return lib
    os.makedirs(out_dir, exist_ok=True)
    opt = ["-O2"] if not sanitize else ["-O1", "-g", "-fno-omit-frame-pointer", "-fsanitize=" + sanitize, "-fno-sanitize=vptr"]
    jobs, objs = [], []
    for name, src, defs in UNITS:
        obj = os.path.join(out_dir, name + ".o")
        objs.append(obj)
        if force or _stale(obj, _unit_deps(src) + [hdr]):
            jobs.append((obj, ["g++", "-x", "c++", "-std=c++20"] + opt + ["-fPIC", "-pthread", "-DQSB_EMU"] +
                         (["-DQSB_EMU_DEFINE"] if name == "api" else []) + defs +
                         ["-include", hdr, "-c", "-o", obj, os.path.join(CSRC, src)]))
    _compile_all(jobs, False)
    subprocess.run(["g++", "-shared", "-pthread"] + (["-fsanitize=" + sanitize] if sanitize else []) + ["-o", lib] + objs, check=True)
    return lib


if __name__ == "__main__":
    if "--emu" in sys.argv:
        san = sys.argv[sys.argv.index("--sanitize") + 1] if "--sanitize" in sys.argv else None
        print(build_emu(force="--force" in sys.argv, sanitize=san))
    else:
        print(build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv))
