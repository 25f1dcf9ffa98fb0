"""Times the UNMODIFIED reference (NumPy backend, its own SearchCompiler / LeapCompiler, its own
MultiprocessingParallelizer on all host cores) on the end-to-end synthesis configs and records
CNOT count, final distance and wall seconds per repetition.

Runs only in the build container, where /root/reference exists (it does not travel to the GPU box);
the output tests/golden/reference_synthesis.json is committed and read by tests/test_gpu_synthesis.py
and DESIGN.md §8.  Reference call sites exercised: qsearch/compiler.py:54-164 (SearchCompiler.compile),
qsearch/leap_compiler.py:78-195 (LeapCompiler.compile), README.md:20-35 (the table being reproduced),
examples/leap_qft4.py:7-15 (QFT4 options: min_depth=4).

Usage: python tests/golden/make_reference_synthesis.py [--reps3 10] [--reps4 3] [--only name,...]
Results are appended to the JSON after every repetition, so a partial run is still usable.
"""
import argparse
import json
import os
import platform
import sys
import time

sys.path.insert(0, "/root/reference")
import numpy as np  # noqa: E402
import qsearch  # noqa: E402
from qsearch import unitaries, gatesets, leap_compiler, options as qopt, utils, gates  # noqa: E402
from qsearch.compiler import SearchCompiler  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_synthesis.json")


def cnots(structure):
    # every search layer of the CNOT gatesets holds exactly one CNOT-type gate
    s = repr(structure)
    return s.count("CNOTGate(")


def cases3():
    t = {"qft3": unitaries.qft(8), "toffoli": unitaries.toffoli, "fredkin": unitaries.fredkin,
         "peres": unitaries.peres, "or": unitaries.logical_or}
    for topo, gs in (("linear", gatesets.QubitCNOTLinear), ("ring", gatesets.QubitCNOTRing)):
        for name, U in t.items():
            yield "%s_%s" % (name, topo), U, gs


def load():
    if os.path.exists(OUT):
        with open(OUT) as f:
            return json.load(f)
    cpu = "unknown"
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    cpu = line.split(":", 1)[1].strip()
                    break
    except OSError:
        pass
    return {"host": {"cpu": cpu, "cores": os.cpu_count(), "platform": platform.platform(),
                     "numpy": np.__version__, "qsearch": getattr(qsearch, "__version__", "?"),
                     "backend": "NumPy (qsrs unavailable: no Rust toolchain)",
                     "parallelizer": "MultiprocessingParallelizer, num_tasks = cores"},
            "runs": {}}


def save(db):
    tmp = OUT + ".tmp"
    with open(tmp, "w") as f:
        json.dump(db, f, indent=1, sort_keys=True)
    os.replace(tmp, OUT)


def one(compiler_cls, U, extra, seed):
    np.random.seed(seed)
    opts = qopt.Options(target=U, verbosity=0, **extra)
    comp = compiler_cls()
    t0 = time.time()
    res = comp.compile(opts)
    wall = time.time() - t0
    c, v = res["structure"], res["parameters"]
    d = float(utils.matrix_distance_squared(c.matrix(v), U))
    return {"seed": seed, "cnots": cnots(c), "distance": d, "wall_s": round(wall, 3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps3", type=int, default=10)
    ap.add_argument("--reps4", type=int, default=3)
    ap.add_argument("--repsq", type=int, default=3)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    only = set(x for x in a.only.split(",") if x)
    db = load()
    for name, U, gs in cases3():
        if only and name not in only:
            continue
        runs = db["runs"].setdefault(name, [])
        while len(runs) < a.reps3:
            r = one(SearchCompiler, U, {"gateset": gs()}, 1000 + len(runs))
            runs.append(r)
            save(db)
            print(name, r, flush=True)
    # two qutrits, the targets of examples/qutrit_csum.py with its QutritCNOTLinear gateset ("cnots" = two-qutrit gates)
    qt = {"qt_csum": gates.CSUMGate().matrix([]), "qt_bswap": utils.upgrade_qudits(unitaries.swap),
          "qt_tswap": unitaries.general_swap(3), "qt_qft2": unitaries.qft(9)}
    for name, U in qt.items():
        if name not in only:  # only on request: minutes to hours each
            continue
        runs = db["runs"].setdefault(name, [])
        while len(runs) < a.repsq:
            r = one(SearchCompiler, U, {"gateset": gatesets.QutritCNOTLinear()}, 3000 + len(runs))
            runs.append(r)
            save(db)
            print(name, r, flush=True)
    name = "qft4_leap"
    if not only or name in only:
        runs = db["runs"].setdefault(name, [])
        while len(runs) < a.reps4:
            r = one(leap_compiler.LeapCompiler, unitaries.qft(16), {"min_depth": 4}, 2000 + len(runs))
            runs.append(r)
            save(db)
            print(name, r, flush=True)


if __name__ == "__main__":
    main()
