#!/usr/bin/env python
"""Run the emulator suites against a sanitizer build of the kernel sources (TEST INFRASTRUCTURE; CPU only):

    python tests/emu/run_sanitized.py address     [pytest args]      # heap / dynamic-shared-memory / scratch overruns
    python tests/emu/run_sanitized.py undefined   [pytest args]      # misaligned vector loads, shifts, signed overflow

The script re-executes itself with the sanitizer runtime preloaded (the interpreter is not instrumented).  Default selection:
tests/test_emu_parity.py tests/test_emu_solvers.py.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "address"
    args = sys.argv[2:] or [os.path.join(ROOT, "tests", "test_emu_parity.py"), os.path.join(ROOT, "tests", "test_emu_solvers.py")]
    from qsearch_b200 import build
    lib = build.build_emu(sanitize=kind)
    if os.environ.get("QSB_SANITIZED_CHILD") != kind:
        rt = subprocess.run(["gcc", "-print-file-name=lib{}.so".format("asan" if kind == "address" else "ubsan")],
                            capture_output=True, text=True).stdout.strip()
        env = dict(os.environ, LD_PRELOAD=rt, QSB_SANITIZED_CHILD=kind,
                   ASAN_OPTIONS="detect_leaks=0:detect_stack_use_after_return=0:halt_on_error=1",
                   UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1")
        sys.exit(subprocess.run([sys.executable, os.path.abspath(__file__), kind] + args, env=env).returncode)
    build.build_emu = lambda force=False, sanitize=None: lib  # tests.emu.emulated_library() picks this build up
    import pytest
    sys.exit(pytest.main(["-x", "-q", "-p", "no:cacheprovider", "-m", "not gpu"] + args))


if __name__ == "__main__":
    main()
