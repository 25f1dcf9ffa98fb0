"""ctypes binding of libqsb200.so (include/qsb200.h).  No PyTorch, no CPU fallback.

The library is located next to this file (built in-tree by ``__graft_entry__.build()`` /
``python -m qsearch_b200.build``).  Nothing CUDA-related happens at import: the shared object is loaded and the
context created on first use, so the package can be imported before a fork (hard part 4).
If the library is missing, or no CUDA device is visible, every entry point raises -- there is no other path.
"""
import ctypes
import time
import os
import threading

import numpy as np

from .lowering import OP_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_NAME = "libqsb200.so"

RUN_MATRIX, RUN_MATJAC, RUN_HS, RUN_RESID, RUN_DIST, RUN_LMSYS = 1, 2, 3, 4, 5, 6

_c_dp = ctypes.POINTER(ctypes.c_double)
_c_ip = ctypes.POINTER(ctypes.c_int32)
_vp = ctypes.c_void_p


class QsbError(RuntimeError):
    pass


class SolverOpts(ctypes.Structure):
    _fields_ = [("max_iters", ctypes.c_int32), ("lbfgs_memory", ctypes.c_int32), ("ftol", ctypes.c_double),
                ("xtol", ctypes.c_double), ("gtol", ctypes.c_double), ("stop_value", ctypes.c_double),
                ("straggler_factor", ctypes.c_double), ("sibling_stop", ctypes.c_double),
                ("straggler_min_iters", ctypes.c_double), ("gram_tf32", ctypes.c_double)]


def _dp(a):
    return None if a is None else a.ctypes.data_as(_c_dp)


def _f64(a):
    return np.ascontiguousarray(a, dtype="float64")


class Library:
    """One loaded copy of the C ABI + one context over the chosen devices."""

    def __init__(self, path=None, devices=None):
        path = path or os.path.join(_HERE, LIB_NAME)
        if not os.path.exists(path):
            raise QsbError("{} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(qsearch_b200 has no CPU fallback)".format(path))
        self.path = path
        self.dll = ctypes.CDLL(path)
        self._declare()
        self._ctx = None
        self._devices = devices
        self._lock = threading.Lock()
        self._plock = threading.RLock()
        self._programs = {}
        self._pid = None

    # ------------------------------------------------------------------ plumbing
    def _declare(self):
        d = self.dll
        d.qsb_last_error.restype = ctypes.c_char_p
        d.qsb_version.restype = ctypes.c_char_p
        pp = ctypes.POINTER(_vp)
        d.qsb_init.argtypes = [ctypes.c_int, _c_ip, pp]
        d.qsb_shutdown.argtypes = [_vp]
        d.qsb_device_count.argtypes = [ctypes.POINTER(ctypes.c_int)]
        d.qsb_ctx_num_devices.argtypes = [_vp]
        d.qsb_host_alloc.argtypes = [_vp, ctypes.c_size_t, pp]
        d.qsb_host_free.argtypes = [_vp, _vp]
        d.qsb_ctx_launch_count.argtypes = [_vp, ctypes.POINTER(ctypes.c_uint64)]
        d.qsb_program_create.argtypes = [_vp, ctypes.c_int, ctypes.c_int, _vp, ctypes.c_int, _c_dp, ctypes.c_size_t, pp]
        d.qsb_program_create_qudits.argtypes = [_vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, _vp, ctypes.c_int, _c_dp, ctypes.c_size_t, pp]
        d.qsb_program_radix.argtypes = [_vp]
        d.qsb_program_destroy.argtypes = [_vp, _vp]
        common = [_vp, ctypes.c_int, pp, ctypes.c_int, _c_ip, _c_dp]
        d.qsb_eval_matrix.argtypes = common + [_c_dp]
        d.qsb_eval_mat_jac.argtypes = common + [_c_dp, _c_dp]
        d.qsb_eval_hs_grad.argtypes = common + [_c_dp, _c_dp, _c_dp]
        d.qsb_eval_distance.argtypes = common + [_c_dp, _c_dp]
        d.qsb_eval_residuals_jac.argtypes = common + [_c_dp, _c_dp, _c_dp]
        d.qsb_batch_create.argtypes = [_vp, ctypes.c_int, ctypes.c_int, pp, ctypes.c_int, _c_ip, pp]
        d.qsb_batch_destroy.argtypes = [_vp]
        d.qsb_batch_set_params.argtypes = [_vp, _c_dp, ctypes.c_size_t]
        d.qsb_batch_set_kernel.argtypes = [_vp, ctypes.c_int]
        d.qsb_batch_set_target.argtypes = [_vp, _c_dp]
        d.qsb_batch_run.argtypes = [_vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
        d.qsb_batch_sync.argtypes = [_vp]
        d.qsb_batch_fetch.argtypes = [_vp, ctypes.c_int, _c_dp, _c_dp, _c_dp]
        d.qsb_batch_output_bytes.argtypes = [_vp, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)]
        solve = common + [_c_dp, ctypes.POINTER(SolverOpts), _c_dp, _c_dp, _c_ip, _c_ip]
        d.qsb_solve_lm.argtypes = solve
        d.qsb_solve_lbfgs.argtypes = solve
        d.qsb_measure_fp64_peak.argtypes = [_vp, ctypes.c_int, ctypes.c_int, _c_dp]
        d.qsb_measure_hbm_write.argtypes = [_vp, ctypes.c_int, ctypes.c_int, _c_dp]
        d.qsb_eval_lm_system_static.argtypes = [_vp, ctypes.c_int, _vp, ctypes.c_int, _c_dp, _c_dp, _c_dp, _c_dp, _c_dp]

    def check(self, rc):
        if rc != 0:
            raise QsbError("qsb200 error {}: {}".format(rc, self.dll.qsb_last_error().decode()))

    @property
    def ctx(self):
        pid = os.getpid()
        if self._ctx is None or self._pid != pid:
            with self._lock:
                if self._ctx is None or self._pid != pid:  # a forked child gets its own context
                    ctx = _vp()
                    if self._devices is None:
                        self.check(self.dll.qsb_init(0, None, ctypes.byref(ctx)))
                    else:
                        ids = (ctypes.c_int32 * len(self._devices))(*self._devices)
                        self.check(self.dll.qsb_init(len(self._devices), ids, ctypes.byref(ctx)))
                    self._ctx, self._pid = ctx, pid
                    self._programs = {}  # key -> [handle, GateProgram, reference count], least recently used first
        return self._ctx

    @property
    def num_devices(self):
        return self.dll.qsb_ctx_num_devices(self.ctx)

    def pinned_empty(self, shape, dtype="float64"):
        """NumPy array over page-locked host memory (qsb_host_alloc); freed when the array is collected."""
        import weakref
        dt = np.dtype(dtype)
        n = int(np.prod(shape)) * dt.itemsize
        p = _vp()
        self.check(self.dll.qsb_host_alloc(self.ctx, n, ctypes.byref(p)))
        buf = (ctypes.c_char * max(1, n)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dt, count=int(np.prod(shape))).reshape(shape)
        weakref.finalize(buf, self.dll.qsb_host_free, self._ctx, p)
        return arr

    def version(self):
        return self.dll.qsb_version().decode()

    def launch_count(self):
        n = ctypes.c_uint64()
        self.check(self.dll.qsb_ctx_launch_count(self.ctx, ctypes.byref(n)))
        return n.value

    # ------------------------------------------------------------------ programs
    PROGRAM_CACHE_LIMIT = 4096

    def _acquire(self, gps):
        """Device handles of GatePrograms (uploaded once per context, cached by identity key), each with one more
        reference.  Entries are only ever destroyed when nobody references them: a live Batch and every call in flight
        hold references, so eviction can never free a program a kernel may still read (the cache is an LRU bounded at
        PROGRAM_CACHE_LIMIT idle entries + whatever is referenced)."""
        ctx = self.ctx
        with self._plock:
            keys = [gp.key if gp.key is not None else id(gp) for gp in gps]
            if len(self._programs) + len(keys) > self.PROGRAM_CACHE_LIMIT:
                need = set(keys)
                for k in list(self._programs):
                    if len(self._programs) <= self.PROGRAM_CACHE_LIMIT // 2:
                        break
                    ent = self._programs[k]
                    if ent[2] == 0 and k not in need:
                        self.dll.qsb_program_destroy(ctx, ent[0])
                        del self._programs[k]
            handles = []
            for gp, key in zip(gps, keys):
                ent = self._programs.get(key)
                if ent is None:
                    ops = np.ascontiguousarray(gp.ops, dtype=OP_DTYPE)
                    h = _vp()
                    if getattr(gp, "radix", 2) == 2:
                        self.check(self.dll.qsb_program_create(ctx, gp.n_qubits, len(ops), ops.ctypes.data_as(_vp), gp.n_params,
                                                               _dp(gp.consts), gp.consts.size, ctypes.byref(h)))
                    else:  # qutrit programs (radix 3)
                        self.check(self.dll.qsb_program_create_qudits(ctx, gp.radix, gp.n_qubits, len(ops), ops.ctypes.data_as(_vp),
                                                                      gp.n_params, _dp(gp.consts), gp.consts.size, ctypes.byref(h)))
                    ent = [h, gp, 0]  # keep gp alive so an id()-based key cannot be reused
                else:
                    del self._programs[key]  # re-insert below: most recently used last
                self._programs[key] = ent
                ent[2] += 1
                handles.append(ent[0])
        return keys, handles

    def _release(self, keys):
        with self._plock:
            for k in keys:
                ent = self._programs.get(k)
                if ent is not None and ent[2] > 0:
                    ent[2] -= 1

    def program(self, gp):
        """Device handle for one GateProgram (no reference kept: use _acquire / _release around launches)."""
        keys, handles = self._acquire([gp])
        self._release(keys)
        return handles[0]

    class _Packed:
        """Arguments of one host-buffer call; holds program references until the call returns."""

        def __init__(self, lib, gps, prog_of_instance, params):
            self.lib = lib
            self.keys, hs = lib._acquire(gps)
            self.handles = (_vp * len(gps))(*hs)
            self.poi = np.ascontiguousarray(prog_of_instance, dtype="int32")
            self.params = _f64(params).ravel()
            self.sizes = np.array([g.n_params for g in gps], dtype="int64")[self.poi]
            if self.params.size != int(self.sizes.sum()):
                lib._release(self.keys)
                raise ValueError("expected {} parameters in total, got {}".format(int(self.sizes.sum()), self.params.size))

        def __enter__(self):
            return self.handles, self.poi, self.params, self.sizes

        def __exit__(self, *exc):
            self.lib._release(self.keys)
            return False

    def _pack(self, gps, prog_of_instance, params):
        return Library._Packed(self, gps, prog_of_instance, params)

    # ------------------------------------------------------------------ host-buffer evaluation
    def eval_matrix(self, gps, prog_of_instance, params):
        with self._pack(gps, prog_of_instance, params) as (handles, poi, params, _):
            N = gps[0].dim
            U = np.empty((len(poi), N, N), dtype="complex128")
            self.check(self.dll.qsb_eval_matrix(self.ctx, len(gps), handles, len(poi), poi.ctypes.data_as(_c_ip), _dp(params),
                                                U.ctypes.data_as(_c_dp)))
        return U

    def eval_mat_jac(self, gps, prog_of_instance, params):
        with self._pack(gps, prog_of_instance, params) as (handles, poi, params, sizes):
            N = gps[0].dim
            U = np.empty((len(poi), N, N), dtype="complex128")
            J = np.empty((int(sizes.sum()), N, N), dtype="complex128")
            self.check(self.dll.qsb_eval_mat_jac(self.ctx, len(gps), handles, len(poi), poi.ctypes.data_as(_c_ip), _dp(params),
                                                 U.ctypes.data_as(_c_dp), J.ctypes.data_as(_c_dp)))
        return U, J

    def eval_hs_grad(self, gps, prog_of_instance, params, target):
        with self._pack(gps, prog_of_instance, params) as (handles, poi, params, sizes):
            T = np.ascontiguousarray(target, dtype="complex128")
            f = np.empty(len(poi), dtype="float64")
            g = np.empty(int(sizes.sum()), dtype="float64")
            self.check(self.dll.qsb_eval_hs_grad(self.ctx, len(gps), handles, len(poi), poi.ctypes.data_as(_c_ip), _dp(params),
                                                 T.ctypes.data_as(_c_dp), _dp(f), _dp(g)))
        return f, g

    def eval_distance(self, gps, prog_of_instance, params, target):
        with self._pack(gps, prog_of_instance, params) as (handles, poi, params, _):
            T = np.ascontiguousarray(target, dtype="complex128")
            f = np.empty(len(poi), dtype="float64")
            self.check(self.dll.qsb_eval_distance(self.ctx, len(gps), handles, len(poi), poi.ctypes.data_as(_c_ip), _dp(params),
                                                  T.ctypes.data_as(_c_dp), _dp(f)))
        return f

    def eval_residuals_jac(self, gps, prog_of_instance, params, target, out_r=None, out_Jr=None):
        """r: [inst, 2N^2]; Jr: [sum P, 2N^2] (row k = parameter k; the reference's array is its transpose view)."""
        with self._pack(gps, prog_of_instance, params) as (handles, poi, params, sizes):
            N = gps[0].dim
            T = np.ascontiguousarray(target, dtype="complex128")
            r = out_r if out_r is not None else np.empty((len(poi), 2 * N * N), dtype="float64")
            Jr = out_Jr if out_Jr is not None else np.empty((int(sizes.sum()), 2 * N * N), dtype="float64")
            self.check(self.dll.qsb_eval_residuals_jac(self.ctx, len(gps), handles, len(poi), poi.ctypes.data_as(_c_ip),
                                                       _dp(params), T.ctypes.data_as(_c_dp), _dp(r), _dp(Jr)))
        return r, Jr

    # ------------------------------------------------------------------ solvers
    def solve(self, method, gps, prog_of_instance, x0, target, max_iters=0, ftol=0.0, xtol=0.0, gtol=0.0,
              stop_value=0.0, lbfgs_memory=10, straggler_factor=0.0, sibling_stop=0.0, straggler_min_iters=0.0,
              gram_tf32=0.0):
        with self._pack(gps, prog_of_instance, x0) as (handles, poi, x0, sizes):
            T = np.ascontiguousarray(target, dtype="complex128")
            opts = SolverOpts(max_iters=max_iters, lbfgs_memory=lbfgs_memory, ftol=ftol, xtol=xtol, gtol=gtol,
                              stop_value=stop_value, straggler_factor=straggler_factor, sibling_stop=sibling_stop,
                              straggler_min_iters=straggler_min_iters, gram_tf32=gram_tf32)
            x = np.empty_like(x0)
            f = np.empty(len(poi), dtype="float64")
            iters = np.empty(len(poi), dtype="int32")
            status = np.empty(len(poi), dtype="int32")
            fn = self.dll.qsb_solve_lm if method == "lm" else self.dll.qsb_solve_lbfgs
            t0 = time.perf_counter()
            self.check(fn(self.ctx, len(gps), handles, len(poi), poi.ctypes.data_as(_c_ip), _dp(x0), T.ctypes.data_as(_c_dp),
                          ctypes.byref(opts), _dp(x), _dp(f), iters.ctypes.data_as(_c_ip), status.ctypes.data_as(_c_ip)))
            self.last_solve_call_ms = 1e3 * (time.perf_counter() - t0)  # the C call alone (QSB_TRACE_SOLVES prints it)
        return x, f, iters, status

    def eval_lm_system_static(self, gp, params, target, slot=0):
        """qsb_eval_lm_system_static: (r'r [n], J_r'r [n, P], J_r'J_r [n, P, P]) of n parameter vectors of ONE program, computed
        without any device allocation (static scratch array, include/qsb200.h)."""
        keys, hs = self._acquire([gp])
        try:
            V = np.ascontiguousarray(np.atleast_2d(_f64(params)))
            n, P = V.shape
            if P != gp.n_params:
                raise ValueError("expected {} parameters per vector, got {}".format(gp.n_params, P))
            T = np.ascontiguousarray(target, dtype="complex128")
            rr, g, A = np.empty(n), np.empty((n, P)), np.empty((n, P, P))
            self.check(self.dll.qsb_eval_lm_system_static(self.ctx, slot, hs[0], n, _dp(V), T.ctypes.data_as(_c_dp), _dp(rr), _dp(g), _dp(A)))
        finally:
            self._release(keys)
        return rr, g, A

    def measure_hbm_write(self, slot=0, mode=0):
        out = ctypes.c_double()
        self.check(self.dll.qsb_measure_hbm_write(self.ctx, slot, int(mode), ctypes.byref(out)))
        return out.value

    def measure_fp64_peak(self, slot=0, use_dmma=False):
        out = ctypes.c_double()
        self.check(self.dll.qsb_measure_fp64_peak(self.ctx, slot, int(use_dmma), ctypes.byref(out)))
        return out.value

    # ------------------------------------------------------------------ device-resident batches
    def batch(self, gps, prog_of_instance, slot=0):
        return Batch(self, gps, prog_of_instance, slot)


class Batch:
    """Device-resident batch (qsb_batch_*): parameters and outputs stay in HBM between calls."""

    def __init__(self, lib, gps, prog_of_instance, slot=0):
        self.lib = lib
        self.gps = list(gps)
        self.poi = np.ascontiguousarray(prog_of_instance, dtype="int32")
        self.sizes = np.array([g.n_params for g in self.gps], dtype="int64")[self.poi]
        self.n_qubits = self.gps[0].n_qubits
        self.h = None
        self._keys, hs = lib._acquire(self.gps)  # the batch keeps its programs referenced until close()
        handles = (_vp * len(self.gps))(*hs)
        h = _vp()
        try:
            lib.check(lib.dll.qsb_batch_create(lib.ctx, slot, len(self.gps), handles, len(self.poi),
                                               self.poi.ctypes.data_as(_c_ip), ctypes.byref(h)))
        except Exception:
            lib._release(self._keys)
            self._keys = []
            raise
        self.h = h

    def set_kernel(self, kind):
        """0 = automatic (DMMA warp kernels where they apply), 1 = shared-memory interpreter."""
        self.lib.check(self.lib.dll.qsb_batch_set_kernel(self.h, int(kind)))

    def set_params(self, params):
        params = _f64(params).ravel()
        self.lib.check(self.lib.dll.qsb_batch_set_params(self.h, _dp(params), params.size))

    def set_target(self, target):
        T = np.ascontiguousarray(target, dtype="complex128")
        self.lib.check(self.lib.dll.qsb_batch_set_target(self.h, T.ctypes.data_as(_c_dp)))

    def run(self, what, first=0, count=None, timed=False):
        count = len(self.poi) - first if count is None else count
        ms = ctypes.c_float()
        self.lib.check(self.lib.dll.qsb_batch_run(self.h, what, first, count, ctypes.byref(ms) if timed else None))
        self._last = (what, first, count)
        return ms.value if timed else None

    def sync(self):
        self.lib.check(self.lib.dll.qsb_batch_sync(self.h))

    def output_bytes(self, what, count):
        n = ctypes.c_uint64()
        self.lib.check(self.lib.dll.qsb_batch_output_bytes(self.h, what, count, ctypes.byref(n)))
        return n.value

    def fetch(self):
        what, first, count = self._last
        N = self.gps[0].dim
        npar = int(self.sizes[first:first + count].sum())
        a = b = c = None
        if what == RUN_MATRIX:
            a = np.empty((count, N, N), dtype="complex128")
        elif what == RUN_MATJAC:
            a, b = np.empty((count, N, N), dtype="complex128"), np.empty((npar, N, N), dtype="complex128")
        elif what == RUN_RESID:
            a, b = np.empty((count, 2 * N * N)), np.empty((npar, 2 * N * N))
        elif what == RUN_HS:
            a, b = np.empty(count), np.empty(npar)
        elif what == RUN_DIST:
            a = np.empty(count)
        elif what == RUN_LMSYS:
            a, b = np.empty(count), np.empty(npar)
            c = np.empty(int((self.sizes[first:first + count] ** 2).sum()))
        ptr = [None if x is None else x.ctypes.data_as(_c_dp) for x in (a, b, c)]
        self.lib.check(self.lib.dll.qsb_batch_fetch(self.h, what, *ptr))
        return a, b, c

    def close(self):
        if self.h is not None and self.h.value:
            self.lib.dll.qsb_batch_destroy(self.h)
            self.h = None
            self.lib._release(self._keys)
            self._keys = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_LIB = None
_LIB_LOCK = threading.Lock()


def get_library():
    """The process-wide Library over the in-tree libqsb200.so (raises if it is missing)."""
    global _LIB
    if _LIB is None:
        with _LIB_LOCK:
            if _LIB is None:
                devs = os.environ.get("QSB200_DEVICES")
                _LIB = Library(devices=[int(x) for x in devs.split(",")] if devs else None)
    return _LIB


def use_library(lib):
    """Select an explicitly constructed Library (e.g. one context per rank; the CPU test-suite passes the
    host-emulation build of the same kernels from tests/emu here).  Returns the previous one."""
    global _LIB
    prev, _LIB = _LIB, lib
    return prev
