/* qsb200.h -- C ABI of libqsb200.so: the B200-native (sm_100a) qsearch hot path.
 *
 * This is the boundary the reference crosses with PyO3 in qsrs/src/lib.rs (native_from_object :183-277,
 * :543-554; PyGateWrapper.matrix / mat_jac :280-369; the native solver classes :391-539; the free functions
 * matrix_distance_squared(_jac) / matrix_residuals(_jac) :558-629).  Here it is a flat C ABI: plain pointers and
 * sizes, no torch / numpy types, loaded with ctypes (qsearch_b200/_lib.py).  INTEGRATION.md shows the binding a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - complex128 matrices are row-major, interleaved (re, im) doubles: exactly NumPy's complex128 layout.
 *   - qubit 0 is the most significant index bit (big endian, qsearch/unitaries.py:160).
 *   - every function returns QSB_OK (0) or a negative error code; qsb_last_error() gives the message of the
 *     last failure on the calling thread.  Nothing throws or aborts across the ABI.
 *   - all host buffers are caller-owned; the library owns device memory behind opaque handles.
 *   - a batch of "instances" = (program, parameter vector) pairs; parameters of a batch are one ragged
 *     concatenation params[param_offset[i] .. +P_i), in instance order; gradients/Jacobians use the same order.
 */
#ifndef QSB200_H
#define QSB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QSB_OK 0
#define QSB_ERR_INVALID (-1)     /* bad argument / malformed program */
#define QSB_ERR_CUDA (-2)        /* CUDA runtime failure (message has the CUDA error string) */
#define QSB_ERR_NOMEM (-3)
#define QSB_ERR_UNSUPPORTED (-4) /* e.g. more than QSB_MAX_QUBITS qubits */

#define QSB_MAX_QUBITS 5
#define QSB_MAX_QUTRITS 3 /* radix-3 programs (qsb_program_create_qudits): N = 3^n <= 27 */

/* Flat gate program: the lowered form of a ProductGate / KroneckerGate tree (qsearch/gates.py:845-1051).
 * Ops are applied in array order, each one LEFT-multiplying the running unitary (gates.py:959). */
enum qsb_opcode {
  QSB_OP_RX = 1,    /* XGate      gates.py:132  1 param  */
  QSB_OP_RY = 2,    /* YGate      gates.py:151  1 param  */
  QSB_OP_RZ = 3,    /* ZGate      gates.py:170  1 param  */
  QSB_OP_U1 = 4,    /* U1Gate     gates.py:387  1 param  */
  QSB_OP_U2 = 5,    /* U2Gate     gates.py:358  2 params */
  QSB_OP_U3 = 6,    /* U3Gate     gates.py:318  3 params */
  QSB_OP_XZXZ = 7,  /* XZXZGate   gates.py:270  2 params */
  QSB_OP_ZXZXZ = 8, /* ZXZXZGate  gates.py:208  3 params */
  QSB_OP_CONST1 = 9, /* constant 2x2 on qubit q0; const_off -> 8 doubles in the pool */
  QSB_OP_CNOT = 10,  /* CNOT(control q0, target q1): CNOTGate :604 / NonadjacentCNOTGate :696 (a row permutation) */
  QSB_OP_CONSTK = 11, /* constant 2^k x 2^k on contiguous qubits q0..q0+k-1 (k = q1); const_off -> 2*4^k doubles */
  /* two-qubit rotations exp(-i x/2 P (x) P) on qubits (q0, q1): the native backend's RXX / RYY / RZZ gates
   * (qsrs/src/circuits.rs:457-633; no Python twin in qsearch/gates.py).  1 param each. */
  QSB_OP_RXX = 12, /* circuits.rs:457-514 */
  QSB_OP_RYY = 13, /* circuits.rs:516-578 */
  QSB_OP_RZZ = 14, /* circuits.rs:580-633 */
  /* radix-3 programs only (qsb_program_create_qudits with radix 3): the arbitrary single-qutrit gate, 8 params
   * (three angles, five phases), on qutrit q0.  Constants of a radix-3 program (CSUM, CPI, CPIPhase, upgraded qubit
   * gates: gates.py:525-602, 761-788) are QSB_OP_CONSTK ops whose matrix is 3^k x 3^k. */
  QSB_OP_QUTRIT1 = 15 /* SingleQutritGate gates.py:411-523 */
};

typedef struct qsb_op {
  int32_t opcode;    /* enum qsb_opcode */
  int32_t q0;        /* qubit (1-qubit ops), control (CNOT), first qubit (CONSTK) */
  int32_t q1;        /* target (CNOT), k (CONSTK), else 0 */
  int32_t param_off; /* first parameter of this op within the circuit's parameter vector (param ops) */
  int32_t const_off; /* offset in doubles into the constant pool (CONST1 / CONSTK) */
  int32_t reserved;  /* ignored on input; the library stores the op's gate-table slot here in its own copy */
} qsb_op;

typedef struct qsb_ctx qsb_ctx;
typedef struct qsb_program qsb_program;
typedef struct qsb_batch qsb_batch;

/* what a batch run computes */
enum qsb_what {
  QSB_RUN_MATRIX = 1, /* U                       (Gate.matrix,  gates.py:946)                          */
  QSB_RUN_MATJAC = 2, /* U, J_k = dU/dx_k        (Gate.mat_jac, gates.py:965)                          */
  QSB_RUN_HS = 3,     /* f, grad f               (comparison.matrix_distance_squared_jac :37)           */
  QSB_RUN_RESID = 4,  /* r, J_r                  (comparison.matrix_residuals :57 / _residuals_jac :66) */
  QSB_RUN_DIST = 5,   /* f only                  (comparison.matrix_distance_squared :8)                */
  QSB_RUN_LMSYS = 6   /* r'r, J_r'J_r, J_r'r     (what a Levenberg-Marquardt step consumes; J_r stays in L2).
                         Device-resident batches only; one-qubit-leaf + CNOT programs on 3..5 qubits. */
};

/* ---- context ------------------------------------------------------------------------------------------ */
const char* qsb_last_error(void);
int qsb_device_count(int* out_count);
/* n_devices <= 0: use every visible device.  CUDA is first touched here (lazy; importing the shim is fork-safe). */
int qsb_init(int n_devices, const int* device_ids, qsb_ctx** out_ctx);
int qsb_shutdown(qsb_ctx* ctx);
int qsb_ctx_num_devices(const qsb_ctx* ctx);
/* Page-locked host memory for the buffers of the host-buffer calls (optional: any host pointer is accepted, pinned
 * ones let the device->host copy run at PCIe speed and overlap with the next chunk's kernel). */
int qsb_host_alloc(qsb_ctx* ctx, size_t bytes, void** out_ptr);
int qsb_host_free(qsb_ctx* ctx, void* ptr);
const char* qsb_version(void);

/* ---- programs (replaces native_from_object, qsrs/src/lib.rs:543-554) ------------------------------------ */
int qsb_program_create(qsb_ctx* ctx, int n_qubits, int n_ops, const qsb_op* ops, int n_params,
                       const double* const_pool, size_t const_len, qsb_program** out_prog);
/* The same for qudits of size `radix` (2: identical to qsb_program_create; 3: qutrits, gates.py:411-602 -- the reference's
 * native backend half-supports them, qsrs/src/circuits.rs:968-1100, its Python backend fully).  A radix-3 program has
 * n_qudits <= QSB_MAX_QUTRITS, ops QSB_OP_QUTRIT1 and QSB_OP_CONSTK only (q0 / q1 count qutrits; a CONSTK constant is
 * 3^k x 3^k, 2*9^k doubles), N = 3^n_qudits everywhere an output size says N, and runs in the interpreter kernels: every
 * evaluation call and both solvers, not QSB_RUN_LMSYS. */
int qsb_program_create_qudits(qsb_ctx* ctx, int radix, int n_qudits, int n_ops, const qsb_op* ops, int n_params,
                              const double* const_pool, size_t const_len, qsb_program** out_prog);
int qsb_program_destroy(qsb_ctx* ctx, qsb_program* prog);
int qsb_program_radix(const qsb_program* prog);
int qsb_program_num_params(const qsb_program* prog);
int qsb_program_num_qubits(const qsb_program* prog);

/* ---- host-buffer evaluation calls (replace PyGateWrapper.matrix / mat_jac and the objective functions) ---
 * progs[n_programs]; prog_of_instance[n_instances] indexes progs; all programs must have the same qubit count (and radix).
 * params: ragged concatenation, instance i uses the next P_i doubles.  Work is sharded over the context's
 * devices in contiguous instance ranges; host<->device copies are part of the call. */
int qsb_eval_matrix(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                    const int32_t* prog_of_instance, const double* params,
                    double* out_U /* [inst][N][N][2] */);
int qsb_eval_mat_jac(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                     const int32_t* prog_of_instance, const double* params,
                     double* out_U /* [inst][N][N][2] */, double* out_J /* [sum P][N][N][2] */);
int qsb_eval_hs_grad(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                     const int32_t* prog_of_instance, const double* params, const double* target /* [N][N][2] */,
                     double* out_f /* [inst] */, double* out_grad /* [sum P] */);
int qsb_eval_distance(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                      const int32_t* prog_of_instance, const double* params, const double* target,
                      double* out_f /* [inst] */);
/* out_Jr is [sum P][2 N^2]: row k = concat(ravel(Re K_k), ravel(Im K_k)), K_k = J_k T^dagger.  That is the memory
 * the reference builds before its final .T (comparison.py:66-72); the shim returns the transposed view. */
int qsb_eval_residuals_jac(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                           const int32_t* prog_of_instance, const double* params, const double* target,
                           double* out_r /* [inst][2 N^2] */, double* out_Jr /* [sum P][2 N^2] */);

/* ---- device-resident batches (what bench.py times with inputs already in HBM; also used by the solvers) -- */
int qsb_batch_create(qsb_ctx* ctx, int device_slot, int n_programs, const qsb_program* const* progs,
                     int n_instances, const int32_t* prog_of_instance, qsb_batch** out_batch);
int qsb_batch_destroy(qsb_batch* batch);
/* Kernel family of a batch: QSB_KERNEL_AUTO picks the register-resident DMMA warp kernels where they apply (one-qubit
 * leaves + CNOT, 3..5 qubits) and the shared-memory interpreter otherwise; QSB_KERNEL_GENERIC forces the interpreter
 * (parity tests and profiling compare the two).  QSB_KERNEL=generic in the environment sets the default of a context. */
#define QSB_KERNEL_AUTO 0
#define QSB_KERNEL_GENERIC 1
int qsb_batch_set_kernel(qsb_batch* batch, int kind);
int qsb_batch_set_params(qsb_batch* batch, const double* params, size_t n_doubles); /* H2D */
int qsb_batch_set_target(qsb_batch* batch, const double* target);                   /* H2D, [N][N][2] */
/* Launch on the batch's stream over instances [first, first+count); outputs land in device buffers sized for
 * `count` instances.  If elapsed_ms != NULL the launch is bracketed by CUDA events on that stream and the call
 * waits for completion; otherwise it returns after enqueueing. */
int qsb_batch_run(qsb_batch* batch, int what, int first, int count, float* elapsed_ms);
int qsb_batch_sync(qsb_batch* batch);
/* Copy the outputs of the last run of `what` to host.  Unused pointers may be NULL.
 *   MATRIX: a=U            MATJAC: a=U, b=J       HS: a=f, b=grad      DIST: a=f
 *   RESID:  a=r, b=J_r     LMSYS:  a=r'r [inst], b=J'r [sum P], c=J'J [sum P_i^2] */
int qsb_batch_fetch(qsb_batch* batch, int what, double* a, double* b, double* c);
/* bytes of device output the last run of `what` wrote (for bandwidth accounting) */
int qsb_batch_output_bytes(qsb_batch* batch, int what, int count, uint64_t* out_bytes);
/* number of kernel launches issued by this context so far */
int qsb_ctx_launch_count(const qsb_ctx* ctx, uint64_t* out);

/* ---- batched on-device solvers (replace BFGS_Jac_SolverNative / LeastSquares_Jac_SolverNative,
 *      qsrs/src/lib.rs:391-539, qsrs/src/solvers.rs:13-122; scipy drivers qsearch/solvers.py:138-163) -------- */
typedef struct qsb_solver_opts {
  int32_t max_iters;   /* LM: max Jacobian evaluations; <=0 -> 100*P like max_nfev (scipy) / max_num_iterations (Ceres) */
  int32_t lbfgs_memory; /* L-BFGS history length, default 10 (qsrs/src/lib.rs:402-413) */
  double ftol, xtol, gtol; /* <=0 -> 1e-8 defaults of scipy.least_squares(method='lm') */
  double stop_value;   /* stop as soon as the objective (HS distance for L-BFGS, r'r/(2N) proxy for LM) < this; <=0 off */
  /* Batch-level stop rules of the LM solver (all off at 0, which reproduces the reference's per-solve semantics:
   * every start runs to its own convergence, qsearch/solvers.py:146-163).  A frontier expansion waits for its slowest
   * instance, so the search front-end turns them on (qsearch_b200/parallelizers.py). */
  double straggler_factor;    /* > 0: once half of the batch has terminated (for any reason, the sibling rule included), stop
                                 an instance whose Jacobian-evaluation count exceeds straggler_factor x the largest count among
                                 the instances of that first half that converged by themselves (out_status 6) */
  double sibling_stop;        /* > 0: stop an instance once another instance of the SAME program finished with a distance
                                 below this value (out_status 7): the other starts of a solved child are not needed */
  double straggler_min_iters; /* the straggler rule never fires before this many evaluations (<= 0: 32) */
  double gram_tf32;           /* > 0 (body-frame LM kernel only): form J_r'J_r, which only steers the optimiser, from TF32-rounded rows on
                                 the TF32 tensor path; J_r'r, the objective and every returned value stay fp64.  0: everything fp64 */
} qsb_solver_opts;

/* One independent solve per instance, parameters stay on the device for the whole solve.
 * x0 / out_x: ragged like params.  out_f: final HS distance 1-|Tr(T^dagger U)|/N per instance.
 * out_iters: iterations used.  out_status: >0 converged reason, 0 iteration limit. */
int qsb_solve_lm(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                 const int32_t* prog_of_instance, const double* x0, const double* target,
                 const qsb_solver_opts* opts, double* out_x, double* out_f, int32_t* out_iters,
                 int32_t* out_status);
int qsb_solve_lbfgs(qsb_ctx* ctx, int n_programs, const qsb_program* const* progs, int n_instances,
                    const int32_t* prog_of_instance, const double* x0, const double* target,
                    const qsb_solver_opts* opts, double* out_x, double* out_f, int32_t* out_iters,
                    int32_t* out_status);

/* ---- allocation-free variant of QSB_RUN_LMSYS --------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Computes, for n_instances parameter vectors of ONE program, what qsb_batch_run(QSB_RUN_LMSYS) + qsb_batch_fetch return:
 *   out_rr [n] = r'r,   out_g [n][P] = J_r'r,   out_A [n][P][P] = J_r'J_r (row-major).
 * params [n][P], target [N][N][2] (must be unitary) and the outputs are caller-owned and may be HOST or DEVICE memory
 * (copied with cudaMemcpyDefault).  The call allocates nothing on the device: its whole working set is one static
 * __device__ array of QSB_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until the process exits.
 * Limits (QSB_ERR_UNSUPPORTED with a message otherwise): one-qubit-leaf + CNOT programs on 3..5 qubits; staged inputs +
 * outputs <= 16 MiB, i.e. about 8 (P^2 + 2 P + 1) n bytes (n = 4, P = 62: ~500 instances); per-instance scratch <= 48 MiB.
 * Every call shares the array and it keeps the previous call's contents (the kernel zeroes what it accumulates into).
 * Calls on one device are serialised inside this library and the call returns after the stream has drained; work that
 * reaches the same GPU from another process or library copy while a call runs would race on the array. */
#define QSB_STATIC_SCRATCH_BYTES ((size_t)64 << 20)
int qsb_eval_lm_system_static(qsb_ctx* ctx, int device_slot, const qsb_program* prog, int n_instances,
                              const double* params, const double* target, double* out_rr, double* out_g, double* out_A);

/* ---- device micro-benchmarks (FP64 pipe / DMMA peak measured on the box; bench.py records them) ---------- */
int qsb_measure_fp64_peak(qsb_ctx* ctx, int device_slot, int use_dmma, double* out_tflops);
/* store-only HBM bandwidth in GB/s over a 4 GiB buffer (mode 0: 16-byte coalesced stores from a kernel; 1: cudaMemsetAsync):
 * the roof of a kernel that, like the LM-form evaluation, writes its output and reads almost nothing */
int qsb_measure_hbm_write(qsb_ctx* ctx, int device_slot, int mode, double* out_gbs);

#ifdef __cplusplus
}
#endif
#endif /* QSB200_H */
