// Batched on-device L-BFGS on the Hilbert-Schmidt objective f = 1 - |Tr(T^dagger U)|/N with the adjoint gradient
// (qsb_hs_instance).  Plays the role of BFGS_Jac_Solver (qsearch/solvers.py:138-144, scipy BFGS, jac=True,
// tol = options.threshold) and of the NLopt LD_LBFGS driver (qsrs/src/solvers.rs:33-74: memory 10, stopval 1e-16,
// maxeval 15000).  One CTA per instance; history, iterate and gradient live in shared memory.
#pragma once
#include "qsb_solver.cuh"

// nn = N^2, trig = QsbDims::TRIG of the geometry
static inline size_t qsb_lbfgs_smem_bytes(size_t nn, int trig, int max_ops, int max_params, int mem) {
  const size_t P = (size_t)max_params;
  return (2 * nn + 32 + QSB_HS_WPART + P) * 16 + (size_t)max_ops * trig * 8 + (5 * P + 2 * (size_t)mem * P + 2 * (size_t)mem) * 8 + 64;
}

QSB_DEVICE double qsb_dot(const double* u, const double* v, int P, cplx* red) {
  cplx part = cmk(0, 0);
  for (int k = threadIdx.x; k < P; k += blockDim.x) part.x = fma(u[k], v[k], part.x);
  return qsb_block_sum(part, red).x;
}

// f only (forward pass): 1 - |<T, U>| / N
template <int NQ>
QSB_DEVICE double qsb_hs_value_at(cplx* X, double* trig, const QsbDevProgram& P, const double* x, const double* target,
                                  cplx* red) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N;
  __syncthreads();
  qsb_fill_trig(P, x, trig, QsbDims<NQ>::TRIG);
  qsb_load_init<NQ>(X, nullptr, false);
  __syncthreads();
  qsb_forward<NQ>(X, P, trig);
  cplx part = cmk(0, 0);
  for (int e = threadIdx.x; e < NN; e += blockDim.x) {
    cplx t = cmk(target[2 * e], target[2 * e + 1]);
    part = cfma_conj(t, X[e], part);
  }
  const cplx tr = qsb_block_sum(part, red);
  return 1.0 - sqrt(tr.x * tr.x + tr.y * tr.y) / N;
}

template <int NQ>
QSB_GLOBAL __launch_bounds__(QsbGeom<NQ>::T) qsb_k_lbfgs(QsbSolveArgs a) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N;
  QSB_DYN_SMEM(smem_raw);
  cplx* X = (cplx*)smem_raw;
  cplx* L = X + NN;
  cplx* red = L + NN;
  cplx* wpart = red + 32;
  cplx* acc = wpart + QSB_HS_WPART;
  double* trig = (double*)(acc + a.max_params);
  double* x = trig + (size_t)a.max_ops * QsbDims<NQ>::TRIG;
  double* g = x + a.max_params;
  double* xn = g + a.max_params;
  double* gn = xn + a.max_params;
  double* dir = gn + a.max_params;
  const int mem = a.lbfgs_memory;
  double* Sh = dir + a.max_params;                 // [mem][maxP]
  double* Yh = Sh + (size_t)mem * a.max_params;    // [mem][maxP]
  double* rho = Yh + (size_t)mem * a.max_params;   // [mem]
  double* alpha = rho + mem;                       // [mem]
  (void)N;
  for (int inst = blockIdx.x; inst < a.count; inst += gridDim.x) {
    const QsbDevProgram prog = a.progs[a.prog_idx[inst]];
    const int P = prog.n_params, ld = a.max_params;
    const double* x0 = a.x0 + a.param_off[inst];
    const int max_evals = a.max_iters > 0 ? a.max_iters : 15000;
    const double stop_f = a.stop_f > 0.0 ? a.stop_f : 1e-16;
    __syncthreads();
    for (int k = threadIdx.x; k < P; k += blockDim.x) x[k] = x0[k];
    __syncthreads();
    int evals = 1, status = 0, hist = 0, head = 0, stall = 0;
    double f = qsb_hs_instance<NQ>(X, L, red, wpart, acc, trig, prog, x, a.target, g);
    while (P > 0) {
      if (f <= stop_f) { status = 4; break; }
      double gm = 0.0;
      for (int k = threadIdx.x; k < P; k += blockDim.x) gm = fmax(gm, fabs(g[k]));
      gm = qsb_block_max(gm, red);
      if (!(gm > a.gtol)) { status = 1; break; }
      if (evals >= max_evals) { status = 0; break; }
      // two-loop recursion: dir = -H g
      for (int k = threadIdx.x; k < P; k += blockDim.x) dir[k] = g[k];
      for (int i = 0; i < hist; ++i) {
        const int slot = (head - 1 - i + 2 * mem) % mem;
        const double al = rho[slot] * qsb_dot(Sh + (size_t)slot * ld, dir, P, red);
        if (threadIdx.x == 0) alpha[slot] = al;
        for (int k = threadIdx.x; k < P; k += blockDim.x) dir[k] = fma(-al, Yh[(size_t)slot * ld + k], dir[k]);
      }
      if (hist > 0) {
        const int slot = (head - 1 + mem) % mem;
        const double yy = qsb_dot(Yh + (size_t)slot * ld, Yh + (size_t)slot * ld, P, red);
        const double gamma = 1.0 / (rho[slot] * yy);
        for (int k = threadIdx.x; k < P; k += blockDim.x) dir[k] *= gamma;
      }
      for (int i = hist - 1; i >= 0; --i) {
        const int slot = (head - 1 - i + 2 * mem) % mem;
        const double be = rho[slot] * qsb_dot(Yh + (size_t)slot * ld, dir, P, red);
        const double al = alpha[slot];
        for (int k = threadIdx.x; k < P; k += blockDim.x) dir[k] = fma(al - be, Sh[(size_t)slot * ld + k], dir[k]);
      }
      double dg = -qsb_dot(g, dir, P, red);  // directional derivative along -dir
      double t = 1.0;
      if (!(dg < 0.0)) {  // not a descent direction: restart from steepest descent
        hist = 0;
        __syncthreads();
        for (int k = threadIdx.x; k < P; k += blockDim.x) dir[k] = g[k];
        dg = -qsb_dot(g, g, P, red);
      }
      if (hist == 0) t = fmin(1.0, 1.0 / sqrt(-dg));
      // backtracking Armijo line search on forward-only evaluations
      double fn = f;
      bool ok = false;
      for (int ls = 0; ls < 40; ++ls) {
        __syncthreads();
        for (int k = threadIdx.x; k < P; k += blockDim.x) xn[k] = fma(-t, dir[k], x[k]);
        fn = qsb_hs_value_at<NQ>(X, trig, prog, xn, a.target, red);
        ++evals;
        if (fn <= f + 1e-4 * t * dg) { ok = true; break; }
        // safeguarded quadratic interpolation
        const double tq = -dg * t * t / (2.0 * (fn - f - dg * t));
        t = (tq > 0.1 * t && tq < 0.5 * t) ? tq : 0.5 * t;
      }
      if (!ok) { status = 5; break; }
      const double fg = qsb_hs_instance<NQ>(X, L, red, wpart, acc, trig, prog, xn, a.target, gn);
      ++evals;
      // history update
      double* sv = Sh + (size_t)head * ld;
      double* yv = Yh + (size_t)head * ld;
      for (int k = threadIdx.x; k < P; k += blockDim.x) { sv[k] = xn[k] - x[k]; yv[k] = gn[k] - g[k]; }
      const double sy = qsb_dot(sv, yv, P, red);
      const double ss = qsb_dot(sv, sv, P, red), yy = qsb_dot(yv, yv, P, red);
      if (sy > 1e-10 * sqrt(ss * yy)) {
        if (threadIdx.x == 0) rho[head] = 1.0 / sy;
        head = (head + 1) % mem;
        if (hist < mem) ++hist;
      }
      stall = (f - fg <= a.ftol * fabs(fg)) ? stall + 1 : 0;
      __syncthreads();
      for (int k = threadIdx.x; k < P; k += blockDim.x) { x[k] = xn[k]; g[k] = gn[k]; }
      f = fg;
      __syncthreads();
      if (stall >= 3) { status = 2; break; }
    }
    __syncthreads();
    double* ox = a.out_x + a.param_off[inst];
    for (int k = threadIdx.x; k < P; k += blockDim.x) ox[k] = x[k];
    if (threadIdx.x == 0) {
      a.out_f[inst] = fabs(f);
      a.out_iters[inst] = evals;
      a.out_status[inst] = status;
    }
    __syncthreads();
  }
}
