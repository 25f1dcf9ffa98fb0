// Batched on-device L-BFGS on the register-resident HS kernels: one CTA of N/8 warps (one per column block) per
// (circuit, start) instance, so that no branch depends on the warp index (uniform control flow, see qsb_k_warp_jac).
// Same algorithm, settings and stopping rules as qsb_k_lbfgs (qsb_lbfgs.cuh: BFGS_Jac_Solver, qsearch/solvers.py:138-144;
// NLopt LD_LBFGS driver, qsrs/src/solvers.rs:33-74); the objective and gradient come from qsb_warp_hs_cols
// (N/8 warps, column blocks) and the line-search evaluations from the forward-only sweep qsb_warp_forward8.
#pragma once
#include "qsb_lbfgs.cuh"
#include "qsb_hs_warp.cuh"
#include "qsb_lm2.cuh"

struct QsbLbfgs2Smem {
  cplx* gtab; double* trig; double* trp; double* cpart; double* part;
};

// f and gradient at x (all threads of the CTA call; result uniform)
template <int NQ>
QSB_DEVICE double qsb_cta_hs_grad(const QsbDevProgram& P, const double* x, const double* target, const QsbLbfgs2Smem& sm,
                                  double* grad) {
  constexpr int NCB = (1 << NQ) / 8;
  const int lane = threadIdx.x & 31, warp = NCB == 1 ? 0 : (int)(threadIdx.x >> 5);
  __syncthreads();
  for (int p = threadIdx.x; p < P.n_params; p += blockDim.x) {
    double s, c;
    sincos(P.pscale[p] * x[p], &s, &c);
    sm.trig[2 * p] = c; sm.trig[2 * p + 1] = s;
  }
  __syncthreads();
  qsb_fill_gates(P, sm.trig, sm.gtab, threadIdx.x, blockDim.x);
  __syncthreads();
  qsb_warp_hs_cols<NQ>(P, target, sm.gtab, warp, NCB, lane, sm.trp + 2 * warp, sm.cpart);  // blockDim.x == 32 NCB
  __syncthreads();
  const double f = qsb_hs_finish<NQ>(P, sm.gtab, sm.trp, sm.cpart, NCB, threadIdx.x, blockDim.x, grad);
  __syncthreads();
  return f;
}

// f only: 1 - |Tr(U T^dagger)| / N  (= 1 - |<T, U>| / N)
template <int NQ>
QSB_DEVICE double qsb_cta_hs_value(const QsbDevProgram& P, const double* x, const double* target, const QsbLbfgs2Smem& sm) {
  cplx tr;
  qsb_cta_half_norm<NQ, true>(P, x, target, sm.trig, sm.gtab, sm.part, &tr);
  return 1.0 - sqrt(tr.x * tr.x + tr.y * tr.y) / (1 << NQ);
}

static inline size_t qsb_lbfgs2_smem_bytes(int nq, int max_ops, int max_params, int mem) {
  const size_t ncb = ((size_t)1 << nq) / 8, P = (size_t)max_params;
  return 32 * 16 + (8 * (size_t)(max_ops + max_params) + 2 * P + 2 * ncb + 8 * (size_t)max_ops * ncb + 16 + (size_t)3 * max_ops +
                    5 * P + 2 * (size_t)mem * P + 2 * (size_t)mem + 4) * 8 + 64;
}

template <int NQ>
QSB_GLOBAL __launch_bounds__(32 * ((1 << NQ) / 8)) qsb_k_lbfgs2(QsbSolveArgs a) {
  constexpr int NCB = (1 << NQ) / 8;
  QSB_DYN_SMEM(smem_raw);
  cplx* red = (cplx*)smem_raw;  // 32
  QsbLbfgs2Smem sm;
  sm.gtab = red + 32;
  sm.trig = (double*)(sm.gtab + (size_t)4 * (a.max_ops + a.max_params));
  sm.trp = sm.trig + 2 * a.max_params;
  sm.cpart = sm.trp + 2 * NCB;
  sm.part = sm.cpart + 8 * (size_t)a.max_ops * NCB;
  qsb_op* sops = (qsb_op*)(sm.part + 16);
  double* x = (double*)(sops + a.max_ops);
  double* g = x + a.max_params;
  double* xn = g + a.max_params;
  double* gn = xn + a.max_params;
  double* dir = gn + a.max_params;
  const int mem = a.lbfgs_memory;
  double* Sh = dir + a.max_params;                 // [mem][maxP]
  double* Yh = Sh + (size_t)mem * a.max_params;    // [mem][maxP]
  double* rho = Yh + (size_t)mem * a.max_params;   // [mem]
  double* alpha = rho + mem;                       // [mem]
  for (int inst = blockIdx.x; inst < a.count; inst += gridDim.x) {
    QsbDevProgram prog = a.progs[a.prog_idx[inst]];
    const int P = prog.n_params, ld = a.max_params;
    const double* x0 = a.x0 + a.param_off[inst];
    const int max_evals = a.max_iters > 0 ? a.max_iters : 15000;
    const double stop_f = a.stop_f > 0.0 ? a.stop_f : 1e-16;
    __syncthreads();
    for (int e = threadIdx.x; e < prog.n_ops * 6; e += blockDim.x) ((int32_t*)sops)[e] = ((const int32_t*)prog.ops)[e];
    prog.ops = sops;
    for (int k = threadIdx.x; k < P; k += blockDim.x) x[k] = x0[k];
    __syncthreads();
    int evals = 1, status = 0, hist = 0, head = 0, stall = 0;
    double f = qsb_cta_hs_grad<NQ>(prog, x, a.target, sm, g);
    while (P > 0) {
      if (f <= stop_f) { status = 4; break; }
      double gm = 0.0;
      for (int k = threadIdx.x; k < P; k += blockDim.x) gm = fmax(gm, fabs(g[k]));
      gm = qsb_block_max(gm, red);
      if (!(gm > a.gtol)) { status = 1; break; }
      if (evals >= max_evals) { status = 0; break; }
      for (int k = threadIdx.x; k < P; k += blockDim.x) dir[k] = g[k];  // two-loop recursion: dir = H g
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
      double fn = f;
      bool ok = false;
      for (int ls = 0; ls < 40; ++ls) {  // backtracking Armijo line search on forward-only evaluations
        __syncthreads();
        for (int k = threadIdx.x; k < P; k += blockDim.x) xn[k] = fma(-t, dir[k], x[k]);
        fn = qsb_cta_hs_value<NQ>(prog, xn, a.target, sm);
        ++evals;
        if (fn <= f + 1e-4 * t * dg) { ok = true; break; }
        const double tq = -dg * t * t / (2.0 * (fn - f - dg * t));  // safeguarded quadratic interpolation
        t = (tq > 0.1 * t && tq < 0.5 * t) ? tq : 0.5 * t;
      }
      if (!ok) { status = 5; break; }
      const double fg = qsb_cta_hs_grad<NQ>(prog, xn, a.target, sm, gn);
      ++evals;
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
