// Batched on-device Levenberg-Marquardt: one CTA per (circuit, start vector) instance, the whole solve inside one
// kernel launch, parameters never leave the device.  Plays the role of LeastSquares_Jac_Solver
// (qsearch/solvers.py:146-163: scipy least_squares(method='lm') = MINPACK lmder) and of the Ceres DENSE_QR driver
// (qsrs/src/solvers.rs:88-121).  Trajectory parity with those third-party optimisers is unpinned;
// the end state (distance below threshold) is what the reference's tests check.
//
// Per iteration:  r, J_r  (Jacobian sweep, qsb_jac_instance)  ->  A = J_r^T J_r, g = J_r^T r  (tiled SYRK)
//                 (A + mu D^2) h = -g  (Cholesky; D = running max of column norms, MINPACK's scaling)
//                 gain ratio on a forward-only trial evaluation; Nielsen damping update.
#pragma once
#include "qsb_generic.cuh"

#define QSB_SYRK_MC 32

// device-side control block of one solve launch (ints, zeroed by the host before the launch)
#define QSB_CTL_NEXT 0      // next instance of the queue
#define QSB_CTL_FINISHED 1  // instances finished so far (any status)
#define QSB_CTL_MEDIAN 2    // max Jacobian-evaluation count among the naturally finished instances of the first half to finish
#define QSB_CTL_SOLVED 4    // + program index: 1 once an instance of that program finished below the sibling threshold

struct QsbSolveArgs {
  const QsbDevProgram* progs;
  const int32_t* prog_idx;
  const long long* param_off;
  const double* x0;
  const double* target;
  int count, max_ops, max_params;
  int max_iters;  // <= 0: 100 * P (LM) / 15000 evaluations (L-BFGS)
  int lbfgs_memory;
  double ftol, xtol, gtol, stop_f;  // stop_f: stop once F = |r|^2 / 2 <= stop_f (<= 0: off)
  double* scratch;           // per CTA: R [(maxP+1)][2NN], A [(maxP+1)^2], L [maxP^2]
  size_t scratch_per_cta;    // doubles
  double* out_x;
  double* out_f;
  int32_t* out_iters;
  int32_t* out_status;
  int cluster;      // qsb_k_lm2: CTAs per instance (thread-block cluster size; 1 in the interpreter kernels)
  int chol_smem;    // qsb_k_lm2: packed Cholesky factor in shared memory (fits) or in global scratch
  long long* prof;  // optional [12] cycle counters (QSB_LM_PROFILE=1): jac, syrk, cholesky, solve, trial, iterations
  // qsb_k_lm3: device-side queue and batch-level stop rules (see QSB_CTL_* in qsb_lm3.cuh)
  int* ctl;             // [4 + n_programs] ints, zeroed before the launch
  double straggler_c;   // > 0: once half of the batch has finished, stop an instance whose Jacobian-evaluation count
                        // exceeds straggler_c x the largest count among that first half (status 6)
  int straggler_min;    // ... but never before this many evaluations
  double sibling_f;     // > 0: stop an instance once another instance of the same program finished below this distance (status 7)
  int gram_tf32;        // qsb_k_lm3: J_r^T J_r from TF32-rounded rows on the mma.sync TF32 path (the gradient and the objective stay fp64)
  // Damping schedule of the LM kernels.  Accepted step: mu *= max(mu_down, 1 - (2 rho - 1)^3), with no decrease at all for the first
  // mu_streak accepted steps after a rejection; rejected step: mu *= nu, nu *= mu_up.  (1/3, 0, 2) is Nielsen's rule, the default:
  // on these objectives almost every accepted step is then followed by a rejected one (1.7-2.0 factorisations + trial evaluations
  // per Jacobian evaluation), but the instances that crawl along a valley shed their damping fastest, and without a cap the wall
  // time of a batch is its slowest instance.  Batches that run under the straggler cap (frontier expansions) use (0.7, 1, 2):
  // 1.2-1.25 attempts per Jacobian evaluation at the same number of Jacobian evaluations, 11-18 % less work per solve and
  // 15 % less time per iteration, at the price of a longer tail that the cap cuts anyway (profiles/r2/zd_mu_sweep.log,
  // zg_bench_*.json: +17 % solves/s under the frontier rules at n = 4, -19 % without them).  mu_fast > 0: from the mu_fast-th
  // consecutive accepted step on the floor is 1/3 again (tried, no consistent gain: zf_mu_sweep.log).
  // QSB_LM_MU_DOWN / _UP / _STREAK / _FAST in the environment override (experiments).
  double mu_down = 1.0 / 3.0, mu_up = 2.0;
  int mu_streak = 0, mu_fast = 0;
};

QSB_DEVICE double qsb_block_max(double v, cplx* red) {
  for (int m = 16; m > 0; m >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, m));
  int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w].x = v;
  __syncthreads();
  double s = red[0].x;
  for (int i = 1; i < nw; ++i) s = fmax(s, red[i].x);
  return s;
}

// F = |U T^dagger - I|_F^2 / 2 at x (forward only).  Leaves X = U T^dagger.
template <int NQ>
QSB_DEVICE double qsb_half_norm_at(cplx* X, double* trig, const QsbDevProgram& P, const double* x, const double* target,
                                   cplx* red) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N;
  __syncthreads();
  qsb_fill_trig(P, x, trig, QsbDims<NQ>::TRIG);
  qsb_load_init<NQ>(X, target, true);
  __syncthreads();
  qsb_forward<NQ>(X, P, trig);
  cplx part = cmk(0, 0);
  for (int e = threadIdx.x; e < NN; e += blockDim.x) {
    double re = X[e].x - ((e / N == e % N) ? 1.0 : 0.0), im = X[e].y;
    part.x = fma(re, re, part.x);
    part.x = fma(im, im, part.x);
  }
  return 0.5 * qsb_block_sum(part, red).x;
}

// A[k][l] = sum_m R[k][m] R[l][m], 0 <= l <= k < rows (both triangles written), R row-major [rows][M] in global.
// tile: shared [4*ceil(rows/4)][QSB_SYRK_MC + 1].
QSB_DEVICE void qsb_syrk(const double* R, int rows, int M, double* A, double* tile) {
  const int nb = (rows + 3) >> 2, ntiles = nb * (nb + 1) / 2, ldt = QSB_SYRK_MC + 1, T = blockDim.x;
  for (int base = 0; base < ntiles; base += T) {
    const int t = base + threadIdx.x;
    const bool valid = t < ntiles;
    int bk = 0, bl = 0;
    if (valid) {
      bk = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
      while ((bk + 1) * (bk + 2) / 2 <= t) ++bk;
      while (bk * (bk + 1) / 2 > t) --bk;
      bl = t - bk * (bk + 1) / 2;
    }
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
    for (int m0 = 0; m0 < M; m0 += QSB_SYRK_MC) {
      __syncthreads();
      for (int idx = threadIdx.x; idx < 4 * nb * QSB_SYRK_MC; idx += T) {
        int r = idx / QSB_SYRK_MC, mm = idx % QSB_SYRK_MC;
        tile[r * ldt + mm] = (r < rows && m0 + mm < M) ? R[(size_t)r * M + m0 + mm] : 0.0;
      }
      __syncthreads();
      if (valid) {
        const double* ta = tile + (4 * bk) * ldt;
        const double* tb = tile + (4 * bl) * ldt;
#pragma unroll 4
        for (int mm = 0; mm < QSB_SYRK_MC; ++mm) {
          double av[4], bv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) { av[i] = ta[i * ldt + mm]; bv[i] = tb[i * ldt + mm]; }
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
        }
      }
    }
    if (valid) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          int k = 4 * bk + i, l = 4 * bl + j;
          if (k < rows && l < rows) { A[k * rows + l] = acc[i][j]; A[l * rows + k] = acc[i][j]; }
        }
    }
  }
  __syncthreads();
}

// Cholesky of A[0:P,0:P] + mu diag(d2) -> L (row-major, ld = P, lower).  Uniform return value.
QSB_DEVICE bool qsb_cholesky(const double* A, int ldA, const double* d2, double mu, int P, double* L, double* col) {
  for (int j = 0; j < P; ++j) {
    for (int i = j + threadIdx.x; i < P; i += blockDim.x) {
      double s = A[i * ldA + j] + (i == j ? mu * d2[j] : 0.0);
      const double* li = L + (size_t)i * P;
      const double* lj = L + (size_t)j * P;
      for (int k = 0; k < j; ++k) s = fma(-li[k], lj[k], s);
      col[i] = s;
    }
    __syncthreads();
    const double piv = col[j];
    if (!(piv > 0.0) || !(piv < 1e300)) { __syncthreads(); return false; }
    const double inv = 1.0 / sqrt(piv);
    for (int i = j + threadIdx.x; i < P; i += blockDim.x) L[(size_t)i * P + j] = col[i] * inv;
    __syncthreads();
  }
  return true;
}

// Solve L L^T h = b.  b is overwritten (work), y is scratch, result in h.  All in shared memory except L.
QSB_DEVICE void qsb_chol_solve(const double* L, int P, double* b, double* y, double* h) {
  for (int j = 0; j < P; ++j) {
    __syncthreads();
    const double yj = b[j] / L[(size_t)j * P + j];
    if (threadIdx.x == 0) y[j] = yj;
    for (int i = j + 1 + threadIdx.x; i < P; i += blockDim.x) b[i] = fma(-L[(size_t)i * P + j], yj, b[i]);
  }
  for (int j = P - 1; j >= 0; --j) {
    __syncthreads();
    const double hj = y[j] / L[(size_t)j * P + j];
    if (threadIdx.x == 0) h[j] = hj;
    for (int i = threadIdx.x; i < j; i += blockDim.x) y[i] = fma(-L[(size_t)j * P + i], hj, y[i]);
  }
  __syncthreads();
}

// shared memory of qsb_k_lm (bytes) and its scratch per CTA (doubles); nn = N^2, trig = QsbDims::TRIG
static inline size_t qsb_lm_smem_bytes(size_t nn, int trig, int max_ops, size_t maxP) {
  const size_t nb = (maxP + 1 + 3) / 4;
  return (3 * nn + 32) * 16 + 7 * maxP * 8 + (size_t)max_ops * trig * 8 + 4 * nb * (QSB_SYRK_MC + 1) * 8 + 64;
}
static inline size_t qsb_lm_scratch_doubles(size_t nn, size_t maxP) {
  return (maxP + 1) * 2 * nn + (maxP + 1) * (maxP + 1) + maxP * maxP;
}

template <int NQ>
QSB_GLOBAL __launch_bounds__(QsbGeom<NQ>::T) qsb_k_lm(QsbSolveArgs a) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N, M = 2 * NN;
  QSB_DYN_SMEM(smem_raw);
  cplx* X = (cplx*)smem_raw;
  cplx* S = X + NN;
  cplx* Y = S + NN;
  cplx* red = Y + NN;  // 32
  double* x = (double*)(red + 32);
  double* xn = x + a.max_params;
  double* g = xn + a.max_params;
  double* h = g + a.max_params;
  double* d2 = h + a.max_params;
  double* wk = d2 + a.max_params;   // work vector (rhs / column buffer)
  double* wy = wk + a.max_params;   // work vector
  double* trig = wy + a.max_params;
  double* tile = trig + (size_t)a.max_ops * QsbDims<NQ>::TRIG;
  double* Rg = a.scratch + (size_t)blockIdx.x * a.scratch_per_cta;
  for (int inst = blockIdx.x; inst < a.count; inst += gridDim.x) {
    const QsbDevProgram prog = a.progs[a.prog_idx[inst]];
    const int P = prog.n_params, rows = P + 1;
    double* A = Rg + (size_t)(a.max_params + 1) * M;
    double* L = A + (size_t)(a.max_params + 1) * (a.max_params + 1);
    const double* x0 = a.x0 + a.param_off[inst];
    const int max_iters = a.max_iters > 0 ? a.max_iters : 100 * (P > 0 ? P : 1);
    __syncthreads();
    for (int k = threadIdx.x; k < P; k += blockDim.x) x[k] = x0[k];
    __syncthreads();
    int iters = 0, status = 0;
    double F = 0.0;
    if (P > 0) {
      // r -> row P of R, J_r -> rows 0..P-1
      qsb_jac_instance<NQ, QSB_RUN_RESID>(X, S, Y, trig, prog, x, a.target, Rg + (size_t)P * M, Rg);
      iters = 1;
      qsb_syrk(Rg, rows, M, A, tile);
      for (int k = threadIdx.x; k < P; k += blockDim.x) {
        g[k] = A[P * rows + k];
        double akk = A[k * rows + k];
        d2[k] = akk > 0.0 ? akk : 1.0;
      }
      F = 0.5 * A[P * rows + P];
      __syncthreads();
      double mu = 1e-3, nu = a.mu_up;
      int streak = 0;
      int rejects = 0;
      while (true) {
        if (!(F > 1e-300) || (a.stop_f > 0.0 && F <= a.stop_f)) { status = 4; break; }
        {  // MINPACK's gtol: max cosine between r and the columns of J_r
          double gm = 0.0;
          const double rn = sqrt(2.0 * F);
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            double akk = A[k * rows + k];
            if (akk > 0.0) gm = fmax(gm, fabs(g[k]) / (sqrt(akk) * rn));
          }
          gm = qsb_block_max(gm, red);
          if (gm <= a.gtol) { status = 1; break; }
        }
        if (!qsb_cholesky(A, rows, d2, mu, P, L, wk)) {
          mu *= nu; nu *= a.mu_up; streak = 0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
          continue;
        }
        for (int k = threadIdx.x; k < P; k += blockDim.x) wk[k] = -g[k];
        qsb_chol_solve(L, P, wk, wy, h);
        cplx part = cmk(0, 0), part2 = cmk(0, 0);
        for (int k = threadIdx.x; k < P; k += blockDim.x) {
          xn[k] = x[k] + h[k];
          part.x = fma(d2[k] * h[k], h[k], part.x);          // |D h|^2
          part.y = fma(d2[k] * x[k], x[k], part.y);          // |D x|^2
          part2.x = fma(h[k], mu * d2[k] * h[k] - g[k], part2.x);  // 2 * predicted reduction
        }
        part = qsb_block_sum(part, red);
        part2 = qsb_block_sum(part2, red);
        const double hnorm = sqrt(part.x), xnorm = sqrt(part.y), pred = 0.5 * part2.x;
        const double Fn = qsb_half_norm_at<NQ>(X, trig, prog, xn, a.target, red);
        const double dF = F - Fn;
        const double rho = (pred > 0.0) ? dF / pred : -1.0;
        if (rho > 1e-4 && Fn == Fn) {
          rejects = 0;
          __syncthreads();
          for (int k = threadIdx.x; k < P; k += blockDim.x) x[k] = xn[k];
          __syncthreads();
          const bool conv_f = (dF <= a.ftol * F) && (pred <= a.ftol * F);
          const bool conv_x = hnorm <= a.xtol * xnorm;
          F = Fn;
          if (a.stop_f > 0.0 && F <= a.stop_f) { status = 4; break; }
          if (conv_f) { status = 2; break; }
          if (conv_x) { status = 3; break; }
          if (iters >= max_iters) { status = 0; break; }
          qsb_jac_instance<NQ, QSB_RUN_RESID>(X, S, Y, trig, prog, x, a.target, Rg + (size_t)P * M, Rg);
          ++iters;
          qsb_syrk(Rg, rows, M, A, tile);
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            g[k] = A[P * rows + k];
            d2[k] = fmax(d2[k], A[k * rows + k]);
          }
          F = 0.5 * A[P * rows + P];
          __syncthreads();
          const double q = 2.0 * rho - 1.0;
          double fac = fmax((a.mu_fast > 0 && streak + 1 >= a.mu_fast) ? 1.0 / 3.0 : a.mu_down, 1.0 - q * q * q);
          if (++streak <= a.mu_streak) fac = fmax(fac, 1.0);  // damping schedule: see QsbSolveArgs
          mu *= fac;
          nu = a.mu_up;
        } else {
          if (hnorm <= a.xtol * xnorm) { status = 3; break; }
          mu *= nu; nu *= a.mu_up; streak = 0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
        }
      }
    }
    // final distance 1 - |Tr(U T^dagger)| / N at x   (comparison.matrix_distance_squared, comparison.py:8-22)
    qsb_half_norm_at<NQ>(X, trig, prog, x, a.target, red);
    cplx dg = cmk(0, 0);
    if (threadIdx.x < N) dg = X[threadIdx.x * N + threadIdx.x];
    const cplx tr = qsb_block_sum(dg, red);
    double* ox = a.out_x + a.param_off[inst];
    for (int k = threadIdx.x; k < P; k += blockDim.x) ox[k] = x[k];
    if (threadIdx.x == 0) {
      a.out_f[inst] = fabs(1.0 - sqrt(tr.x * tr.x + tr.y * tr.y) / N);
      a.out_iters[inst] = iters;
      a.out_status[inst] = status;
    }
    __syncthreads();
  }
}
