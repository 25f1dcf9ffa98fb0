// Batched on-device Levenberg-Marquardt, tensor-core version: one CTA of QSB_LM_WARPS warps per (circuit, start
// vector) instance, the whole solve in one launch.  Same role and stopping rules as qsb_k_lm (qsb_solver.cuh):
// LeastSquares_Jac_Solver (qsearch/solvers.py:146-163, scipy least_squares(method='lm') = MINPACK lmder) and the
// Ceres DENSE_QR driver (qsrs/src/solvers.rs:88-121).  What changes is how an iteration is computed:
//
//   r, J_r     the warps split the program into segments (and, for n = 5, the matrix into quadrants) and each runs
//              the register-resident DMMA sweep of qsb_warp.cuh on its part: no data exchanged between warps;
//   J^T J      real DMMA SYRK over 16x16 blocks of the (P+1) x (P+1) Gram matrix of [J_r ; r], rows read from L2;
//   trial r    forward-only sweep on 8-column blocks of X, one warp per block;
//   Cholesky   as in qsb_solver.cuh.
#pragma once
#include "qsb_solver.cuh"
#include "qsb_warp.cuh"

#define QSB_LM_WARPS 8

// A[k][l] = sum_m R[k][m] R[l][m] for k, l < rows (both triangles written; ld = rows).  R row-major [rows][K2].
QSB_DEVICE void qsb_syrk_dmma(const double* R, int rows, int K2, double* A, int warp, int nwarps, int lane) {
  const int g = lane >> 2, t = lane & 3;
  const int nb = (rows + 15) >> 4, ntasks = nb * (nb + 1) / 2;
  for (int task = warp; task < ntasks; task += nwarps) {
    int bk = (int)((sqrt(8.0 * task + 1.0) - 1.0) * 0.5);
    while ((bk + 1) * (bk + 2) / 2 <= task) ++bk;
    while (bk * (bk + 1) / 2 > task) --bk;
    const int bl = task - bk * (bk + 1) / 2;
    double acc[2][2][2];
    const double* pa[2];
    const double* pb[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      int ra = 16 * bk + 8 * i + g, rb = 16 * bl + 8 * i + g;
      if (ra > rows - 1) ra = rows - 1;
      if (rb > rows - 1) rb = rows - 1;
      pa[i] = R + (size_t)ra * K2 + t;
      pb[i] = R + (size_t)rb * K2 + t;
#pragma unroll
      for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    }
#pragma unroll 4
    for (int m0 = 0; m0 < K2; m0 += 4) {
      const double a0 = pa[0][m0], a1 = pa[1][m0], b0 = pb[0][m0], b1 = pb[1][m0];
      qsb_dmma884(acc[0][0][0], acc[0][0][1], a0, b0);
      qsb_dmma884(acc[0][1][0], acc[0][1][1], a0, b1);
      qsb_dmma884(acc[1][0][0], acc[1][0][1], a1, b0);
      qsb_dmma884(acc[1][1][0], acc[1][1][1], a1, b1);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int k = 16 * bk + 8 * i + g, l = 16 * bl + 8 * j + 2 * t + e;
          if (k < rows && l < rows) { A[(size_t)k * rows + l] = acc[i][j][e]; A[(size_t)l * rows + k] = acc[i][j][e]; }
        }
  }
}

// Forward-only sweep on the 8 columns [8 cb, 8 cb + 8) of X = U T^dagger: returns this block's share of
// |X - I|_F^2 (every lane) and of Tr(X).
template <int NQ>
QSB_DEVICE void qsb_warp_forward8(const QsbDevProgram& P, const double* target, const double* trig, int cb, int lane,
                                  double& norm2, cplx& tr) {
  constexpr int N = 1 << NQ, KS = N / 4;
  const int g = lane >> 2, t = lane & 3, c = 8 * cb + g;
  QsbFrag<KS, 1> X;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const double2 tv = *reinterpret_cast<const double2*>(target + 2 * (c * N + 4 * ks + t));
    X.re[ks][0] = tv.x; X.im[ks][0] = -tv.y;
  }
  for (int j = 0; j < P.n_ops; ++j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) { qsb_cnot<NQ, 1>(X, op.q0, op.q1, lane); continue; }
    cplx m[4];
    qsb_warp_gate(op, trig, P.consts, m);
    qsb_mix<NQ, 1>(X, op.q0, m, lane);
  }
  double s = 0.0;
  cplx d = cmk(0, 0);
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const bool diag = (4 * ks + t) == c;
    const double re = X.re[ks][0] - (diag ? 1.0 : 0.0), im = X.im[ks][0];
    s = fma(re, re, s); s = fma(im, im, s);
    if (diag) d = cmk(X.re[ks][0], X.im[ks][0]);
  }
  for (int m = 16; m > 0; m >>= 1) {
    s += __shfl_xor_sync(QSB_FULL, s, m);
    d.x += __shfl_xor_sync(QSB_FULL, d.x, m);
    d.y += __shfl_xor_sync(QSB_FULL, d.y, m);
  }
  norm2 = s; tr = d;
}

// CTA-wide: (cos, sin) table of x, then F = |U T^dagger - I|_F^2 / 2 and Tr(U T^dagger).  part: >= 3 * (N/8) doubles.
template <int NQ>
QSB_DEVICE double qsb_cta_half_norm(const QsbDevProgram& P, const double* x, const double* target, double* trig,
                                    double* part, cplx* tr_out) {
  constexpr int NCB = (1 << NQ) / 8;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  for (int p = threadIdx.x; p < P.n_params; p += blockDim.x) {
    double s, c;
    sincos(P.pscale[p] * x[p], &s, &c);
    trig[2 * p] = c; trig[2 * p + 1] = s;
  }
  __syncthreads();
  if (warp < NCB) {
    double n2; cplx tr;
    qsb_warp_forward8<NQ>(P, target, trig, warp, lane, n2, tr);
    if (lane == 0) { part[3 * warp] = n2; part[3 * warp + 1] = tr.x; part[3 * warp + 2] = tr.y; }
  }
  __syncthreads();
  double F = 0.0;
  cplx tr = cmk(0, 0);
  for (int w = 0; w < NCB; ++w) { F += part[3 * w]; tr.x += part[3 * w + 1]; tr.y += part[3 * w + 2]; }
  if (tr_out) *tr_out = tr;
  return 0.5 * F;
}

// first op of each of the NSEG parameter-balanced segments of a program; seg[NSEG] = n_ops
QSB_DEVICE void qsb_cta_segments(const QsbDevProgram& P, int nseg, int* seg) {
  if ((int)threadIdx.x <= nseg) {
    const int s = threadIdx.x;
    int j = 0;
    if (s == nseg) j = P.n_ops;
    else if (s > 0) {
      const int lo = (int)(((long long)s * P.n_params + nseg - 1) / nseg);
      for (j = 0; j < P.n_ops; ++j) {
        const qsb_op op = P.ops[j];
        if (qsb_num_params(op.opcode) > 0 && op.param_off >= lo) break;
      }
    }
    seg[s] = j;
  }
}

// r -> row P of R, J_r -> rows 0..P-1 (R row-major [(P+1)][2 N^2]); trig must hold the table of the current x
template <int NQ>
QSB_DEVICE void qsb_cta_jac(const QsbDevProgram& P, const double* target, const double* trig, const int* seg, double* R) {
  using G = QsbWarpGeom<NQ>;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tile = warp % G::WARPS, s = warp / G::WARPS;
  qsb_warp_jac_range<NQ, QSB_RUN_RESID>(P, target, trig, tile, lane, seg[s], seg[s + 1],
                                        R + (size_t)P.n_params * 2 * G::NN, R);
}

template <int NQ>
QSB_GLOBAL __launch_bounds__(QSB_LM_WARPS * 32) qsb_k_lm2(QsbSolveArgs a) {
  using G = QsbWarpGeom<NQ>;
  constexpr int N = G::N, NN = G::NN, M = 2 * NN, NSEG = QSB_LM_WARPS / G::WARPS;
  QSB_DYN_SMEM(smem_raw);
  cplx* red = (cplx*)smem_raw;  // 32
  double* x = (double*)(red + 32);
  double* xn = x + a.max_params;
  double* g = xn + a.max_params;
  double* h = g + a.max_params;
  double* d2 = h + a.max_params;
  double* wk = d2 + a.max_params;
  double* wy = wk + a.max_params;
  double* trig = wy + a.max_params;         // 2 * maxP
  double* part = trig + 2 * a.max_params;   // 16
  int* seg = (int*)(part + 16);             // NSEG + 1
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
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
    qsb_cta_segments(prog, NSEG, seg);
    int iters = 0, status = 0;
    cplx tr;
    double F = qsb_cta_half_norm<NQ>(prog, x, a.target, trig, part, &tr);  // also fills trig for x
    if (P > 0) {
      qsb_cta_jac<NQ>(prog, a.target, trig, seg, Rg);
      iters = 1;
      __syncthreads();
      qsb_syrk_dmma(Rg, rows, M, A, warp, QSB_LM_WARPS, lane);
      __syncthreads();
      for (int k = threadIdx.x; k < P; k += blockDim.x) {
        g[k] = A[(size_t)P * rows + k];
        const double akk = A[(size_t)k * rows + k];
        d2[k] = akk > 0.0 ? akk : 1.0;
      }
      F = 0.5 * A[(size_t)P * rows + P];
      __syncthreads();
      double mu = 1e-3, nu = 2.0;
      int rejects = 0;
      while (true) {
        if (!(F > 1e-300) || (a.stop_f > 0.0 && F <= a.stop_f)) { status = 4; break; }
        {  // MINPACK's gtol: max cosine between r and the columns of J_r
          double gm = 0.0;
          const double rn = sqrt(2.0 * F);
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            const double akk = A[(size_t)k * rows + k];
            if (akk > 0.0) gm = fmax(gm, fabs(g[k]) / (sqrt(akk) * rn));
          }
          gm = qsb_block_max(gm, red);
          if (gm <= a.gtol) { status = 1; break; }
        }
        if (!qsb_cholesky(A, rows, d2, mu, P, L, wk)) {
          mu *= nu; nu *= 2.0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
          continue;
        }
        for (int k = threadIdx.x; k < P; k += blockDim.x) wk[k] = -g[k];
        qsb_chol_solve(L, P, wk, wy, h);
        cplx p1 = cmk(0, 0), p2 = cmk(0, 0);
        for (int k = threadIdx.x; k < P; k += blockDim.x) {
          xn[k] = x[k] + h[k];
          p1.x = fma(d2[k] * h[k], h[k], p1.x);                 // |D h|^2
          p1.y = fma(d2[k] * x[k], x[k], p1.y);                 // |D x|^2
          p2.x = fma(h[k], mu * d2[k] * h[k] - g[k], p2.x);     // 2 * predicted reduction
        }
        p1 = qsb_block_sum(p1, red);
        p2 = qsb_block_sum(p2, red);
        const double hnorm = sqrt(p1.x), xnorm = sqrt(p1.y), pred = 0.5 * p2.x;
        const double Fn = qsb_cta_half_norm<NQ>(prog, xn, a.target, trig, part, &tr);  // trig now belongs to xn
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
          qsb_cta_jac<NQ>(prog, a.target, trig, seg, Rg);
          ++iters;
          __syncthreads();
          qsb_syrk_dmma(Rg, rows, M, A, warp, QSB_LM_WARPS, lane);
          __syncthreads();
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            g[k] = A[(size_t)P * rows + k];
            d2[k] = fmax(d2[k], A[(size_t)k * rows + k]);
          }
          F = 0.5 * A[(size_t)P * rows + P];
          __syncthreads();
          const double q = 2.0 * rho - 1.0;
          mu *= fmax(1.0 / 3.0, 1.0 - q * q * q);
          nu = 2.0;
        } else {
          if (hnorm <= a.xtol * xnorm) { status = 3; break; }
          mu *= nu; nu *= 2.0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
        }
      }
    }
    // final distance 1 - |Tr(U T^dagger)| / N at x   (comparison.matrix_distance_squared, comparison.py:8-22)
    qsb_cta_half_norm<NQ>(prog, x, a.target, trig, part, &tr);
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

static inline size_t qsb_lm2_smem_bytes(int max_params) {
  return 32 * 16 + (size_t)9 * max_params * 8 + 16 * 8 + 16 * 4 + 64;
}
