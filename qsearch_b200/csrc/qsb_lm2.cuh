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

// Warp index as a value the compiler knows to be the same in all lanes (the broadcast of lane 0, cf. CUTLASS
// canonical_warp_idx_sync): branches on it - each warp of the solver walks its own program segment - are then warp-uniform
// for ptxas, which otherwise wraps every shuffle below them in a WARPSYNC.COLLECTIVE sequence (see qsb_k_warp_jac).
#ifdef QSB_EMU
static inline int qsb_warp_index() { return (int)(threadIdx.x >> 5); }
#else
__device__ __forceinline__ int qsb_warp_index() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }
#endif

#ifdef QSB_EMU
#define QSB_CLOCK() 0ll
#define QSB_LDCG(p) (*(p))
#define QSB_LDCG2(p) (*reinterpret_cast<const double2*>(p))
static inline void qsb_cluster_sync() { qsb_emu_cluster_sync(); }
#else
#define QSB_CLOCK() clock64()
#define QSB_LDCG(p) __ldcg(p)
#define QSB_LDCG2(p) __ldcg(reinterpret_cast<const double2*>(p))
// all threads of all CTAs of the cluster; release/acquire orders the global-memory traffic between them
__device__ __forceinline__ void qsb_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
#endif
#define QSB_LM_MAX_CLUSTER 8
// phase timers: thread 0 of each CTA accumulates cycles into a.prof (only when profiling was requested)
#define QSB_PROF_BEGIN() long long _t0 = a.prof ? QSB_CLOCK() : 0ll
#define QSB_PROF_END(slot) do { if (a.prof && threadIdx.x == 0) atomicAdd((unsigned long long*)&a.prof[slot], (unsigned long long)(QSB_CLOCK() - _t0)); } while (0)

// A[k][l] = sum_m R[k][m] R[l][m] for k, l < rows (both triangles written; ld = rows).  R row-major [rows][K2].
// One warp per TB x TB block of the lower triangle (TB = 8 TW); the rows stream from L2, so bigger blocks mean
// proportionally fewer bytes per flop.  task0 / ntask_stride: this warp's position among all cooperating warps.
// KSPLIT = 2: a block's contraction is cut in two halves done by two warps that ADD their results into a zeroed A (two
// addends: the sum does not depend on which lands first, so the result stays deterministic).
template <int TW>
QSB_DEVICE void qsb_syrk_dmma_t(const double* R, int rows, int K2, double* A, int task0, int stride, int lane, int ksplit) {
  constexpr int TB = 8 * TW;
  const int g = lane >> 2, t = lane & 3;
  const int nb = (rows + TB - 1) / TB, ntasks = nb * (nb + 1) / 2 * ksplit;
  const int klen = K2 / ksplit;
  for (int task = task0; task < ntasks; task += stride) {
    const int tile = task / ksplit, kbeg = (task % ksplit) * klen;
    int bk = (int)((sqrt(8.0 * tile + 1.0) - 1.0) * 0.5);
    while ((bk + 1) * (bk + 2) / 2 <= tile) ++bk;
    while (bk * (bk + 1) / 2 > tile) --bk;
    const int bl = tile - bk * (bk + 1) / 2;
    double acc[TW][TW][2];
    const double* pa[TW];
    const double* pb[TW];
#pragma unroll
    for (int i = 0; i < TW; ++i) {
      int ra = TB * bk + 8 * i + g, rb = TB * bl + 8 * i + g;
      if (ra > rows - 1) ra = rows - 1;
      if (rb > rows - 1) rb = rows - 1;
      pa[i] = R + (size_t)ra * K2 + kbeg + 4 * t;
      pb[i] = R + (size_t)rb * K2 + kbeg + 4 * t;
#pragma unroll
      for (int j = 0; j < TW; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    }
    // The order of the contracted index is free as long as both operands agree: lane t takes the 4 consecutive columns
    // 16 u + 4 t .. + 3 of a 16-column chunk (two 16-byte loads) and feeds them to 4 successive DMMAs, instead of one
    // 8-byte load per DMMA.
#pragma unroll 2
    for (int m0 = 0; m0 < klen; m0 += 16) {
      double2 av[TW][2], bv[TW][2];
#pragma unroll
      for (int i = 0; i < TW; ++i) {
        av[i][0] = QSB_LDCG2(pa[i] + m0); av[i][1] = QSB_LDCG2(pa[i] + m0 + 2);
        bv[i][0] = QSB_LDCG2(pb[i] + m0); bv[i][1] = QSB_LDCG2(pb[i] + m0 + 2);
      }
#pragma unroll
      for (int s4 = 0; s4 < 4; ++s4)
#pragma unroll
        for (int i = 0; i < TW; ++i)
#pragma unroll
          for (int j = 0; j < TW; ++j) {
            const double x = (s4 & 1) ? av[i][s4 >> 1].y : av[i][s4 >> 1].x, y = (s4 & 1) ? bv[j][s4 >> 1].y : bv[j][s4 >> 1].x;
            qsb_dmma884(acc[i][j][0], acc[i][j][1], x, y);
          }
    }
#pragma unroll
    for (int i = 0; i < TW; ++i)
#pragma unroll
      for (int j = 0; j < TW; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int k = TB * bk + 8 * i + g, l = TB * bl + 8 * j + 2 * t + e;
          if (k < rows && l < rows) {
            if (ksplit == 1) { A[(size_t)k * rows + l] = acc[i][j][e]; A[(size_t)l * rows + k] = acc[i][j][e]; }
            else {
              atomicAdd(&A[(size_t)k * rows + l], acc[i][j][e]);
              if (bk != bl) atomicAdd(&A[(size_t)l * rows + k], acc[i][j][e]);  // a diagonal block holds both triangles itself
            }
          }
        }
  }
}

// Block size (16 or 32) and K-split (1 or 2) that finish soonest on nwarps warps: rounds x DMMAs per round, 16-wide
// blocks charged 1.5x for their doubled L2 bytes per flop.  Returns tw | ksplit << 8.
QSB_DEVICE int qsb_syrk_plan(int rows, int nwarps) {
  int best = 0;
  long long best_cost = 0;
  for (int c = 0; c < 4; ++c) {
    const int tw = (c & 1) ? 4 : 2, ks = (c & 2) ? 2 : 1;
    const int nb = (rows + 8 * tw - 1) / (8 * tw);
    const long long tasks = (long long)nb * (nb + 1) / 2 * ks, rounds = (tasks + nwarps - 1) / nwarps;
    const long long cost = rounds * (tw * tw) * (tw == 2 ? 3 : 2) / ks * 8 + (ks == 2 ? 1 : 0);
    if (c == 0 || cost < best_cost) { best_cost = cost; best = tw | (ks << 8); }
  }
  return best;
}
QSB_DEVICE bool qsb_syrk_needs_zero(int rows, int nwarps) { return (qsb_syrk_plan(rows, nwarps) >> 8) == 2; }

// warp = index of this warp among the nwarps cooperating warps (all CTAs of the cluster).  When the plan splits K the
// caller must have zeroed A (qsb_syrk_needs_zero) and made that visible to all cooperating warps.
QSB_DEVICE void qsb_syrk_dmma(const double* R, int rows, int K2, double* A, int warp, int nwarps, int lane) {
  const int plan = qsb_syrk_plan(rows, nwarps), ks = plan >> 8;
  if ((plan & 255) == 4) qsb_syrk_dmma_t<4>(R, rows, K2, A, warp, nwarps, lane, ks);
  else qsb_syrk_dmma_t<2>(R, rows, K2, A, warp, nwarps, lane, ks);
}

// Forward-only sweep on the 8 columns [8 cb, 8 cb + 8) of X = U T^dagger: returns this block's share of
// |X - I|_F^2 (every lane) and of Tr(X).
template <int NQ>
QSB_DEVICE void qsb_warp_forward8(const QsbDevProgram& P, const double* target, const cplx* gtab, int cb, int lane,
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
    qsb_warp_gate(op, gtab, 0, m);
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

// CTA-wide: (cos, sin) table and gate table of x, then F = |U T^dagger - I|_F^2 / 2 and Tr(U T^dagger).  part: >= 3 * (N/8) doubles.
template <int NQ, bool ALLW = false>  // ALLW: the CTA has exactly N/8 warps, all of them take a column block (qsb_k_lbfgs2)
QSB_DEVICE double qsb_cta_half_norm(const QsbDevProgram& P, const double* x, const double* target, double* trig,
                                    cplx* gtab, double* part, cplx* tr_out) {
  constexpr int NCB = (1 << NQ) / 8;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  __syncthreads();
  for (int p = threadIdx.x; p < P.n_params; p += blockDim.x) {
    double s, c;
    sincos(P.pscale[p] * x[p], &s, &c);
    trig[2 * p] = c; trig[2 * p + 1] = s;
  }
  __syncthreads();
  qsb_fill_gates(P, trig, gtab, threadIdx.x, blockDim.x, true);  // rank-one derivative slots: qsb_cta_jac reads this table
  __syncthreads();
  if (ALLW || warp < NCB) {
    double n2; cplx tr;
    qsb_warp_forward8<NQ>(P, target, gtab, warp, lane, n2, tr);
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

// r -> row P of R, J_r -> rows 0..P-1 (R row-major [(P+1)][2 N^2]); gtab must hold the gate table of the current x
// wtiles: QSB_LM_WARPS * QSB_WTILE_DOUBLES of shared memory (16-byte aligned)
template <int NQ>
QSB_DEVICE void qsb_cta_jac(const QsbDevProgram& P, const double* target, const cplx* gtab, double* wtiles, const int* seg,
                            int rank, double* R) {
  using G = QsbWarpGeom<NQ>;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  const int tile = warp % G::WARPS, s = rank * (QSB_LM_WARPS / G::WARPS) + warp / G::WARPS;
  qsb_warp_jac_range<NQ, QSB_RUN_RESID>(P, target, gtab, wtiles + (size_t)warp * QSB_WTILE_DOUBLES, tile, lane, seg[s], seg[s + 1],
                                        R + (size_t)P.n_params * 2 * G::NN, R);
}

// ---------------------------------------------------------------------------- blocked Cholesky on the tensor cores
// Factor the augmented damped normal matrix  B = [A + mu diag(d2), g ; g^T, r^T r]  ((P+1) x (P+1), A = J^T J and
// g = J^T r are rows/columns of the Gram matrix of [J_r ; r]) in the work buffer Lb (ld = rows = P+1):
//   lower triangle of rows 0..P-1 : L,   upper triangle : L^T (so both solves read contiguous rows),
//   row P                         : y^T with L y = g  (the forward substitution comes for free).
// Right-looking, 16-wide panels: diagonal block by one warp in shared memory, panel solve one row per thread,
// trailing update C -= P P^T as 16x16 DMMA tiles.  Returns false (uniformly) if a pivot is not positive.
#define QSB_CH_NB 16
#define QSB_CH_LD 17
#define QSB_CH_MAXROWS 272  // P + 1 <= 272 (host checks)

struct QsbCholSmem {
  double D[QSB_CH_NB * QSB_CH_LD];
  double rinv[QSB_CH_NB];
  double Pn[QSB_CH_MAXROWS * QSB_CH_LD];
  int ok;
};

QSB_DEVICE bool qsb_chol_blocked(const double* A, int rows, const double* d2, double mu, int P, double* Lb, QsbCholSmem* cs) {
  constexpr int NB = QSB_CH_NB, LD = QSB_CH_LD;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index(), nwarps = blockDim.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  __syncthreads();
  for (int e = threadIdx.x; e < rows * rows; e += blockDim.x) {
    const int i = e / rows, j = e - i * rows;
    if (j <= i) Lb[e] = QSB_LDCG(A + e) + ((i == j && i < P) ? mu * d2[i] : 0.0);
  }
  if (threadIdx.x == 0) cs->ok = 1;
  __syncthreads();
  for (int kb = 0; kb < P; kb += NB) {
    const int nbk = (P - kb < NB) ? P - kb : NB;
    // 1. diagonal block (one warp, shared memory)
    if (warp == 0) {
      for (int e = lane; e < NB * NB; e += 32) {
        const int i = e / NB, j = e % NB;
        double v = (i == j) ? 1.0 : 0.0;
        if (i < nbk && j < nbk && j <= i) v = Lb[(size_t)(kb + i) * rows + kb + j];
        else if (i < nbk && j < nbk) v = 0.0;
        cs->D[i * LD + j] = v;
      }
      __syncwarp();
      for (int j = 0; j < nbk; ++j) {
        const double piv = cs->D[j * LD + j];
        if (!(piv > 0.0) || !(piv < 1e300)) { if (lane == 0) cs->ok = 0; break; }
        const double s = sqrt(piv), inv = 1.0 / s;
        __syncwarp();
        if (lane == j) { cs->D[j * LD + j] = s; cs->rinv[j] = inv; }
        else if (lane > j && lane < nbk) cs->D[lane * LD + j] *= inv;
        __syncwarp();
        if (lane > j && lane < nbk) {
          const double lij = cs->D[lane * LD + j];
          for (int k = j + 1; k <= lane; ++k) cs->D[lane * LD + k] = fma(-lij, cs->D[k * LD + j], cs->D[lane * LD + k]);
        }
        __syncwarp();
      }
      for (int j = nbk + lane; j < NB; j += 32) cs->rinv[j] = 1.0;
      for (int e = lane; e < nbk * nbk; e += 32) {  // write L and L^T of the block
        const int i = e / nbk, j = e % nbk;
        if (j <= i) {
          const double v = cs->D[i * LD + j];
          Lb[(size_t)(kb + i) * rows + kb + j] = v;
          Lb[(size_t)(kb + j) * rows + kb + i] = v;
        }
      }
    }
    __syncthreads();
    if (!cs->ok) { __syncthreads(); return false; }
    // 2. panel: rows below the block, X L_kk^T = A_panel, one row per thread
    const int R0 = kb + nbk, nrow = rows - R0;
    for (int ri = threadIdx.x; ri < nrow; ri += blockDim.x) {
      const int i = R0 + ri;
      double x[NB];
      double* src = Lb + (size_t)i * rows + kb;
#pragma unroll
      for (int j = 0; j < NB; ++j) x[j] = (j < nbk) ? src[j] : 0.0;
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        double s = x[j];
#pragma unroll
        for (int m = 0; m < j; ++m) s = fma(-x[m], cs->D[j * LD + m], s);
        x[j] = (j < nbk) ? s * cs->rinv[j] : 0.0;
      }
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        cs->Pn[ri * LD + j] = x[j];
        if (j < nbk) { src[j] = x[j]; Lb[(size_t)(kb + j) * rows + i] = x[j]; }
      }
    }
    __syncthreads();
    // 3. trailing update of the rows below R0 (the augmented row P rides along); nothing to do if only row P is left
    if (R0 < P) {
      const int nbt = (nrow + 15) >> 4, ntasks = nbt * (nbt + 1) / 2;
      for (int task = warp; task < ntasks; task += nwarps) {
        int bi = (int)((sqrt(8.0 * task + 1.0) - 1.0) * 0.5);
        while ((bi + 1) * (bi + 2) / 2 <= task) ++bi;
        while (bi * (bi + 1) / 2 > task) --bi;
        const int bj = task - bi * (bi + 1) / 2;
        double acc[2][2][2];
        int ra[2], rb[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          ra[i] = 16 * bi + 8 * i + g; if (ra[i] > nrow - 1) ra[i] = nrow - 1;
          rb[i] = 16 * bj + 8 * i + g; if (rb[i] > nrow - 1) rb[i] = nrow - 1;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int r = 16 * bi + 8 * i + g, c = 16 * bj + 8 * j + 2 * t + e;
              acc[i][j][e] = (r < nrow && c < nrow) ? Lb[(size_t)(R0 + r) * rows + R0 + c] : 0.0;
            }
#pragma unroll
        for (int ks = 0; ks < NB / 4; ++ks) {
          const double a0 = -cs->Pn[ra[0] * LD + 4 * ks + t], a1 = -cs->Pn[ra[1] * LD + 4 * ks + t];
          const double b0 = cs->Pn[rb[0] * LD + 4 * ks + t], b1 = cs->Pn[rb[1] * LD + 4 * ks + t];
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
              const int r = 16 * bi + 8 * i + g, c = 16 * bj + 8 * j + 2 * t + e;
              if (r < nrow && c < nrow && c <= r) Lb[(size_t)(R0 + r) * rows + R0 + c] = acc[i][j][e];
            }
      }
    }
    __syncthreads();
  }
  return true;
}

// h = -(L^T)^{-1} y with y = row P of Lb (see qsb_chol_blocked); one warp, rows of L^T are contiguous in Lb.
QSB_DEVICE void qsb_chol_back(const double* Lb, int rows, int P, double* h) {
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  __syncthreads();
  if (warp == 0) {
    const double* y = Lb + (size_t)P * rows;
    for (int j = P - 1; j >= 0; --j) {
      const double* u = Lb + (size_t)j * rows;  // row j of L^T: u[i] = L[i][j], i >= j
      double s = 0.0;
      for (int i = j + 1 + lane; i < P; i += 32) s = fma(u[i], h[i], s);
      for (int m = 16; m > 0; m >>= 1) s += __shfl_xor_sync(QSB_FULL, s, m);
      if (lane == 0) h[j] = (-y[j] - s) / u[j];
      __syncwarp();
    }
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------- the same factorisation, shared memory
// Packed lower triangle Ls[i (i+1) / 2 + j] of the augmented matrix (rows = P + 1) entirely in shared memory: no
// L2 round trips inside the factorisation.  Used whenever it fits (P <= ~220); otherwise qsb_chol_blocked.
#define QSB_TRI(i, j) ((size_t)(i) * ((i) + 1) / 2 + (j))

// diagonal block [kb, kb + nbk) of the packed factor: one warp, lane = row, the row in registers, columns exchanged by
// shuffle.  A short block is padded with identity rows, and the entries above the diagonal are left to hold garbage
// that nothing reads, so the 16 column steps are straight-line code: broadcast, rsqrt, 15 - j shuffle + FMA pairs.
QSB_DEVICE void qsb_chp_diag(double* Ls, int kb, int nbk, double* rinv, double* rdiag, int* okflag, int lane) {
  constexpr int NB = QSB_CH_NB;
  const int i = kb + lane;
  double d[NB];
#pragma unroll
  for (int j = 0; j < NB; ++j) d[j] = (lane < nbk && j <= lane) ? Ls[QSB_TRI(i, kb + j)] : ((j == lane) ? 1.0 : 0.0);
  bool good = true;
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    const double piv = __shfl_sync(QSB_FULL, d[j], j);
    if (!(piv > 0.0) || !(piv < 1e300)) { good = false; break; }
    const double inv = rsqrt(piv);  // one reciprocal-square-root sequence instead of sqrt + divide
    if (lane == j && j < nbk) { rinv[j] = inv; rdiag[kb + j] = inv; }
    d[j] = (lane == j) ? piv * inv : d[j] * inv;
#pragma unroll
    for (int k = j + 1; k < NB; ++k) d[k] = fma(-d[j], __shfl_sync(QSB_FULL, d[j], k), d[k]);
  }
  if (!good && lane == 0) *okflag = 0;
  if (lane < nbk) {
#pragma unroll
    for (int j = 0; j < NB; ++j) if (j <= lane) Ls[QSB_TRI(i, kb + j)] = d[j];
  }
}

// one 16x16 tile (bi, bj) of the trailing update C -= P P^T behind panel [kb, kb + nbk); the trailing matrix starts at R0
QSB_DEVICE void qsb_chp_tile(double* Ls, int R0, int kb, int nbk, int nrow, int bi, int bj, int lane) {
  constexpr int NB = QSB_CH_NB;
  const int g = lane >> 2, t = lane & 3;
  double acc[2][2][2];
  const double* pa[2];
  const double* pb[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    int ra = 16 * bi + 8 * i + g, rb = 16 * bj + 8 * i + g;
    if (ra > nrow - 1) ra = nrow - 1;
    if (rb > nrow - 1) rb = nrow - 1;
    pa[i] = Ls + QSB_TRI(R0 + ra, kb) + t;
    pb[i] = Ls + QSB_TRI(R0 + rb, kb) + t;
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int r = 16 * bi + 8 * i + g, c = 16 * bj + 8 * j + 2 * t + e;
        acc[i][j][e] = (r < nrow && c <= r) ? Ls[QSB_TRI(R0 + r, R0 + c)] : 0.0;
      }
#pragma unroll
  for (int ks = 0; ks < NB / 4; ++ks) {
    const bool in = 4 * ks + t < nbk;
    const double a0 = in ? -pa[0][4 * ks] : 0.0, a1 = in ? -pa[1][4 * ks] : 0.0;
    const double b0 = in ? pb[0][4 * ks] : 0.0, b1 = in ? pb[1][4 * ks] : 0.0;
    qsb_dmma884(acc[0][0][0], acc[0][0][1], a0, b0);
    qsb_dmma884(acc[0][1][0], acc[0][1][1], a0, b1);
    qsb_dmma884(acc[1][0][0], acc[1][0][1], a1, b0);
    qsb_dmma884(acc[1][1][0], acc[1][1][1], a1, b1);
  }
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int r = 16 * bi + 8 * i + g, c = 16 * bj + 8 * j + 2 * t + e;
        if (r < nrow && c <= r) Ls[QSB_TRI(R0 + r, R0 + c)] = acc[i][j][e];
      }
}

QSB_DEVICE bool qsb_chol_packed(const double* A, int rows, const double* d2, double mu, int P, double* Ls, double* rinv, double* rdiag, int* okflag,
                                long long* prof) {
#define QSB_CH_T(slot) do { if (prof && threadIdx.x == 0) { const long long _n = QSB_CLOCK(); atomicAdd((unsigned long long*)&prof[slot], (unsigned long long)(_n - _tc)); _tc = _n; } } while (0)
  long long _tc = prof ? QSB_CLOCK() : 0ll;
  constexpr int NB = QSB_CH_NB;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index(), nwarps = blockDim.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  __syncthreads();
  // copy in the lower triangle, one row per warp and round; a row's loads (<= 9 per lane) are all issued before the
  // first store, so the L2 latency is paid once per row, not once per element
  for (int i0 = 2 * warp; i0 < rows; i0 += 2 * nwarps) {  // two rows per warp and round: up to 18 loads in flight per lane
    const int i1 = (i0 + 1 < rows) ? i0 + 1 : i0;
    const double* ar0 = A + (size_t)i0 * rows;
    const double* ar1 = A + (size_t)i1 * rows;
    double v0[9], v1[9];
#pragma unroll
    for (int m = 0; m < 9; ++m) {
      const int j = lane + 32 * m;
      v0[m] = (j <= i0) ? QSB_LDCG(ar0 + j) : 0.0;
      v1[m] = (j <= i1) ? QSB_LDCG(ar1 + j) : 0.0;
    }
    double* lr0 = Ls + QSB_TRI(i0, 0);
    double* lr1 = Ls + QSB_TRI(i1, 0);
#pragma unroll
    for (int m = 0; m < 9; ++m) {
      const int j = lane + 32 * m;
      if (j <= i0) lr0[j] = v0[m] + ((j == i0 && i0 < P) ? mu * d2[i0] : 0.0);
      if (j <= i1 && i1 != i0) lr1[j] = v1[m] + ((j == i1 && i1 < P) ? mu * d2[i1] : 0.0);
    }
  }
  if (threadIdx.x == 0) *okflag = 1;
  __syncthreads();
  QSB_CH_T(6);
  // Look-ahead: the diagonal block of panel k+1 is factored by warp 0 while the other warps finish the trailing update of
  // panel k (only the first block column of the trailing matrix has to be done before that), so the serial column chain
  // of the diagonal block is hidden behind tensor-core work except for the very first block.
  if (warp == 0) qsb_chp_diag(Ls, 0, (P < NB) ? P : NB, rinv, rdiag, okflag, lane);
  __syncthreads();
  QSB_CH_T(7);
  for (int kb = 0; kb < P; kb += NB) {
    const int nbk = (P - kb < NB) ? P - kb : NB;
    if (!*okflag) { __syncthreads(); return false; }
    const int R0 = kb + nbk, nrow = rows - R0;
    for (int ri = threadIdx.x; ri < nrow; ri += blockDim.x) {  // 2. panel, one row per thread
      double* src = Ls + QSB_TRI(R0 + ri, kb);
      double x[NB];
#pragma unroll
      for (int j = 0; j < NB; ++j) x[j] = (j < nbk) ? src[j] : 0.0;
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        if (j < nbk) {
          double s = x[j];
          const double* lj = Ls + QSB_TRI(kb + j, kb);
#pragma unroll
          for (int m = 0; m < j; ++m) s = fma(-x[m], lj[m], s);
          x[j] = s * rinv[j];
          src[j] = x[j];
        }
      }
    }
    __syncthreads();
    QSB_CH_T(8);
    if (R0 < P) {  // 3. trailing update, 16x16 DMMA tiles of the lower triangle
      const int nbt = (nrow + 15) >> 4;
      for (int bi = warp; bi < nbt; bi += nwarps) qsb_chp_tile(Ls, R0, kb, nbk, nrow, bi, 0, lane);  // 3a. first block column
      __syncthreads();
      const int nbk1 = (P - R0 < NB) ? P - R0 : NB;
      if (warp == 0) qsb_chp_diag(Ls, R0, nbk1, rinv, rdiag, okflag, lane);  // 3b. next diagonal block ...
      if (warp != 0 || nwarps == 1) {                                         // ... under the rest of the update
        const int nw = (nwarps == 1) ? 1 : nwarps - 1, w = (nwarps == 1) ? 0 : warp - 1;
        int bi = 1, bj = 1;
        for (int skip = w; skip > 0; --skip) { if (++bj > bi) { ++bi; bj = 1; } }
        while (bi < nbt) {
          qsb_chp_tile(Ls, R0, kb, nbk, nrow, bi, bj, lane);
          for (int skip = nw; skip > 0; --skip) { if (++bj > bi) { ++bi; bj = 1; } }
        }
      }
      __syncthreads();
    }
    QSB_CH_T(9);
  }
  if (!*okflag) { __syncthreads(); return false; }
  return true;
#undef QSB_CH_T
}

// h = -(L^T)^{-1} y, y = row P of the packed factor.  One warp, column sweep: once h_j is known every lane folds it
// into its own entries of the right-hand side (row j of L is contiguous), so there is no reduction on the critical path.
QSB_DEVICE void qsb_chol_back_packed(const double* Ls, int P, const double* rdiag, double* h) {
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  __syncthreads();
  if (warp == 0) {
    // the right-hand side lives in registers (lane l owns rows l, l + 32, ...): a step is one multiply, one shuffle
    // broadcast and one FMA per owned row; the factor's row j is read from shared memory off the critical path
    constexpr int NR = (QSB_CH_MAXROWS + 31) / 32;
    const double* y = Ls + QSB_TRI(P, 0);
    double hr[NR], rd[NR];
#pragma unroll
    for (int b = 0; b < NR; ++b) {
      const int i = 32 * b + lane;
      hr[b] = (i < P) ? -y[i] : 0.0;
      rd[b] = (i < P) ? rdiag[i] : 0.0;
    }
#pragma unroll
    for (int b = NR - 1; b >= 0; --b) {
      if (32 * b < P) {
        const int jhi = (P - 1 < 32 * b + 31) ? P - 1 : 32 * b + 31;
        for (int j = jhi; j >= 32 * b; --j) {
          const double hj = __shfl_sync(QSB_FULL, hr[b] * rd[b], j & 31);
          if (lane == (j & 31)) hr[b] = hj;
          const double* lj = Ls + QSB_TRI(j, 0);
#pragma unroll
          for (int c = 0; c <= b; ++c) {
            const int i = 32 * c + lane;
            if (i < j) hr[c] = fma(-lj[i], hj, hr[c]);
          }
        }
      }
    }
#pragma unroll
    for (int b = 0; b < NR; ++b) {
      const int i = 32 * b + lane;
      if (i < P) h[i] = hr[b];
    }
  }
  __syncthreads();
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
  cplx* gtab = (cplx*)(part + 16 + (a.max_params & 1));  // [max_ops + max_params][4], 16-byte aligned
  int* seg = (int*)(gtab + (size_t)4 * (a.max_ops + a.max_params));  // cluster * NSEG + 1 (<= 72 ints)
  qsb_op* sops = (qsb_op*)(seg + 72);                                 // the program's ops, staged once per instance
  QsbCholSmem* cs = (QsbCholSmem*)(sops + a.max_ops);  // global-memory factorisation (large P)
  double* Ls = (double*)(sops + a.max_ops);             // or: packed factor in shared memory (a.chol_smem); during the
                                                        // Jacobian phase the same bytes are the warps' W tiles
  // aligned through the offset, not an integer cast of the pointer (which would turn the W-tile accesses into generic LD / ST)
  double* wtiles = (double*)((unsigned char*)smem_raw + (((size_t)((unsigned char*)Ls - (unsigned char*)smem_raw) + 15) & ~(size_t)15));
  double* rinv = Ls + (size_t)(a.max_params + 1) * (a.max_params + 2) / 2;
  int* okflag = (int*)(rinv + QSB_CH_NB);
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  // a cluster of a.cluster CTAs works on one instance: they split the Jacobian segments and the SYRK blocks, share
  // R and A through global memory (two cluster barriers per iteration) and repeat the small dense algebra redundantly,
  // so every CTA takes the same decisions without a broadcast.
  const int csize = a.cluster, rank = blockIdx.x % csize, cid = blockIdx.x / csize, nclusters = gridDim.x / csize;
  const int cwarps = csize * QSB_LM_WARPS, cwarp = rank * QSB_LM_WARPS + warp;
  double* Rg = a.scratch + (size_t)cid * a.scratch_per_cta;
  for (int inst = cid; inst < a.count; inst += nclusters) {
    const int pidx = a.prog_idx[inst];
    QsbDevProgram prog = a.progs[pidx];
    const int P = prog.n_params, rows = P + 1;
    __syncthreads();
    for (int e = threadIdx.x; e < prog.n_ops * 6; e += blockDim.x) ((int32_t*)sops)[e] = ((const int32_t*)prog.ops)[e];
    prog.ops = sops;  // every sweep of this solve reads the ops from shared memory
    // two Gram buffers used alternately: the one the next SYRK adds into is zeroed during the Jacobian phase, when no
    // CTA of the cluster can still be reading it (they all passed the barriers of the evaluation before)
    double* Abuf = Rg + (size_t)(a.max_params + 1) * M;
    const size_t asz = (size_t)(a.max_params + 1) * (a.max_params + 1);
    double* A = Abuf;
    double* L = Abuf + 2 * asz + (size_t)rank * asz;
    const bool kzero = qsb_syrk_needs_zero(rows, cwarps);
    if (kzero && csize > 1) qsb_cluster_sync();  // nobody is still reading the previous instance's Gram matrix
    const double* x0 = a.x0 + a.param_off[inst];
    const int max_iters = a.max_iters > 0 ? a.max_iters : 100 * (P > 0 ? P : 1);
    __syncthreads();
    for (int k = threadIdx.x; k < P; k += blockDim.x) x[k] = x0[k];
    qsb_cta_segments(prog, NSEG * csize, seg);
    int iters = 0, status = 0;
    cplx tr;
    double F = qsb_cta_half_norm<NQ>(prog, x, a.target, trig, gtab, part, &tr);  // also fills trig for x
    if (P > 0) {
      {
        QSB_PROF_BEGIN();
        if (kzero) for (int e = rank * blockDim.x + threadIdx.x; e < rows * rows; e += csize * blockDim.x) A[e] = 0.0;
        qsb_cta_jac<NQ>(prog, a.target, gtab, wtiles, seg, rank, Rg); qsb_cluster_sync(); QSB_PROF_END(0);
      }
      iters = 1;
      { QSB_PROF_BEGIN(); qsb_syrk_dmma(Rg, rows, M, A, cwarp, cwarps, lane); qsb_cluster_sync(); QSB_PROF_END(1); }
      for (int k = threadIdx.x; k < P; k += blockDim.x) {
        g[k] = QSB_LDCG(&A[(size_t)P * rows + k]);
        const double akk = QSB_LDCG(&A[(size_t)k * rows + k]);
        d2[k] = akk > 0.0 ? akk : 1.0;
      }
      F = 0.5 * QSB_LDCG(&A[(size_t)P * rows + P]);
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
            const double akk = QSB_LDCG(&A[(size_t)k * rows + k]);
            if (akk > 0.0) gm = fmax(gm, fabs(g[k]) / (sqrt(akk) * rn));
          }
          // batch-level stop rules as in qsb_k_lm3 (circuits too long for its shared-memory factor run here: P > ~190 at
          // n = 5, the deep rounds of a LEAP search).  Only without a cluster: the CTAs of a cluster take their decisions
          // from identical data, and a flag read from global memory at different moments would let them diverge.
          if (csize == 1 && threadIdx.x == 0) {
            if (a.sibling_f > 0.0 && QSB_LDCG(&a.ctl[QSB_CTL_SOLVED + pidx]) != 0) gm = 1e301;
            else if (a.straggler_c > 0.0 && 2 * QSB_LDCG(&a.ctl[QSB_CTL_FINISHED]) >= a.count && QSB_LDCG(&a.ctl[QSB_CTL_MEDIAN]) > 0 &&
                     (double)iters > a.straggler_c * (double)QSB_LDCG(&a.ctl[QSB_CTL_MEDIAN]) && iters >= a.straggler_min) gm = 1e300;
          }
          gm = qsb_block_max(gm, red);
          if (gm >= 1e301) { status = 7; break; }
          if (gm >= 1e300) { status = 6; break; }
          if (gm <= a.gtol) { status = 1; break; }
        }
        bool chol_ok;
        {
          QSB_PROF_BEGIN();
          chol_ok = a.chol_smem ? qsb_chol_packed(A, rows, d2, mu, P, Ls, rinv, wk, okflag, a.prof) : qsb_chol_blocked(A, rows, d2, mu, P, L, cs);
          QSB_PROF_END(2);
        }
        if (!chol_ok) {
          mu *= nu; nu *= a.mu_up; streak = 0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
          continue;
        }
        { QSB_PROF_BEGIN(); if (a.chol_smem) qsb_chol_back_packed(Ls, P, wk, h); else qsb_chol_back(L, rows, P, h); QSB_PROF_END(3); }
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
        double Fn;
        { QSB_PROF_BEGIN(); Fn = qsb_cta_half_norm<NQ>(prog, xn, a.target, trig, gtab, part, &tr); QSB_PROF_END(4); }  // trig now belongs to xn
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
          {
            QSB_PROF_BEGIN();
            A = Abuf + (size_t)(iters & 1) * asz;
            if (kzero) for (int e = rank * blockDim.x + threadIdx.x; e < rows * rows; e += csize * blockDim.x) A[e] = 0.0;
            qsb_cta_jac<NQ>(prog, a.target, gtab, wtiles, seg, rank, Rg); qsb_cluster_sync(); QSB_PROF_END(0);
          }
          ++iters;
          { QSB_PROF_BEGIN(); qsb_syrk_dmma(Rg, rows, M, A, cwarp, cwarps, lane); qsb_cluster_sync(); QSB_PROF_END(1); }
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            g[k] = QSB_LDCG(&A[(size_t)P * rows + k]);
            d2[k] = fmax(d2[k], QSB_LDCG(&A[(size_t)k * rows + k]));
          }
          F = 0.5 * QSB_LDCG(&A[(size_t)P * rows + P]);
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
    qsb_cta_half_norm<NQ>(prog, x, a.target, trig, gtab, part, &tr);
    double* ox = a.out_x + a.param_off[inst];
    for (int k = threadIdx.x; k < P; k += blockDim.x) ox[k] = x[k];
    if (a.prof && threadIdx.x == 0) atomicAdd((unsigned long long*)&a.prof[5], (unsigned long long)iters);
    if (threadIdx.x == 0 && rank == 0) {
      const double dist = fabs(1.0 - sqrt(tr.x * tr.x + tr.y * tr.y) / N);
      a.out_f[inst] = dist;
      a.out_iters[inst] = iters;
      a.out_status[inst] = status;
      if (csize == 1) {  // bookkeeping of the stop rules (see qsb_k_lm3)
        if (a.sibling_f > 0.0 && dist < a.sibling_f) atomicAdd(&a.ctl[QSB_CTL_SOLVED + pidx], 1);
        const int nfin = atomicAdd(&a.ctl[QSB_CTL_FINISHED], 1) + 1;
        if (status != 6 && status != 7 && (2 * nfin <= a.count + 1 || QSB_LDCG(&a.ctl[QSB_CTL_MEDIAN]) == 0))
          atomicMax(&a.ctl[QSB_CTL_MEDIAN], iters);
      }
    }
    __syncthreads();
  }
}

static inline size_t qsb_lm2_smem_bytes(int max_ops, int max_params, bool chol_smem) {
  const size_t base = 32 * 16 + (size_t)9 * max_params * 8 + 16 * 8 + (size_t)64 * (max_ops + max_params) + 72 * 4 + (size_t)24 * max_ops + 88;
  const size_t packed = ((size_t)(max_params + 1) * (max_params + 2) / 2 + QSB_CH_NB + 2) * 8;
  const size_t wtiles = (size_t)QSB_LM_WARPS * QSB_WTILE_DOUBLES * 8 + 16;
  return base + std::max(chol_smem ? packed : sizeof(QsbCholSmem), wtiles);
}


// ---------------------------------------------------------------------------- QSB_RUN_LMSYS
// What one Levenberg-Marquardt step consumes, for every instance of a batch: r^T r, g = J_r^T r and A = J_r^T J_r,
// with J_r only ever written to a per-CTA slab that is immediately re-read by the SYRK (it stays in L2; nothing of
// size 2 N^2 P goes to HBM or to the host).  One CTA of QSB_LM_WARPS warps per instance, same phases as qsb_k_lm2.
// out_a[inst] = r^T r, out_b[param_off ..] = g, out_c[sum P_i^2 ..] = A (row-major P x P).
template <int NQ>
QSB_GLOBAL __launch_bounds__(QSB_LM_WARPS * 32) qsb_k_lmsys(QsbBatchArgs a, double* scratch, size_t scratch_per_cta,
                                                            const long long* sq_off) {
  using G = QsbWarpGeom<NQ>;
  constexpr int M = 2 * G::NN, NSEG = QSB_LM_WARPS / G::WARPS;
  QSB_DYN_SMEM(smem_raw);
  double* trig = (double*)smem_raw;                       // 2 * maxP
  double* part = trig + 2 * a.max_params;                 // 16
  cplx* gtab = (cplx*)(part + 16);
  double* wtiles = (double*)(gtab + (size_t)4 * (a.max_ops + a.max_params));
  int* seg = (int*)(wtiles + (size_t)QSB_LM_WARPS * QSB_WTILE_DOUBLES);
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  double* Rg = scratch + (size_t)blockIdx.x * scratch_per_cta;
  const long long p_base = a.param_off[a.first], s_base = sq_off[a.first];
  for (int li = blockIdx.x; li < a.count; li += gridDim.x) {
    const int inst = a.first + li;
    const QsbDevProgram prog = a.progs[a.prog_idx[inst]];
    const int P = prog.n_params, rows = P + 1;
    double* A = Rg + (size_t)(a.max_params + 1) * M;
    __syncthreads();
    qsb_cta_segments(prog, NSEG, seg);
    cplx tr;
    qsb_cta_half_norm<NQ>(prog, a.params + a.param_off[inst], a.target, trig, gtab, part, &tr);  // trig + gate table
    if (qsb_syrk_needs_zero(rows, QSB_LM_WARPS)) for (int e = threadIdx.x; e < rows * rows; e += blockDim.x) A[e] = 0.0;
    qsb_cta_jac<NQ>(prog, a.target, gtab, wtiles, seg, 0, Rg);
    __syncthreads();
    qsb_syrk_dmma(Rg, rows, M, A, warp, QSB_LM_WARPS, lane);
    __syncthreads();
    double* og = a.out_b + (a.param_off[inst] - p_base);
    double* oA = a.out_c + (sq_off[inst] - s_base);
    for (int k = threadIdx.x; k < P; k += blockDim.x) og[k] = A[(size_t)P * rows + k];
    for (int e = threadIdx.x; e < P * P; e += blockDim.x) oA[e] = A[(size_t)(e / P) * rows + (e % P)];
    if (threadIdx.x == 0) a.out_a[li] = A[(size_t)P * rows + P];
  }
}

static inline size_t qsb_lmsys_smem_bytes(int max_ops, int max_params) {
  return ((size_t)2 * max_params + 16 + 8 * (size_t)(max_ops + max_params) + (size_t)QSB_LM_WARPS * QSB_WTILE_DOUBLES) * 8 + 16 * 4 + 64;
}
