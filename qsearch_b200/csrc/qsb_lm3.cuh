// Batched on-device Levenberg-Marquardt, body-frame version: one CTA per (circuit, start vector) instance, the whole
// solve in one launch, instances handed out from a device-side queue.  Same role and stopping rules as qsb_k_lm2 /
// qsb_k_lm: LeastSquares_Jac_Solver (qsearch/solvers.py:146-163, scipy least_squares(method='lm') = MINPACK lmder) and
// the Ceres DENSE_QR driver (qsrs/src/solvers.rs:88-121).  What changes is the form of the normal equations.
//
// For a UNITARY target T the residual Jacobian of comparison.matrix_residuals_jac (comparison.py:66-72) factors as
//     K_k = dU/dx_k T^dagger = V A_k,   V = U T^dagger,   A_k = X_k^dagger (g_k^dagger dg_k (x) I) X_k = i Q_k,
// with X_k = g_{k-1} .. g_0 T^dagger the prefix BEFORE the op of parameter k and Q_k Hermitian (g^dagger dg is a 2x2
// anti-Hermitian generator).  V is unitary, so it drops out of every inner product:
//     (J_r^T J_r)[k][l] = Re Tr(K_k^dagger K_l) = <Q_k, Q_l>          (independent of the target)
//     (J_r^T r)[k]      = Re Tr(K_k^dagger (V - I)) = -<Q_k, C>,   C = (V^dagger - V) / (2i)   (Hermitian)
//     r^T r             = |V - I|_F^2
// <.,.> is the real Frobenius inner product of Hermitian matrices, which needs only N^2 reals per matrix (lower tiles):
// half the row length of [Re K ; Im K], so the Gram product (the dominant cost of an LM iteration) halves.  With the
// spectral form  -i g^dagger dg = gamma I + mu v v^dagger  every Q_k = gamma I + mu Y^dagger Y, Y = (v^dagger (x) I) X_k
// (N/2 x N): one Hermitian rank-N/2 update on the tensor cores, lower 8x8 tiles only, from a FORWARD-ONLY sweep -
// no suffix matrix, no peel, no dense V product.  Diagonal tiles are stored as Z = Re + Im (Re symmetric, Im
// antisymmetric: <Z_k, Z_l> = <Re_k, Re_l> + <Im_k, Im_l> exactly, the cross terms cancel), which takes 2 DMMAs per
// k-step instead of 4; off-diagonal tiles are stored once, scaled by sqrt(2).
// Non-unitary targets (nothing in the reference forbids them) keep the general kernel qsb_k_lm2: the host checks.
#pragma once
#include "qsb_lm2.cuh"

template <int NQ>
struct QsbQGeom {
  static constexpr int N = 1 << NQ;
  static constexpr int KS = N / 4;               // k-steps of the m8n8k4 instruction over the rows of X
  static constexpr int TT = N / 8;               // 8-wide column tiles: a warp holds ALL columns of X
  static constexpr int NOFF = TT * (TT - 1) / 2; // strictly-lower 8x8 tiles of a Hermitian N x N
  static constexpr int K2 = N * N;               // packed reals per Hermitian matrix: TT * 64 + NOFF * 128
  static constexpr int WARPS = 8;                 // warps per CTA (= per instance)
};

// device-side control block of one solve launch: QSB_CTL_* in qsb_solver.cuh

// slot of a parameter in the gate table: {sqrt(mu) conj(v0), sqrt(mu) conj(v1), (gamma, 0), 0} with
// -i g^dagger dg = gamma I + mu v v^dagger, |v| = 1, mu >= 0 (the eigenvector of the LARGER eigenvalue)
QSB_DEVICE void qsb_body_generator(const cplx* g, const cplx* d, cplx* out) {
  const cplx z = cmk(0, 0);
  const cplx G00 = cfma_conj(g[2], d[2], cfma_conj(g[0], d[0], z));
  const cplx G01 = cfma_conj(g[2], d[3], cfma_conj(g[0], d[1], z));
  const cplx G10 = cfma_conj(g[3], d[2], cfma_conj(g[1], d[0], z));
  const cplx G11 = cfma_conj(g[3], d[3], cfma_conj(g[1], d[1], z));
  // H = -i G = [[a, b], [conj b, dd]]
  const double a = G00.y, dd = G11.y;
  const cplx b = cmk(0.5 * (G01.y + G10.y), 0.5 * (G10.x - G01.x));
  const double mean = 0.5 * (a + dd), del = 0.5 * (a - dd);
  const double rad = sqrt(del * del + b.x * b.x + b.y * b.y);
  cplx v0, v1;
  if (del >= 0.0) { v0 = cmk(del + rad, 0.0); v1 = cmk(b.x, -b.y); }
  else { v0 = b; v1 = cmk(rad - del, 0.0); }
  const double n2 = v0.x * v0.x + v0.y * v0.y + v1.x * v1.x + v1.y * v1.y;
  const double s = n2 > 0.0 ? sqrt(2.0 * rad / n2) : 0.0;
  out[0] = cmk(s * v0.x, -s * v0.y);
  out[1] = cmk(s * v1.x, -s * v1.y);
  out[2] = cmk(mean - rad, 0.0);
  out[3] = z;
}

// gate table for the body-frame sweep: 2x2 of every one-qubit op, spectral factors of every parameter's generator
QSB_DEVICE void qsb_fill_gates_body(const QsbDevProgram& P, const double* trig, cplx* gtab, int tid, int nthreads) {
  const int n_slots = P.n_ops + P.n_params;
  for (int s = tid; s < n_slots; s += nthreads) {
    const int code = P.slotmap[s];
    if (code < 0) continue;
    const qsb_op op = P.ops[code >> 2];
    cplx g[4];
    if (op.opcode == QSB_OP_CONST1) {
      const double* c = P.consts + op.const_off;
      g[0] = cmk(c[0], c[1]); g[1] = cmk(c[2], c[3]); g[2] = cmk(c[4], c[5]); g[3] = cmk(c[6], c[7]);
    } else if ((code & 3) == 0) {
      qsb_gate2x2(op.opcode, trig + 2 * op.param_off, 0, g);
    } else {
      cplx g0[4], d[4];
      qsb_gate2x2(op.opcode, trig + 2 * op.param_off, 0, g0);
      qsb_gate2x2(op.opcode, trig + 2 * op.param_off, code & 3, d);
      qsb_body_generator(g0, d, g);
    }
    gtab[4 * s] = g[0]; gtab[4 * s + 1] = g[1]; gtab[4 * s + 2] = g[2]; gtab[4 * s + 3] = g[3];
  }
}

// ---------------------------------------------------------------------------- Gram product into a PACKED lower triangle
// A[TRI(k, l)] = sum_m R[k][m] R[l][m], l <= k < rows.  Same tiling as qsb_syrk_dmma_t (qsb_lm2.cuh), but
//   * the result goes to the packed layout the factorisation works in (a linear copy brings it to shared memory);
//   * 8-row tiles that lie entirely beyond `rows`, and the upper-right quarter of a diagonal block, are skipped
//     (rows = 166 with 32-wide blocks: 60 instead of 84 16x16 tile products).
template <int TW>
QSB_DEVICE void qsb_syrk_packed_t(const double* R, int rows, int K2, double* A, int task0, int stride, int lane, int ksplit) {
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
    const int ni = (rows - TB * bk + 7) / 8 < TW ? (rows - TB * bk + 7) / 8 : TW;  // 8-row tiles of block bk that hold rows
    const bool diag = bk == bl;
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
#pragma unroll 2
    for (int m0 = 0; m0 < klen; m0 += 16) {
      double2 av[TW][2], bv[TW][2];
#pragma unroll
      for (int i = 0; i < TW; ++i) {
        if (i < ni) { av[i][0] = QSB_LDCG2(pa[i] + m0); av[i][1] = QSB_LDCG2(pa[i] + m0 + 2); }
        else { av[i][0] = av[i][1] = make_double2(0.0, 0.0); }
        bv[i][0] = QSB_LDCG2(pb[i] + m0); bv[i][1] = QSB_LDCG2(pb[i] + m0 + 2);
      }
#pragma unroll
      for (int s4 = 0; s4 < 4; ++s4)
#pragma unroll
        for (int i = 0; i < TW; ++i) {
          if (i >= ni) continue;  // warp-uniform
#pragma unroll
          for (int j = 0; j < TW; ++j) {
            if (diag && j > i) continue;  // only the lower 8x8 tiles of a diagonal block
            const double x = (s4 & 1) ? av[i][s4 >> 1].y : av[i][s4 >> 1].x, y = (s4 & 1) ? bv[j][s4 >> 1].y : bv[j][s4 >> 1].x;
            qsb_dmma884(acc[i][j][0], acc[i][j][1], x, y);
          }
        }
    }
#pragma unroll
    for (int i = 0; i < TW; ++i)
#pragma unroll
      for (int j = 0; j < TW; ++j)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int k = TB * bk + 8 * i + g, l = TB * bl + 8 * j + 2 * t + e;
          if (k < rows && l <= k) {
            if (ksplit == 1) A[QSB_TRI(k, l)] = acc[i][j][e];
            else atomicAdd(&A[QSB_TRI(k, l)], acc[i][j][e]);
          }
        }
  }
}

QSB_DEVICE void qsb_syrk_packed(const double* R, int rows, int K2, double* A, int warp, int nwarps, int lane) {
  const int plan = qsb_syrk_plan(rows, nwarps), ks = plan >> 8;
  if ((plan & 255) == 4) qsb_syrk_packed_t<4>(R, rows, K2, A, warp, nwarps, lane, ks);
  else qsb_syrk_packed_t<2>(R, rows, K2, A, warp, nwarps, lane, ks);
}

// ---------------------------------------------------------------------------- factorisation with inverted diagonal blocks
// Same augmented, packed, right-looking Cholesky as qsb_chol_packed (qsb_lm2.cuh), with three changes that take the
// serial work off the critical path of an LM iteration:
//   * the matrix arrives packed (qsb_syrk_packed), so the copy into shared memory is linear and vectorised;
//   * the warp that factors a 16x16 diagonal block also inverts it (W = L_kk^-1, interleaved with the pivot chain), so
//     the panel below it is ONE tensor-core product X = A_panel W^T instead of a 136-step substitution per row;
//   * the inverses are kept ([block][16][16] in shared memory): the back-substitution is one 16x16 matrix-vector
//     product per block instead of one reduction per row.
#define QSB_CH3_WSZ 256  // doubles per inverted diagonal block

// diagonal block [kb, kb + nbk): lane = row, the row in registers; leaves L in Ls, W (row-major 16x16, zero above the
// diagonal, identity in the padding) in Wb, 1 / L_jj in rdiag[kb + j]
QSB_DEVICE void qsb_ch3_diag(double* Ls, int kb, int nbk, double* Wb, double* rdiag, int* okflag, int lane) {
  constexpr int NB = QSB_CH_NB;
  const int i = kb + lane;
  double d[NB], w[NB];
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    d[j] = (lane < nbk && j <= lane) ? Ls[QSB_TRI(i, kb + j)] : ((j == lane) ? 1.0 : 0.0);
    w[j] = (j == lane) ? 1.0 : 0.0;
  }
  bool good = true;
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    const double piv = __shfl_sync(QSB_FULL, d[j], j);
    if (!(piv > 0.0) || !(piv < 1e300)) { good = false; break; }
    const double inv = rsqrt(piv);
    if (lane == j && j < nbk) rdiag[kb + j] = inv;
    d[j] = (lane == j) ? piv * inv : d[j] * inv;
#pragma unroll
    for (int k = j + 1; k < NB; ++k) d[k] = fma(-d[j], __shfl_sync(QSB_FULL, d[j], k), d[k]);
    // row j of W is final once scaled by 1 / L_jj; rows below subtract L_ij times it (columns c <= j)
#pragma unroll
    for (int c = 0; c <= j; ++c) {
      if (lane == j) w[c] *= inv;
      const double wj = __shfl_sync(QSB_FULL, w[c], j);
      if (lane > j) w[c] = fma(-d[j], wj, w[c]);
    }
  }
  if (!good && lane == 0) *okflag = 0;
  if (lane < NB) {
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      if (lane < nbk && j <= lane) Ls[QSB_TRI(i, kb + j)] = d[j];
      Wb[lane * NB + j] = (j <= lane) ? w[j] : 0.0;
    }
  }
}

// panel below the diagonal block: rows R0 .. rows-1, X = A_panel W^T, one 8-row tile per warp and round
QSB_DEVICE void qsb_ch3_panel(double* Ls, int R0, int kb, int nbk, int nrow, const double* Wb, int warp, int nwarps, int lane) {
  constexpr int NB = QSB_CH_NB;
  const int g = lane >> 2, t = lane & 3;
  double bw[2][NB / 4];  // B fragments of W^T: B[k][n] = W[n][k]
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int ks = 0; ks < NB / 4; ++ks) bw[nt][ks] = Wb[(8 * nt + g) * NB + 4 * ks + t];
  const int nmt = (nrow + 7) >> 3;
  for (int mt = warp; mt < nmt; mt += nwarps) {
    int r = 8 * mt + g;
    const bool rv = r < nrow;
    if (!rv) r = nrow - 1;
    double* src = Ls + QSB_TRI(R0 + r, kb);
    double a[NB / 4];
#pragma unroll
    for (int ks = 0; ks < NB / 4; ++ks) a[ks] = (4 * ks + t < nbk) ? src[4 * ks + t] : 0.0;
    double c0[2] = {0.0, 0.0}, c1[2] = {0.0, 0.0};
#pragma unroll
    for (int ks = 0; ks < NB / 4; ++ks) {
      qsb_dmma884(c0[0], c0[1], a[ks], bw[0][ks]);
      qsb_dmma884(c1[0], c1[1], a[ks], bw[1][ks]);
    }
    __syncwarp();  // every lane has read its row before anybody overwrites it
    if (rv) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        if (2 * t + e < nbk) src[2 * t + e] = c0[e];
        if (8 + 2 * t + e < nbk) src[8 + 2 * t + e] = c1[e];
      }
    }
  }
}

// Apk: packed Gram matrix of [Q ; -C] in global memory (rows = P + 1).  Returns false (uniformly) if a pivot is not positive.
QSB_DEVICE bool qsb_chol3(const double* Apk, int rows, const double* d2, double mu, int P, double* Ls, double* Winv, double* rdiag,
                          int* okflag, long long* prof) {
#define QSB_CH_T(slot) do { if (prof && threadIdx.x == 0) { const long long _n = QSB_CLOCK(); atomicAdd((unsigned long long*)&prof[slot], (unsigned long long)(_n - _tc)); _tc = _n; } } while (0)
  long long _tc = prof ? QSB_CLOCK() : 0ll;
  constexpr int NB = QSB_CH_NB;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index(), nwarps = blockDim.x >> 5;
  const int tri = rows * (rows + 1) / 2;
  __syncthreads();
  for (int e = 2 * threadIdx.x; e + 1 < tri; e += 2 * blockDim.x) {
    const double2 v = QSB_LDCG2(Apk + e);
    *reinterpret_cast<double2*>(Ls + e) = v;
  }
  if ((tri & 1) && threadIdx.x == 0) Ls[tri - 1] = QSB_LDCG(Apk + tri - 1);
  if (threadIdx.x == 0) *okflag = 1;
  __syncthreads();
  for (int k = threadIdx.x; k < P; k += blockDim.x) Ls[QSB_TRI(k, k)] += mu * d2[k];
  __syncthreads();
  QSB_CH_T(6);
  if (warp == 0) qsb_ch3_diag(Ls, 0, (P < NB) ? P : NB, Winv, rdiag, okflag, lane);
  __syncthreads();
  QSB_CH_T(7);
  for (int kb = 0; kb < P; kb += NB) {
    const int nbk = (P - kb < NB) ? P - kb : NB;
    if (!*okflag) { __syncthreads(); return false; }
    const int R0 = kb + nbk, nrow = rows - R0;
    qsb_ch3_panel(Ls, R0, kb, nbk, nrow, Winv + (size_t)(kb / NB) * QSB_CH3_WSZ, warp, nwarps, lane);
    __syncthreads();
    QSB_CH_T(8);
    if (R0 < P) {  // trailing update, 16x16 DMMA tiles of the lower triangle
      const int nbt = (nrow + 15) >> 4;
      for (int bi = warp; bi < nbt; bi += nwarps) qsb_chp_tile(Ls, R0, kb, nbk, nrow, bi, 0, lane);  // first block column
      __syncthreads();
      const int nbk1 = (P - R0 < NB) ? P - R0 : NB;
      if (warp == 0) qsb_ch3_diag(Ls, R0, nbk1, Winv + (size_t)(R0 / NB) * QSB_CH3_WSZ, rdiag, okflag, lane);  // next diagonal block ...
      if (warp != 0 || nwarps == 1) {                                                                          // ... under the rest of the update
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

// h = -(L^T)^-1 y, y = row P of the packed factor: one warp, one 16-row block per step.  With W_b = L_bb^-1:
//   h_b = W_b^T (z_b),   z = -y - sum over later blocks c of L_cb^T h_c  (folded into z as soon as h_c is known)
// zs: 32 doubles of shared memory.
QSB_DEVICE void qsb_ch3_back(const double* Ls, int P, const double* Winv, double* zs, double* h) {
  constexpr int NB = QSB_CH_NB;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  __syncthreads();
  if (warp == 0) {
    constexpr int NR = (QSB_CH_MAXROWS + 31) / 32;
    const double* y = Ls + QSB_TRI(P, 0);
    double z[NR];
#pragma unroll
    for (int b = 0; b < NR; ++b) {
      const int i = 32 * b + lane;
      z[b] = (i < P) ? -y[i] : 0.0;
    }
    for (int blk = (P + NB - 1) / NB - 1; blk >= 0; --blk) {
      const int kb = NB * blk;
      const double* Wb = Winv + (size_t)blk * QSB_CH3_WSZ;
      const int chunk = kb >> 5, off = kb & 31, li = lane - off;  // the block's rows sit in z[chunk], lanes off .. off + 15
      double zb = 0.0;
#pragma unroll
      for (int b = 0; b < NR; ++b) if (b == chunk) zb = z[b];
      if (li >= 0 && li < NB) zs[li] = zb;
      __syncwarp();
      double hb = 0.0;
      if (li >= 0 && li < NB) {
#pragma unroll
        for (int j = 0; j < NB; ++j) hb = fma(Wb[j * NB + li], zs[j], hb);  // W is zero above its diagonal: j >= li only
        zs[NB + li] = hb;
        if (kb + li < P) h[kb + li] = hb;
      }
      __syncwarp();
      // fold h_b into the right-hand side of the earlier rows: z[i] -= sum_j L[kb + j][i] h_b[j], i < kb
      const int nbk = (P - kb < NB) ? P - kb : NB;
#pragma unroll
      for (int b = 0; b < NR; ++b) {
        const int i = 32 * b + lane;
        if (i < kb) {
          double s = z[b];
          for (int j = 0; j < nbk; ++j) s = fma(-Ls[QSB_TRI(kb + j, i)], zs[NB + j], s);
          z[b] = s;
        }
      }
      __syncwarp();
    }
  }
  __syncthreads();
}

// One parameter whose qubit sits at k-step bit KB: packed Q = gamma I + Y^dagger Y, Y = (w^T (x) I) X, into `row`
// (K2 doubles).  Layout of a row: TT diagonal tiles (Z = Re + Im, 64 doubles each, C-fragment order: lane-major pairs),
// then the NOFF tiles (mt > nt, index mt (mt - 1) / 2 + nt): sqrt(2) Re (64 doubles), sqrt(2) Im (64 doubles).
template <int NQ, int KB>
QSB_DEVICE void qsb_q_param(const QsbFrag<(1 << NQ) / 4, (1 << NQ) / 8>& X, const cplx* slot, double* row, int lane) {
  using G = QsbQGeom<NQ>;
  constexpr int KS = G::KS, TT = G::TT, NOFF = G::NOFF, NO1 = NOFF > 0 ? NOFF : 1;
  const int g = lane >> 2, t = lane & 3;
  const cplx w0 = slot[0], w1 = slot[1];
  const double gam = slot[2].x;
  double zd[TT][2], hr[NO1][2], hi[NO1][2];
#pragma unroll
  for (int tt = 0; tt < TT; ++tt) { zd[tt][0] = (2 * t == g) ? gam : 0.0; zd[tt][1] = (2 * t + 1 == g) ? gam : 0.0; }
#pragma unroll
  for (int o = 0; o < NO1; ++o) { hr[o][0] = hr[o][1] = hi[o][0] = hi[o][1] = 0.0; }
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (ks & KB) continue;
    double yr[TT], yi[TT];
#pragma unroll
    for (int tt = 0; tt < TT; ++tt) {
      const double ar = X.re[ks][tt], ai = X.im[ks][tt], br = X.re[ks | KB][tt], bi = X.im[ks | KB][tt];
      yr[tt] = fma(w1.x, br, fma(-w1.y, bi, fma(w0.x, ar, -w0.y * ai)));
      yi[tt] = fma(w1.x, bi, fma(w1.y, br, fma(w0.x, ai, w0.y * ar)));
    }
    // term by term over all tiles, so that two DMMAs into the same accumulator are several instructions apart
#pragma unroll
    for (int tt = 0; tt < TT; ++tt) qsb_dmma884(zd[tt][0], zd[tt][1], yr[tt], yr[tt] + yi[tt]);
    if (NOFF > 0) {
#pragma unroll
      for (int mt = 1; mt < TT; ++mt)
#pragma unroll
        for (int nt = 0; nt < mt; ++nt) { const int o = mt * (mt - 1) / 2 + nt; qsb_dmma884(hr[o][0], hr[o][1], yr[mt], yr[nt]); }
#pragma unroll
      for (int mt = 1; mt < TT; ++mt)
#pragma unroll
        for (int nt = 0; nt < mt; ++nt) { const int o = mt * (mt - 1) / 2 + nt; qsb_dmma884(hi[o][0], hi[o][1], yr[mt], yi[nt]); }
    }
#pragma unroll
    for (int tt = 0; tt < TT; ++tt) qsb_dmma884(zd[tt][0], zd[tt][1], yi[tt], yi[tt] - yr[tt]);
    if (NOFF > 0) {
#pragma unroll
      for (int mt = 1; mt < TT; ++mt)
#pragma unroll
        for (int nt = 0; nt < mt; ++nt) { const int o = mt * (mt - 1) / 2 + nt; qsb_dmma884(hr[o][0], hr[o][1], yi[mt], yi[nt]); }
#pragma unroll
      for (int mt = 1; mt < TT; ++mt)
#pragma unroll
        for (int nt = 0; nt < mt; ++nt) { const int o = mt * (mt - 1) / 2 + nt; qsb_dmma884(hi[o][0], hi[o][1], yi[mt], -yr[nt]); }
    }
  }
  const double r2 = 1.4142135623730951;
#pragma unroll
  for (int tt = 0; tt < TT; ++tt) *reinterpret_cast<double2*>(row + tt * 64 + 2 * lane) = make_double2(zd[tt][0], zd[tt][1]);
  if (NOFF > 0) {
#pragma unroll
    for (int o = 0; o < NOFF; ++o) {
      double* p = row + TT * 64 + o * 128 + 2 * lane;
      *reinterpret_cast<double2*>(p) = make_double2(r2 * hr[o][0], r2 * hr[o][1]);
      *reinterpret_cast<double2*>(p + 64) = make_double2(r2 * hi[o][0], r2 * hi[o][1]);
    }
  }
}

// One warp: forward sweep X <- g_j X over ops [0, je) starting from T^dagger, writing the packed Q_k of the parameters
// of ops [jb, je) to slab + k * K2.  The warp that reaches the last op also leaves V = U T^dagger in vout (N x N
// complex, natural row-major order; shared memory).
template <int NQ>
QSB_DEVICE void qsb_warp_qsweep(const QsbDevProgram& P, const double* target, const cplx* gtab, int lane, int jb, int je,
                                double* slab, cplx* vout) {
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, KS = G::KS, TT = G::TT, K2 = G::K2;
  const int g = lane >> 2, t = lane & 3;
  QsbFrag<KS, TT> X;
  QsbKLayout L;
  L.pos = L.log = 0x43210u;
  L.mru = -1; L.mru2 = -1;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks)
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      const int k = 4 * ks + t, c = 8 * o + g;
      const double2 tv = *reinterpret_cast<const double2*>(target + 2 * (c * N + k));
      X.re[ks][o] = tv.x; X.im[ks][o] = -tv.y;
    }
  for (int j = 0; j < je; ++j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) {
      const int kb = qsb_bring_in<NQ, TT, false>(X, X, L, NQ - 1 - op.q1, lane);
      qsb_cnot_slot<NQ, TT>(X, kb, qsb_nib(L.pos, NQ - 1 - op.q0), lane);
      continue;
    }
    const int kb = qsb_bring_in<NQ, TT, false>(X, X, L, NQ - 1 - op.q0, lane);
    const int np = qsb_num_params(op.opcode);
    if (j >= jb && np > 0) {
      for (int k = 0; k < np; ++k) {
        const cplx* slot = gtab + 4 * (op.reserved + 1 + k);
        double* row = slab + (size_t)(op.param_off + k) * K2;
        switch (kb) {
          case 1: qsb_q_param<NQ, 1>(X, slot, row, lane); break;
          case 2: if (NQ >= 4) qsb_q_param<NQ, (NQ >= 4 ? 2 : 1)>(X, slot, row, lane); break;
          default: if (NQ >= 5) qsb_q_param<NQ, (NQ >= 5 ? 4 : 1)>(X, slot, row, lane); break;
        }
      }
    }
    cplx m[4];
    qsb_warp_gate(op, gtab, 0, m);
    qsb_mix_slot<NQ, TT>(X, kb, m);
  }
  if (je == P.n_ops) {
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int kp = 4 * ks + t;
      int kl = 0;
#pragma unroll
      for (int b = 0; b < NQ; ++b) kl |= ((kp >> b) & 1) << qsb_nib(L.log, b);
#pragma unroll
      for (int o = 0; o < TT; ++o) vout[kl * N + 8 * o + g] = cmk(X.re[ks][o], X.im[ks][o]);
    }
  }
}

// row P of the slab: -C packed like a Q row, C = (V^dagger - V) / (2i); all threads of the CTA
template <int NQ>
QSB_DEVICE void qsb_q_crow(const cplx* V, double* row) {
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, TT = G::TT, K2 = G::K2;
  const double r2 = 1.4142135623730951;
  for (int idx = threadIdx.x; idx < K2; idx += blockDim.x) {
    int mt, nt, plane, rem;
    if (idx < TT * 64) { mt = nt = idx >> 6; plane = 2; rem = idx & 63; }
    else {
      const int i2 = idx - TT * 64, o = i2 >> 7;
      mt = 1;
      while ((mt + 1) * mt / 2 <= o) ++mt;
      nt = o - mt * (mt - 1) / 2;
      plane = (i2 >> 6) & 1; rem = i2 & 63;
    }
    const int lane = rem >> 1, e = rem & 1, m = 8 * mt + (lane >> 2), n = 8 * nt + 2 * (lane & 3) + e;
    const cplx a = V[m * N + n], b = V[n * N + m];
    const double cre = -0.5 * (a.y + b.y), cim = 0.5 * (a.x - b.x);
    row[idx] = -(plane == 2 ? cre + cim : r2 * (plane ? cim : cre));
  }
}

// CTA-wide: (cos, sin) table and body-frame gate table of x, then F = |U T^dagger - I|_F^2 / 2 and Tr(U T^dagger)
template <int NQ>
QSB_DEVICE double qsb_cta_half_norm_body(const QsbDevProgram& P, const double* x, const double* target, double* trig,
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
  qsb_fill_gates_body(P, trig, gtab, threadIdx.x, blockDim.x);
  __syncthreads();
  if (warp < NCB) {
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

// Q rows 0..P-1 and -C in row P of the slab (row-major [(P+1)][K2]); gtab = body-frame table of the current x.
// vsm: N x N complex of shared memory.  Ends with a block barrier: the slab is complete for every warp.
template <int NQ>
QSB_DEVICE void qsb_cta_qrows(const QsbDevProgram& P, const double* target, const cplx* gtab, const int* seg, cplx* vsm, double* slab) {
  using G = QsbQGeom<NQ>;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  const int jb = seg[warp], je = seg[warp + 1];
  if (jb < je || je == P.n_ops) qsb_warp_qsweep<NQ>(P, target, gtab, lane, jb, je, slab, vsm);
  __syncthreads();
  qsb_q_crow<NQ>(vsm, slab + (size_t)P.n_params * G::K2);
  __syncthreads();
}

// segments for the forward-only sweep: warp w pays for the whole prefix [0, seg[w + 1]), so later segments get fewer
// parameters.  A structured update of X costs about `mix` of a parameter's tensor work (2 TT k-step mixes vs the DMMAs).
QSB_DEVICE void qsb_cta_segments_fwd(const QsbDevProgram& P, int nseg, int* seg) {
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

template <int NQ>
QSB_GLOBAL __launch_bounds__(QsbQGeom<NQ>::WARPS * 32) qsb_k_lm3(QsbSolveArgs a) {
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, K2 = G::K2, NW = G::WARPS;
  QSB_DYN_SMEM(smem_raw);
  cplx* red = (cplx*)smem_raw;  // 32
  double* x = (double*)(red + 32);
  double* xn = x + a.max_params;
  double* g = xn + a.max_params;
  double* h = g + a.max_params;
  double* d2 = h + a.max_params;
  double* wk = d2 + a.max_params;
  double* trig = wk + a.max_params;         // 2 * maxP
  double* part = trig + 2 * a.max_params;   // 16
  cplx* gtab = (cplx*)(part + 16);  // [max_ops + max_params][4]; 16-byte aligned: 8 maxP + 16 doubles precede it
  int* seg = (int*)(gtab + (size_t)4 * (a.max_ops + a.max_params));  // NW + 1 (<= 12 ints), then [12] = queue slot, [13] = ok flag
  qsb_op* sops = (qsb_op*)(seg + 16);       // the program's ops, staged once per instance
  double* Ls = (double*)(sops + a.max_ops); // packed Cholesky factor; during the sweep the same bytes hold V
  Ls = (double*)(((uintptr_t)Ls + 15) & ~(uintptr_t)15);
  cplx* vsm = (cplx*)Ls;
  const size_t tri = (size_t)(a.max_params + 1) * (a.max_params + 2) / 2;
  double* Winv = Ls + (((tri > (size_t)2 * N * N ? tri : (size_t)2 * N * N) + 1) & ~(size_t)1);  // [ceil(maxP / 16)][256]
  double* zs = Winv + (size_t)((a.max_params + QSB_CH_NB - 1) / QSB_CH_NB) * QSB_CH3_WSZ;  // 32
  int* okflag = seg + 13;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  double* Rg = a.scratch + (size_t)blockIdx.x * a.scratch_per_cta;
  double* A = Rg + (size_t)(a.max_params + 1) * K2;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) seg[12] = atomicAdd(&a.ctl[QSB_CTL_NEXT], 1);
    __syncthreads();
    const int inst = seg[12];
    if (inst >= a.count) break;
    const int pidx = a.prog_idx[inst];
    QsbDevProgram prog = a.progs[pidx];
    const int P = prog.n_params, rows = P + 1;
    __syncthreads();
    for (int e = threadIdx.x; e < prog.n_ops * 6; e += blockDim.x) ((int32_t*)sops)[e] = ((const int32_t*)prog.ops)[e];
    prog.ops = sops;  // every sweep of this solve reads the ops from shared memory
    const bool kzero = qsb_syrk_needs_zero(rows, NW);
    const double* x0 = a.x0 + a.param_off[inst];
    const int max_iters = a.max_iters > 0 ? a.max_iters : 100 * (P > 0 ? P : 1);
    for (int k = threadIdx.x; k < P; k += blockDim.x) x[k] = x0[k];
    qsb_cta_segments_fwd(prog, NW, seg);
    int iters = 0, status = 0;
    cplx tr;
    double F = qsb_cta_half_norm_body<NQ>(prog, x, a.target, trig, gtab, part, &tr);  // also fills trig / gtab for x
    if (P > 0) {
      { QSB_PROF_BEGIN(); qsb_cta_qrows<NQ>(prog, a.target, gtab, seg, vsm, Rg); QSB_PROF_END(0); }
      iters = 1;
      {
        QSB_PROF_BEGIN();
        if (kzero) { for (int e = threadIdx.x; e < rows * (rows + 1) / 2; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
        qsb_syrk_packed(Rg, rows, K2, A, warp, NW, lane);
        __syncthreads();
        QSB_PROF_END(1);
      }
      for (int k = threadIdx.x; k < P; k += blockDim.x) {
        g[k] = QSB_LDCG(&A[QSB_TRI(P, k)]);
        const double akk = QSB_LDCG(&A[QSB_TRI(k, k)]);
        d2[k] = akk > 0.0 ? akk : 1.0;
      }
      __syncthreads();
      double mu = 1e-3, nu = 2.0;
      int rejects = 0;
      while (true) {
        if (!(F > 1e-300) || (a.stop_f > 0.0 && F <= a.stop_f)) { status = 4; break; }
        {  // MINPACK's gtol: max cosine between r and the columns of J_r; the batch-level stop rules ride on the same reduction
          double gm = 0.0;
          const double rn = sqrt(2.0 * F);
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            const double akk = QSB_LDCG(&A[QSB_TRI(k, k)]);
            if (akk > 0.0) gm = fmax(gm, fabs(g[k]) / (sqrt(akk) * rn));
          }
          if (threadIdx.x == 0) {  // encoded as huge values: every thread sees the same maximum
            if (a.sibling_f > 0.0 && QSB_LDCG(&a.ctl[QSB_CTL_SOLVED + pidx]) != 0) gm = 1e301;
            else if (a.straggler_c > 0.0 && 2 * QSB_LDCG(&a.ctl[QSB_CTL_FINISHED]) >= a.count &&
                     (double)iters > a.straggler_c * (double)QSB_LDCG(&a.ctl[QSB_CTL_MEDIAN]) && iters >= a.straggler_min) gm = 1e300;
          }
          gm = qsb_block_max(gm, red);
          if (gm >= 1e301) { status = 7; break; }   // another start of this circuit already reached the target
          if (gm >= 1e300) { status = 6; break; }   // straggler cap
          if (gm <= a.gtol) { status = 1; break; }
        }
        bool chol_ok;
        { QSB_PROF_BEGIN(); chol_ok = qsb_chol3(A, rows, d2, mu, P, Ls, Winv, wk, okflag, a.prof); QSB_PROF_END(2); }
        if (!chol_ok) {
          mu *= nu; nu *= 2.0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
          continue;
        }
        { QSB_PROF_BEGIN(); qsb_ch3_back(Ls, P, Winv, zs, h); QSB_PROF_END(3); }
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
        { QSB_PROF_BEGIN(); Fn = qsb_cta_half_norm_body<NQ>(prog, xn, a.target, trig, gtab, part, &tr); QSB_PROF_END(4); }  // tables now belong to xn
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
          { QSB_PROF_BEGIN(); qsb_cta_qrows<NQ>(prog, a.target, gtab, seg, vsm, Rg); QSB_PROF_END(0); }
          ++iters;
          {
            QSB_PROF_BEGIN();
            if (kzero) { for (int e = threadIdx.x; e < rows * (rows + 1) / 2; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
            qsb_syrk_packed(Rg, rows, K2, A, warp, NW, lane);
            __syncthreads();
            QSB_PROF_END(1);
          }
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            g[k] = QSB_LDCG(&A[QSB_TRI(P, k)]);
            d2[k] = fmax(d2[k], QSB_LDCG(&A[QSB_TRI(k, k)]));
          }
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
    qsb_cta_half_norm_body<NQ>(prog, x, a.target, trig, gtab, part, &tr);
    double* ox = a.out_x + a.param_off[inst];
    for (int k = threadIdx.x; k < P; k += blockDim.x) ox[k] = x[k];
    if (a.prof && threadIdx.x == 0) atomicAdd((unsigned long long*)&a.prof[5], (unsigned long long)iters);
    if (threadIdx.x == 0) {
      const double dist = fabs(1.0 - sqrt(tr.x * tr.x + tr.y * tr.y) / N);
      a.out_f[inst] = dist;
      a.out_iters[inst] = iters;
      a.out_status[inst] = status;
      // batch-level bookkeeping for the stop rules of the instances still running
      if (a.sibling_f > 0.0 && dist < a.sibling_f) atomicAdd(&a.ctl[QSB_CTL_SOLVED + pidx], 1);
      if (status != 6 && status != 7) {
        const int nfin = atomicAdd(&a.ctl[QSB_CTL_FINISHED], 1) + 1;
        if (2 * nfin <= a.count + 1) atomicMax(&a.ctl[QSB_CTL_MEDIAN], iters);
      }
    }
  }
}

static inline size_t qsb_lm3_smem_bytes(int nq, int max_ops, int max_params) {
  const size_t N = (size_t)1 << nq;
  const size_t base = 32 * 16 + (size_t)8 * max_params * 8 + 16 * 8 + 8 + (size_t)64 * (max_ops + max_params) + 16 * 4 + (size_t)24 * max_ops + 16;
  const size_t packed = std::max((size_t)(max_params + 1) * (max_params + 2) / 2, 2 * N * N);
  return base + (packed + 1 + (size_t)((max_params + QSB_CH_NB - 1) / QSB_CH_NB) * QSB_CH3_WSZ + 32) * 8;
}

// ---------------------------------------------------------------------------- QSB_RUN_LMSYS, unitary target
// r^T r, g = J_r^T r and A = J_r^T J_r of every instance of a batch in the body-frame form above: a forward-only sweep
// and a Gram product over rows of N^2 (not 2 N^2) reals; nothing of size N^2 P leaves the CTA's L2-resident slab.
// One CTA per instance.  out_a[inst] = r^T r, out_b[param_off ..] = g, out_c[sum P_i^2 ..] = A (row-major P x P).
template <int NQ>
QSB_GLOBAL __launch_bounds__(QsbQGeom<NQ>::WARPS * 32) qsb_k_lmsys3(QsbBatchArgs a, double* scratch, size_t scratch_per_cta,
                                                                    const long long* sq_off) {
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, K2 = G::K2, NW = G::WARPS;
  QSB_DYN_SMEM(smem_raw);
  double* trig = (double*)smem_raw;                       // 2 * maxP
  double* part = trig + 2 * a.max_params;                 // 16
  cplx* gtab = (cplx*)(part + 16);
  cplx* vsm = gtab + (size_t)4 * (a.max_ops + a.max_params);  // N * N
  int* seg = (int*)(vsm + N * N);                         // 16
  qsb_op* sops = (qsb_op*)(seg + 16);
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  double* Rg = scratch + (size_t)blockIdx.x * scratch_per_cta;
  double* A = Rg + (size_t)(a.max_params + 1) * K2;
  const long long p_base = a.param_off[a.first], s_base = sq_off[a.first];
  int staged = -1;
  for (int li = blockIdx.x; li < a.count; li += gridDim.x) {
    const int inst = a.first + li;
    const int pi = a.prog_idx[inst];
    QsbDevProgram prog = a.progs[pi];
    const int P = prog.n_params, rows = P + 1;
    __syncthreads();
    if (pi != staged) {
      for (int e = threadIdx.x; e < prog.n_ops * 6; e += blockDim.x) ((int32_t*)sops)[e] = ((const int32_t*)prog.ops)[e];
      staged = pi;
    }
    prog.ops = sops;
    __syncthreads();
    qsb_cta_segments_fwd(prog, NW, seg);
    const double F = qsb_cta_half_norm_body<NQ>(prog, a.params + a.param_off[inst], a.target, trig, gtab, part, nullptr);
    qsb_cta_qrows<NQ>(prog, a.target, gtab, seg, vsm, Rg);
    if (qsb_syrk_needs_zero(rows, NW)) { for (int e = threadIdx.x; e < rows * (rows + 1) / 2; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
    qsb_syrk_packed(Rg, rows, K2, A, warp, NW, lane);
    __syncthreads();
    double* og = a.out_b + (a.param_off[inst] - p_base);
    double* oA = a.out_c + (sq_off[inst] - s_base);
    for (int k = threadIdx.x; k < P; k += blockDim.x) og[k] = QSB_LDCG(&A[QSB_TRI(P, k)]);
    for (int e = threadIdx.x; e < P * P; e += blockDim.x) {
      const int k = e / P, l = e - k * P;
      oA[e] = QSB_LDCG(&A[k >= l ? QSB_TRI(k, l) : QSB_TRI(l, k)]);
    }
    if (threadIdx.x == 0) a.out_a[li] = 2.0 * F;
  }
}

static inline size_t qsb_lmsys3_smem_bytes(int nq, int max_ops, int max_params) {
  const size_t N = (size_t)1 << nq;
  return ((size_t)2 * max_params + 16 + 8 * (size_t)(max_ops + max_params) + 2 * N * N) * 8 + 16 * 4 + (size_t)24 * max_ops + 64;
}
