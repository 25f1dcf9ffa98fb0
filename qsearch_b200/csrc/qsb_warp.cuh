// Register-resident warp kernels for the standard search gatesets (one-qubit leaves + CNOT), n = 3, 4, 5 qubits.
//
// One WARP owns one 16x16 (n = 3: 8x8) complex tile of every output matrix of one circuit instance: the whole matrix
// for n <= 4, one quadrant for n = 5 (four independent warps per instance, no communication between them).  The
// running prefix X (= R_j T^dagger), the running suffix S and the accumulators live in registers in the fragment
// layouts of the FP64 tensor-core instruction mma.sync.m8n8k4.f64 (DMMA):
//
//     lane = 4 g + t,  g = lane >> 2 (0..7),  t = lane & 3
//     X[ks][nt]  = X[row 4 ks + t][col c0 + 8 nt + g]       B fragment (k x n), the warp's column block c0
//     S[ks][mt]  = S[row r0 + 8 mt + g][col 4 ks + t]       A fragment (m x k), the warp's row block r0
//     K[mt][nt]  = K[row r0 + 8 mt + g][col c0 + 8 nt + 2 t + {0,1}]          C fragment
//
// The contracted index k = 4 ks + t carries the qubit that a one-qubit gate acts on: bit (n-1-q) of k.  Bits >= 2 are
// k-step bits (partner element in the same thread), bits 0..1 are lane bits (partner one __shfl_xor away), so every
// structured update of X (left multiplication, rows) and of S (right multiplication, columns) is thread-local or a
// warp shuffle: no shared memory tiles, no block barrier.  The dense products K_k = S (dg_k (x) I) X
// (qsearch/gates.py:993-995 with the T^dagger of comparison.py:70 folded into X) run on DMMA: 4 real m8n8k4
// products per complex tile product, Y = (dg_k (x) I) X generated per k-step on the fly.
//
// Algorithm = the same peel sweep as qsb_generic.cuh (restating ProductGate.mat_jac, qsearch/gates.py:965-1004).
#pragma once
#include "qsb_generic.cuh"

#ifdef QSB_EMU
// emulated DMMA (tests only): D[g][2t+j] += sum_k A[g][k] B[k][2t+j];  A[g][k] sits in lane 4g+k, B[k][n] in lane 4n+k
static inline void qsb_dmma884(double& d0, double& d1, double a, double b) {
  auto* st = qsb_emu::cur();
  const int w = threadIdx.x / 32, l = threadIdx.x % 32, g = l >> 2, t = l & 3;
  uint64_t ra, rb;
  std::memcpy(&ra, &a, 8); std::memcpy(&rb, &b, 8);
  st->warp_buf[w][l] = ra; st->warp_buf2[w][l] = rb;
  __syncwarp();
  for (int k = 0; k < 4; ++k) {
    double av, b0, b1;
    std::memcpy(&av, &st->warp_buf[w][4 * g + k], 8);
    std::memcpy(&b0, &st->warp_buf2[w][4 * (2 * t) + k], 8);
    std::memcpy(&b1, &st->warp_buf2[w][4 * (2 * t + 1) + k], 8);
    d0 = std::fma(av, b0, d0); d1 = std::fma(av, b1, d1);
  }
  __syncwarp();
}
#else
__device__ __forceinline__ void qsb_dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
#endif

template <int NQ>
struct QsbWarpGeom {
  static constexpr int N = 1 << NQ;
  static constexpr int NN = N * N;
  static constexpr int KS = N / 4;                 // k-steps of the m8n8k4 instruction
  static constexpr int TT = (NQ >= 4) ? 2 : 1;     // 8-wide tiles per warp along rows and along columns
  static constexpr int WR = N / (8 * TT);          // warps along one side of the matrix
  static constexpr int WARPS = WR * WR;            // warps per instance
};

template <int KS, int TT>
struct QsbFrag {
  double re[KS][TT], im[KS][TT];
};

#define QSB_FULL 0xffffffffu

// dst[k] = m[b][0] src[k with bit = 0] + m[b][1] src[k with bit = 1], b = that bit of k; the bit belongs to qubit q.
// m = {m00, m01, m10, m11}.  In-thread variant: partner k-step = ks ^ KB.
template <int KS, int TT, int KB>
QSB_DEVICE void qsb_mix_thread(QsbFrag<KS, TT>& F, const cplx* m) {
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (ks & KB) continue;
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      const double ar = F.re[ks][o], ai = F.im[ks][o], br = F.re[ks | KB][o], bi = F.im[ks | KB][o];
      F.re[ks][o] = fma(m[1].x, br, fma(-m[1].y, bi, fma(m[0].x, ar, -m[0].y * ai)));
      F.im[ks][o] = fma(m[1].x, bi, fma(m[1].y, br, fma(m[0].x, ai, m[0].y * ar)));
      F.re[ks | KB][o] = fma(m[3].x, br, fma(-m[3].y, bi, fma(m[2].x, ar, -m[2].y * ai)));
      F.im[ks | KB][o] = fma(m[3].x, bi, fma(m[3].y, br, fma(m[2].x, ai, m[2].y * ar)));
    }
  }
}

template <int KS, int TT>
QSB_DEVICE void qsb_mix_lane(QsbFrag<KS, TT>& F, const cplx* m, int lbit, int lane) {
  const bool hi = (lane >> lbit) & 1;
  const cplx ca = hi ? m[3] : m[0], cb = hi ? m[2] : m[1];
  const int mask = 1 << lbit;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks)
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      const double ar = F.re[ks][o], ai = F.im[ks][o];
      const double pr = __shfl_xor_sync(QSB_FULL, ar, mask), pi = __shfl_xor_sync(QSB_FULL, ai, mask);
      F.re[ks][o] = fma(cb.x, pr, fma(-cb.y, pi, fma(ca.x, ar, -ca.y * ai)));
      F.im[ks][o] = fma(cb.x, pi, fma(cb.y, pr, fma(ca.x, ai, ca.y * ar)));
    }
}

template <int NQ, int TT>
QSB_DEVICE void qsb_mix(QsbFrag<(1 << NQ) / 4, TT>& F, int q, const cplx* m, int lane) {
  constexpr int KS = (1 << NQ) / 4;
  const int bit = NQ - 1 - q;
  switch (bit) {
    case 0: case 1: qsb_mix_lane<KS, TT>(F, m, bit, lane); break;
    case 2: qsb_mix_thread<KS, TT, 1>(F, m); break;
    case 3: if (NQ >= 4) qsb_mix_thread<KS, TT, (NQ >= 4 ? 2 : 1)>(F, m); break;
    default: if (NQ >= 5) qsb_mix_thread<KS, TT, (NQ >= 5 ? 4 : 1)>(F, m); break;
  }
}

// CNOT(control c, target tq) as a permutation of the k index: F'[k] = F[k ^ tbit] where the control bit of k is set.
template <int KS, int TT, int KB>
QSB_DEVICE void qsb_cnot_thread(QsbFrag<KS, TT>& F, int cbit, int lane) {
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (ks & KB) continue;
    const bool ctrl = (((4 * ks) | (lane & 3)) >> cbit) & 1;
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      const double ar = F.re[ks][o], ai = F.im[ks][o], br = F.re[ks | KB][o], bi = F.im[ks | KB][o];
      F.re[ks][o] = ctrl ? br : ar; F.im[ks][o] = ctrl ? bi : ai;
      F.re[ks | KB][o] = ctrl ? ar : br; F.im[ks | KB][o] = ctrl ? ai : bi;
    }
  }
}

template <int KS, int TT>
QSB_DEVICE void qsb_cnot_lane(QsbFrag<KS, TT>& F, int cbit, int tbit, int lane) {
  const int mask = 1 << tbit;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const bool ctrl = (((4 * ks) | (lane & 3)) >> cbit) & 1;
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      const double pr = __shfl_xor_sync(QSB_FULL, F.re[ks][o], mask), pi = __shfl_xor_sync(QSB_FULL, F.im[ks][o], mask);
      if (ctrl) { F.re[ks][o] = pr; F.im[ks][o] = pi; }
    }
  }
}

template <int NQ, int TT>
QSB_DEVICE void qsb_cnot(QsbFrag<(1 << NQ) / 4, TT>& F, int c, int tq, int lane) {
  constexpr int KS = (1 << NQ) / 4;
  const int cbit = NQ - 1 - c, tbit = NQ - 1 - tq;
  switch (tbit) {
    case 0: case 1: qsb_cnot_lane<KS, TT>(F, cbit, tbit, lane); break;
    case 2: qsb_cnot_thread<KS, TT, 1>(F, cbit, lane); break;
    case 3: if (NQ >= 4) qsb_cnot_thread<KS, TT, (NQ >= 4 ? 2 : 1)>(F, cbit, lane); break;
    default: if (NQ >= 5) qsb_cnot_thread<KS, TT, (NQ >= 5 ? 4 : 1)>(F, cbit, lane); break;
  }
}

// K = S (m (x) I) X on the tensor cores.  KB > 0: the qubit of m is k-step bit KB; KB == 0: lane bit lbit.
template <int KS, int TT, int KB>
QSB_DEVICE void qsb_dmma_product(double (&cr)[TT][TT][2], double (&ci)[TT][TT][2], const QsbFrag<KS, TT>& S,
                                 const QsbFrag<KS, TT>& X, const cplx* m, int lbit, int lane) {
#pragma unroll
  for (int a = 0; a < TT; ++a)
#pragma unroll
    for (int b = 0; b < TT; ++b) { cr[a][b][0] = cr[a][b][1] = ci[a][b][0] = ci[a][b][1] = 0.0; }
  const bool hi = (lane >> lbit) & 1;
  const int mask = 1 << lbit;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    double yr[TT], yi[TT], ny[TT];
#pragma unroll
    for (int nt = 0; nt < TT; ++nt) {
      cplx ca, cb;
      double pr, pi;
      const double xr = X.re[ks][nt], xi = X.im[ks][nt];
      if (KB > 0) {
        const bool up = (ks & KB) != 0;
        ca = up ? m[3] : m[0]; cb = up ? m[2] : m[1];
        pr = X.re[ks ^ KB][nt]; pi = X.im[ks ^ KB][nt];
      } else {
        ca = hi ? m[3] : m[0]; cb = hi ? m[2] : m[1];
        pr = __shfl_xor_sync(QSB_FULL, xr, mask); pi = __shfl_xor_sync(QSB_FULL, xi, mask);
      }
      yr[nt] = fma(cb.x, pr, fma(-cb.y, pi, fma(ca.x, xr, -ca.y * xi)));
      yi[nt] = fma(cb.x, pi, fma(cb.y, pr, fma(ca.x, xi, ca.y * xr)));
      ny[nt] = -yi[nt];
    }
    // term by term over all tiles: two DMMAs into the same accumulator are 2 TT^2 instructions apart
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], S.re[ks][mt], yr[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], S.re[ks][mt], yi[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], S.im[ks][mt], ny[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], S.im[ks][mt], yr[nt]);
  }
}

template <int NQ, int TT>
QSB_DEVICE void qsb_product(double (&cr)[TT][TT][2], double (&ci)[TT][TT][2], const QsbFrag<(1 << NQ) / 4, TT>& S,
                            const QsbFrag<(1 << NQ) / 4, TT>& X, int q, const cplx* m, int lane) {
  constexpr int KS = (1 << NQ) / 4;
  const int bit = NQ - 1 - q;
  switch (bit) {
    case 0: case 1: qsb_dmma_product<KS, TT, 0>(cr, ci, S, X, m, bit, lane); break;
    case 2: qsb_dmma_product<KS, TT, 1>(cr, ci, S, X, m, 0, lane); break;
    case 3: if (NQ >= 4) qsb_dmma_product<KS, TT, (NQ >= 4 ? 2 : 1)>(cr, ci, S, X, m, 0, lane); break;
    default: if (NQ >= 5) qsb_dmma_product<KS, TT, (NQ >= 5 ? 4 : 1)>(cr, ci, S, X, m, 0, lane); break;
  }
}

QSB_DEVICE void qsb_dagger2x2(cplx* g) {
  const cplx t = g[1];
  g[0] = cconj(g[0]); g[1] = cconj(g[2]); g[2] = cconj(t); g[3] = cconj(g[3]);
}
QSB_DEVICE void qsb_transpose2x2(cplx* g) { const cplx t = g[1]; g[1] = g[2]; g[2] = t; }

// ---------------------------------------------------------------------------- rank-one form of the derivatives
// Every parameter derivative of the one-qubit leaves is  dg = c g + a b^T  with c = -i/2 (generator-type parameters:
// g^dagger dg has the eigenvalues +-i/2) or c = 0 (phase-type parameters: dg is one row or one column of g times i).
// Then  K = S (dg (x) I) X = c V + (S a)(b^T X),  V = S (g (x) I) X = U T^dagger the same for every parameter, and
// (S a)[:, kappa] = a0 S[:, k0] + a1 S[:, k1],  (b^T X)[kappa, :] = b0 X[k0, :] + b1 X[k1, :]  contract over HALF
// the index range (kappa = k without the bit of the gate's qubit): half the tensor-core work of a dense product.
QSB_DEVICE bool qsb_deriv_has_v(int opcode, int k) {
  switch (opcode) {
    case QSB_OP_RX: case QSB_OP_RY: case QSB_OP_RZ: case QSB_OP_XZXZ: case QSB_OP_ZXZXZ: return true;
    case QSB_OP_U3: return k == 0;
    default: return false;  // U1, U2, U3 phi / lambda
  }
}

// dg -> {rho, (row, has_v), b0, b1}: the remainder R = dg - c g is rank one up to rounding; factored around its largest
// entry it is a b^T with b^T = the pivot row of R and a = (1, rho) (pivot row 0: row = 0) or (rho, 1) (row = 1)
QSB_DEVICE void qsb_rank_one(const cplx* g, bool has_v, cplx* d) {
  cplx r[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) r[i] = has_v ? cmk(d[i].x - 0.5 * g[i].y, d[i].y + 0.5 * g[i].x) : d[i];  // R = dg + (i/2) g
  int piv = 0;
  double best = -1.0;
  cplx pv = r[0];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double m = r[i].x * r[i].x + r[i].y * r[i].y;
    if (m > best) { best = m; piv = i; pv = r[i]; }
  }
  const bool row1 = piv >= 2;                                                  // pivot row
  const cplx ot = piv == 0 ? r[2] : piv == 1 ? r[3] : piv == 2 ? r[0] : r[1];  // same column, other row
  const double inv = best > 0.0 ? 1.0 / best : 0.0;
  const cplx ratio = cmk((ot.x * pv.x + ot.y * pv.y) * inv, (ot.y * pv.x - ot.x * pv.y) * inv);  // ot / pv
  const cplx b0 = row1 ? r[2] : r[0], b1 = row1 ? r[3] : r[1];
  d[0] = ratio; d[1] = cmk(row1 ? 1.0 : 0.0, has_v ? 1.0 : 0.0);
  d[2] = b0; d[3] = b1;
}

// Gate table: the 2x2 of every one-qubit op and of every parameter derivative of one instance, built once per
// parameter vector by all cooperating threads (one (op, which) slot per thread and round) and then read as broadcast
// shared-memory loads, instead of every lane rebuilding the same entries from the sin/cos table at every use.
// rank1: derivative slots hold the rank-one form above (qsb_rank_one) instead of the 2x2 itself.
QSB_DEVICE void qsb_fill_gates(const QsbDevProgram& P, const double* trig, cplx* gtab, int tid, int nthreads,
                               bool rank1 = false) {
  const int n_slots = P.n_ops + P.n_params;
  for (int s = tid; s < n_slots; s += nthreads) {
    const int code = P.slotmap[s];
    if (code < 0) continue;
    const qsb_op op = P.ops[code >> 2];
    cplx g[4];
    if (op.opcode == QSB_OP_CONST1) {
      const double* c = P.consts + op.const_off;
      g[0] = cmk(c[0], c[1]); g[1] = cmk(c[2], c[3]); g[2] = cmk(c[4], c[5]); g[3] = cmk(c[6], c[7]);
    } else {
      qsb_gate2x2(op.opcode, trig + 2 * op.param_off, code & 3, g);
      if (rank1 && (code & 3) != 0) {
        cplx g0[4];
        qsb_gate2x2(op.opcode, trig + 2 * op.param_off, 0, g0);
        qsb_rank_one(g0, qsb_deriv_has_v(op.opcode, (code & 3) - 1), g);
      }
    }
    gtab[4 * s] = g[0]; gtab[4 * s + 1] = g[1]; gtab[4 * s + 2] = g[2]; gtab[4 * s + 3] = g[3];
  }
}

// which = 0: the gate of op; which = 1 + k: its derivative with respect to its k-th parameter
QSB_DEVICE void qsb_warp_gate(const qsb_op& op, const cplx* gtab, int which, cplx* g) {
  const cplx* t = gtab + 4 * (op.reserved + which);
  g[0] = t[0]; g[1] = t[1]; g[2] = t[2]; g[3] = t[3];
}

// ---------------------------------------------------------------------------- dynamic layout of the contracted index
// The contracted index k only has to be laid out the SAME way in X (rows) and S (columns), so the kernel is free to
// move the bit of the qubit it is about to work on from a lane position (0..1) into a k-step position (>= 2), where the
// partner element is in the same thread.  One exchange = one __shfl_xor of half the fragment; afterwards every mix, CNOT
// and half-size product on that qubit is thread-local.  pos/log: nibble l of pos = physical bit of logical bit l,
// nibble p of log = logical bit at physical bit p; mru/mru2: the k-step positions used last (not evicted).
struct QsbKLayout {
  unsigned pos, log;
  int mru, mru2;
};
QSB_DEVICE int qsb_nib(unsigned w, int i) { return (int)((w >> (4 * i)) & 7u); }
QSB_DEVICE unsigned qsb_nib_set(unsigned w, int i, int v) { return (w & ~(15u << (4 * i))) | ((unsigned)v << (4 * i)); }

// exchange lane bit lbit with k-step bit KB of a fragment
template <int KS, int TT, int KB>
QSB_DEVICE void qsb_swap_bits(QsbFrag<KS, TT>& F, int lbit, int lane) {
  const bool hi = (lane >> lbit) & 1;
  const int mask = 1 << lbit;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (ks & KB) continue;
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      const double sr = hi ? F.re[ks][o] : F.re[ks | KB][o], si = hi ? F.im[ks][o] : F.im[ks | KB][o];
      const double rr = __shfl_xor_sync(QSB_FULL, sr, mask), ri = __shfl_xor_sync(QSB_FULL, si, mask);
      if (hi) { F.re[ks][o] = rr; F.im[ks][o] = ri; }
      else { F.re[ks | KB][o] = rr; F.im[ks | KB][o] = ri; }
    }
  }
}

// make the logical bit lb a k-step bit of both fragments; returns its k-step mask (1, 2 or 4)
template <int NQ, int TT, bool WITH_S>
QSB_DEVICE int qsb_bring_in(QsbFrag<(1 << NQ) / 4, TT>& X, QsbFrag<(1 << NQ) / 4, TT>& S, QsbKLayout& L, int lb, int lane) {
  constexpr int KS = (1 << NQ) / 4;
  int p = qsb_nib(L.pos, lb);
  if (p < 2) {
    int v = 2;
    if (NQ == 4) v = (L.mru == 2) ? 3 : 2;
    if (NQ == 5) { while (v == L.mru || v == L.mru2) ++v; }
    switch (v) {
      case 2: qsb_swap_bits<KS, TT, 1>(X, p, lane); if (WITH_S) qsb_swap_bits<KS, TT, 1>(S, p, lane); break;
      case 3: if (NQ >= 4) { qsb_swap_bits<KS, TT, (NQ >= 4 ? 2 : 1)>(X, p, lane); if (WITH_S) qsb_swap_bits<KS, TT, (NQ >= 4 ? 2 : 1)>(S, p, lane); } break;
      default: if (NQ >= 5) { qsb_swap_bits<KS, TT, (NQ >= 5 ? 4 : 1)>(X, p, lane); if (WITH_S) qsb_swap_bits<KS, TT, (NQ >= 5 ? 4 : 1)>(S, p, lane); } break;
    }
    const int lv = qsb_nib(L.log, v);
    L.pos = qsb_nib_set(qsb_nib_set(L.pos, lb, v), lv, p);
    L.log = qsb_nib_set(qsb_nib_set(L.log, v, lb), p, lv);
    p = v;
  }
  if (p != L.mru) { L.mru2 = L.mru; L.mru = p; }
  return 1 << (p - 2);
}

template <int NQ, int TT>
QSB_DEVICE void qsb_mix_slot(QsbFrag<(1 << NQ) / 4, TT>& F, int kb, const cplx* m) {
  constexpr int KS = (1 << NQ) / 4;
  switch (kb) {
    case 1: qsb_mix_thread<KS, TT, 1>(F, m); break;
    case 2: if (NQ >= 4) qsb_mix_thread<KS, TT, (NQ >= 4 ? 2 : 1)>(F, m); break;
    default: if (NQ >= 5) qsb_mix_thread<KS, TT, (NQ >= 5 ? 4 : 1)>(F, m); break;
  }
}

// CNOT whose target bit sits at k-step mask kb; the control bit at physical position cp (lane or k-step bit)
template <int NQ, int TT>
QSB_DEVICE void qsb_cnot_slot(QsbFrag<(1 << NQ) / 4, TT>& F, int kb, int cp, int lane) {
  constexpr int KS = (1 << NQ) / 4;
  switch (kb) {
    case 1: qsb_cnot_thread<KS, TT, 1>(F, cp, lane); break;
    case 2: if (NQ >= 4) qsb_cnot_thread<KS, TT, (NQ >= 4 ? 2 : 1)>(F, cp, lane); break;
    default: if (NQ >= 5) qsb_cnot_thread<KS, TT, (NQ >= 5 ? 4 : 1)>(F, cp, lane); break;
  }
}

// acc += (S a)(b^T X) over the half index range; ab = {rho, (row, has_v), b0, b1} as written by qsb_rank_one
template <int KS, int TT, int KB>
QSB_DEVICE void qsb_dmma_half(double (&cr)[TT][TT][2], double (&ci)[TT][TT][2], const QsbFrag<KS, TT>& S,
                              const QsbFrag<KS, TT>& X, const cplx* ab) {
  const cplx rho = ab[0], b0 = ab[2], b1 = ab[3];
  const bool row1 = ab[1].x != 0.0;  // the same for the whole warp
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (ks & KB) continue;
    double sr[TT], si[TT], yr[TT], yi[TT];
    if (row1) {  // a = (rho, 1)
#pragma unroll
      for (int o = 0; o < TT; ++o) {
        sr[o] = fma(rho.x, S.re[ks][o], fma(-rho.y, S.im[ks][o], S.re[ks | KB][o]));
        si[o] = fma(rho.x, S.im[ks][o], fma(rho.y, S.re[ks][o], S.im[ks | KB][o]));
      }
    } else {  // a = (1, rho)
#pragma unroll
      for (int o = 0; o < TT; ++o) {
        sr[o] = fma(rho.x, S.re[ks | KB][o], fma(-rho.y, S.im[ks | KB][o], S.re[ks][o]));
        si[o] = fma(rho.x, S.im[ks | KB][o], fma(rho.y, S.re[ks | KB][o], S.im[ks][o]));
      }
    }
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      yr[o] = fma(b1.x, X.re[ks | KB][o], fma(-b1.y, X.im[ks | KB][o], fma(b0.x, X.re[ks][o], -b0.y * X.im[ks][o])));
      yi[o] = fma(b1.x, X.im[ks | KB][o], fma(b1.y, X.re[ks | KB][o], fma(b0.x, X.im[ks][o], b0.y * X.re[ks][o])));
    }
    // term by term over all tiles: two DMMAs into the same accumulator are 2 TT^2 instructions apart
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], sr[mt], yr[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], sr[mt], yi[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], si[mt], -yi[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], si[mt], yr[nt]);
  }
}

#ifdef QSB_EMU
static inline double qsb_neg(double x) { return -x; }
#else
// sign flip on the integer pipe (the FP64 pipe is the busy one)
__device__ __forceinline__ double qsb_neg(double x) { return __hiloint2double(__double2hiint(x) ^ (int)0x80000000, __double2loint(x)); }
#endif

// Projector-type derivatives need no factor arithmetic.  With P1 = |1><1| on the gate's qubit:
//   output side  dg = c g + i P1 g   (U3 phi: c = 0; XZXZ v1: c = -i/2, since Rz' = -i/2 Z Rz and Z = I - 2 P1)
//                                     ->  K = c V + i S[:, k1] (g X)[k1, :]     with X BEFORE the peel
//   input side   dg = c g + i g P1   (U3 lambda: c = 0)
//                                     ->  K = c V + i (S g)[:, k1] X[k1, :]     with S AFTER its update
// acc += i S[:, k1] X[k1, :] over the half index range (k1 = the k with the gate's bit set)
template <int KS, int TT, int KB>
QSB_DEVICE void qsb_dmma_half_raw(double (&cr)[TT][TT][2], double (&ci)[TT][TT][2], const QsbFrag<KS, TT>& S,
                                  const QsbFrag<KS, TT>& X) {
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (!(ks & KB)) continue;
    double nxr[TT], nxi[TT];
#pragma unroll
    for (int o = 0; o < TT; ++o) { nxr[o] = qsb_neg(X.re[ks][o]); nxi[o] = qsb_neg(X.im[ks][o]); }
    // (sr + i si) i (xr + i xi):  re = -sr xi - si xr,  im = sr xr - si xi
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], S.re[ks][mt], nxi[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], S.re[ks][mt], X.re[ks][nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], S.im[ks][mt], nxr[nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], S.im[ks][mt], nxi[nt]);
  }
}

// V = S X (dense, all k-steps)
template <int KS, int TT>
QSB_DEVICE void qsb_dmma_full(double (&cr)[TT][TT][2], double (&ci)[TT][TT][2], const QsbFrag<KS, TT>& S,
                              const QsbFrag<KS, TT>& X) {
#pragma unroll
  for (int a = 0; a < TT; ++a)
#pragma unroll
    for (int b = 0; b < TT; ++b) { cr[a][b][0] = cr[a][b][1] = ci[a][b][0] = ci[a][b][1] = 0.0; }
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], S.re[ks][mt], X.re[ks][nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], S.re[ks][mt], X.im[ks][nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(cr[mt][nt][0], cr[mt][nt][1], S.im[ks][mt], -X.im[ks][nt]);
#pragma unroll
    for (int nt = 0; nt < TT; ++nt)
#pragma unroll
      for (int mt = 0; mt < TT; ++mt) qsb_dmma884(ci[mt][nt][0], ci[mt][nt][1], S.im[ks][mt], X.re[ks][nt]);
  }
}

// One derivative: K = (with_v ? W : 0) + half-size product, written to o.  RAW: projector type (qsb_dmma_half_raw), else
// the general rank-one form ab = {rho, (row, has_v), b0, b1}.
template <int NQ, int WHAT, int KB, bool RAW>
QSB_DEVICE void qsb_jac_param(const QsbFrag<(1 << NQ) / 4, QsbWarpGeom<NQ>::TT>& X, const QsbFrag<(1 << NQ) / 4, QsbWarpGeom<NQ>::TT>& S,
                              const cplx* ab, bool with_v, const double* wtile, int r0, int c0, int lane, double* o) {
  using G = QsbWarpGeom<NQ>;
  constexpr int N = G::N, NN = G::NN, KS = G::KS, TT = G::TT;
  const int g = lane >> 2, t = lane & 3;
  double cr[TT][TT][2], ci[TT][TT][2];
  if (with_v) {  // generator-type parameter: start from W = -i/2 V
    const double2* w2 = reinterpret_cast<const double2*>(wtile);
#pragma unroll
    for (int mt = 0; mt < TT; ++mt)
#pragma unroll
      for (int nt = 0; nt < TT; ++nt) {
        const int i = (mt * TT + nt) * 2;
        const double2 a = w2[i * 32 + lane], b = w2[(i + 1) * 32 + lane];
        cr[mt][nt][0] = a.x; cr[mt][nt][1] = a.y; ci[mt][nt][0] = b.x; ci[mt][nt][1] = b.y;
      }
  } else {
#pragma unroll
    for (int mt = 0; mt < TT; ++mt)
#pragma unroll
      for (int nt = 0; nt < TT; ++nt) { cr[mt][nt][0] = cr[mt][nt][1] = ci[mt][nt][0] = ci[mt][nt][1] = 0.0; }
  }
  if (RAW) qsb_dmma_half_raw<KS, TT, KB>(cr, ci, S, X);
  else qsb_dmma_half<KS, TT, KB>(cr, ci, S, X, ab);
#pragma unroll
  for (int mt = 0; mt < TT; ++mt)
#pragma unroll
    for (int nt = 0; nt < TT; ++nt) {
      const int e = (r0 + 8 * mt + g) * N + c0 + 8 * nt + 2 * t;
      if (WHAT == QSB_RUN_MATJAC) {
        *reinterpret_cast<double2*>(o + 2 * e) = make_double2(cr[mt][nt][0], ci[mt][nt][0]);
        *reinterpret_cast<double2*>(o + 2 * e + 2) = make_double2(cr[mt][nt][1], ci[mt][nt][1]);
      } else {
        *reinterpret_cast<double2*>(o + e) = make_double2(cr[mt][nt][0], cr[mt][nt][1]);
        *reinterpret_cast<double2*>(o + NN + e) = make_double2(ci[mt][nt][0], ci[mt][nt][1]);
      }
    }
}

// One parameterised (or constant) one-qubit op of the backward sweep whose qubit sits at k-step bit KB:
// X <- g^dagger X (peel), K_k = c V + (S a_k)(b_k^T X) for its parameters, S <- S g.  The two leaves of the search
// gatesets take the projector shortcut for three of their five parameters, as straight-line code (a version that looped
// over stages and table flags lost 30 % to its extra branches, profiles/r1/ah_*):
//   U3    phi on (S, g X) | peel | theta (general) | S <- S g | lambda on (S g, X)
//   XZXZ  v1 on (S, g X)  | peel | v0 (general)    | S <- S g
// Enabled for n = 5 only (A/B on the B200, profiles/r1/ai_ab.log: n = 5 +11 %; n = 4 -13 %: its hot loop just fits the
// 32 KB instruction cache without the two extra product sites and no longer does with them).
template <int NQ, int WHAT, int KB>
QSB_DEVICE void qsb_jac_op(QsbFrag<(1 << NQ) / 4, QsbWarpGeom<NQ>::TT>& X, QsbFrag<(1 << NQ) / 4, QsbWarpGeom<NQ>::TT>& S,
                           const qsb_op& op, int np, const cplx* gtab, const double* wtile, int r0, int c0, int lane,
                           bool update_s, double* oj) {
  using G = QsbWarpGeom<NQ>;
  constexpr int NN = G::NN, KS = G::KS, TT = G::TT;
  constexpr bool PROJ = NQ >= 5;
  const bool u3 = PROJ && op.opcode == QSB_OP_U3, fast = PROJ && (u3 || op.opcode == QSB_OP_XZXZ);
  cplx m[4];
  qsb_warp_gate(op, gtab, 0, m);
  if (fast) qsb_jac_param<NQ, WHAT, KB, true>(X, S, nullptr, !u3, wtile, r0, c0, lane, oj + (size_t)2 * NN);
  {
    cplx md[4] = {m[0], m[1], m[2], m[3]};
    qsb_dagger2x2(md);
    qsb_mix_thread<KS, TT, KB>(X, md);  // X = R_j T^dagger
  }
  const int ngen = fast ? 1 : np;
  for (int k = 0; k < ngen; ++k) {
    const cplx* ab = gtab + 4 * (op.reserved + 1 + k);
    qsb_jac_param<NQ, WHAT, KB, false>(X, S, ab, ab[1].y != 0.0, wtile, r0, c0, lane, oj + (size_t)k * 2 * NN);
  }
  if (update_s || u3) {
    qsb_transpose2x2(m);
    qsb_mix_thread<KS, TT, KB>(S, m);  // S <- S g_j
  }
  if (u3) qsb_jac_param<NQ, WHAT, KB, true>(X, S, nullptr, false, wtile, r0, c0, lane, oj + (size_t)4 * NN);
}

// doubles of per-warp shared memory qsb_warp_jac_range needs for W = -i/2 V (one C-fragment tile, re and im)
#define QSB_WTILE_DOUBLES 512

// One warp, one tile of one instance, ops [jb, je) of the program (the whole program: jb = 0, je = n_ops).
// WHAT = QSB_RUN_MATJAC: out_u = U, out_j[k] = J_k, interleaved complex;
// WHAT = QSB_RUN_RESID: out_u = r, out_j[k] = [Re K_k ; Im K_k] (see qsb_jac_instance in qsb_generic.cuh).
// gtab: gate table of this parameter vector (qsb_fill_gates with rank1 = true); wtile: QSB_WTILE_DOUBLES of this warp's
// shared memory; tile = warp-row * WR + warp-col of this warp's block.  A segment [jb, je) first builds its own prefix
// X = g_{je-1} .. g_0 T^dagger and suffix S = g_{M-1} .. g_{je} with structured updates only, so several warps can
// share the parameters of one instance without exchanging data (used by the LM solver).  The segment that ends at the
// last op writes out_u (every tile its own block of it).
template <int NQ, int WHAT>
QSB_DEVICE void qsb_warp_jac_range(const QsbDevProgram& P, const double* target, const cplx* gtab, double* wtile, int tile,
                                   int lane, int jb, int je, double* out_u, double* out_j) {
  using G = QsbWarpGeom<NQ>;
  constexpr int N = G::N, NN = G::NN, KS = G::KS, TT = G::TT;
  const int g = lane >> 2, t = lane & 3;
  const int r0 = (tile / G::WR) * 8 * TT, c0 = (tile % G::WR) * 8 * TT;
  QsbFrag<KS, TT> X, S;
  QsbKLayout L;
  L.pos = L.log = 0x43210u;
  L.mru = -1; L.mru2 = -1;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks)
#pragma unroll
    for (int o = 0; o < TT; ++o) {
      const int k = 4 * ks + t, c = c0 + 8 * o + g;
      if (WHAT == QSB_RUN_RESID) {  // X = T^dagger
        const double2 tv = *reinterpret_cast<const double2*>(target + 2 * (c * N + k));
        X.re[ks][o] = tv.x; X.im[ks][o] = -tv.y;
      } else {
        X.re[ks][o] = (k == c) ? 1.0 : 0.0; X.im[ks][o] = 0.0;
      }
    }
  // forward sweep: X <- g_j X for the ops below je (S does not exist yet: only X is exchanged)
  for (int j = 0; j < je; ++j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) {
      const int kb = qsb_bring_in<NQ, TT, false>(X, S, L, NQ - 1 - op.q1, lane);
      qsb_cnot_slot<NQ, TT>(X, kb, qsb_nib(L.pos, NQ - 1 - op.q0), lane);
      continue;
    }
    cplx m[4];
    qsb_warp_gate(op, gtab, 0, m);
    const int kb = qsb_bring_in<NQ, TT, false>(X, S, L, NQ - 1 - op.q0, lane);
    qsb_mix_slot<NQ, TT>(X, kb, m);
  }
  // S = the warp's row block of the identity in the layout the forward sweep left behind
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const int kp = 4 * ks + t;
    int kl = 0;
#pragma unroll
    for (int b = 0; b < NQ; ++b) kl |= ((kp >> b) & 1) << qsb_nib(L.log, b);
#pragma unroll
    for (int o = 0; o < TT; ++o) { S.re[ks][o] = (kl == r0 + 8 * o + g) ? 1.0 : 0.0; S.im[ks][o] = 0.0; }
  }
  // suffix of a segment that does not end at the last op: S <- S g_j for j = M-1 .. je
  for (int j = P.n_ops - 1; j >= je; --j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) {
      const int kb = qsb_bring_in<NQ, TT, true>(X, S, L, NQ - 1 - op.q1, lane);
      qsb_cnot_slot<NQ, TT>(S, kb, qsb_nib(L.pos, NQ - 1 - op.q0), lane);
      continue;
    }
    cplx m[4];
    qsb_warp_gate(op, gtab, 0, m);
    qsb_transpose2x2(m);
    const int kb = qsb_bring_in<NQ, TT, true>(X, S, L, NQ - 1 - op.q0, lane);
    qsb_mix_slot<NQ, TT>(S, kb, m);
  }
  {  // V = S X = U T^dagger (this warp's tile of it): output of the last segment, and W = -i/2 V for the derivatives
    double vr[TT][TT][2], vi[TT][TT][2];
    qsb_dmma_full<KS, TT>(vr, vi, S, X);
    double2* w2 = reinterpret_cast<double2*>(wtile);
#pragma unroll
    for (int mt = 0; mt < TT; ++mt)
#pragma unroll
      for (int nt = 0; nt < TT; ++nt) {
        const int i = (mt * TT + nt) * 2;
        w2[i * 32 + lane] = make_double2(0.5 * vi[mt][nt][0], 0.5 * vi[mt][nt][1]);
        w2[(i + 1) * 32 + lane] = make_double2(-0.5 * vr[mt][nt][0], -0.5 * vr[mt][nt][1]);
        if (je == P.n_ops) {
          const int r = r0 + 8 * mt + g, c = c0 + 8 * nt + 2 * t, e = r * N + c;
          if (WHAT == QSB_RUN_MATJAC) {
            *reinterpret_cast<double2*>(out_u + 2 * e) = make_double2(vr[mt][nt][0], vi[mt][nt][0]);
            *reinterpret_cast<double2*>(out_u + 2 * e + 2) = make_double2(vr[mt][nt][1], vi[mt][nt][1]);
          } else {
            *reinterpret_cast<double2*>(out_u + e) = make_double2(vr[mt][nt][0] - (r == c ? 1.0 : 0.0), vr[mt][nt][1] - (r == c + 1 ? 1.0 : 0.0));
            *reinterpret_cast<double2*>(out_u + NN + e) = make_double2(vi[mt][nt][0], vi[mt][nt][1]);
          }
        }
      }
    __syncwarp();
  }
  // backward sweep over the segment
  for (int j = je - 1; j >= jb; --j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) {
      if (j == jb) break;
      const int kb = qsb_bring_in<NQ, TT, true>(X, S, L, NQ - 1 - op.q1, lane);
      const int cp = qsb_nib(L.pos, NQ - 1 - op.q0);
      qsb_cnot_slot<NQ, TT>(X, kb, cp, lane);
      qsb_cnot_slot<NQ, TT>(S, kb, cp, lane);
      continue;
    }
    const int np = qsb_num_params(op.opcode);
    if (np == 0 && j == jb) break;
    const int kb = qsb_bring_in<NQ, TT, true>(X, S, L, NQ - 1 - op.q0, lane);
    double* oj = out_j + (size_t)op.param_off * 2 * NN;
    switch (kb) {  // one dispatch per op: everything below is straight-line code for that k-step bit
      case 1: qsb_jac_op<NQ, WHAT, 1>(X, S, op, np, gtab, wtile, r0, c0, lane, j > jb, oj); break;
      case 2: if (NQ >= 4) qsb_jac_op<NQ, WHAT, (NQ >= 4 ? 2 : 1)>(X, S, op, np, gtab, wtile, r0, c0, lane, j > jb, oj); break;
      default: if (NQ >= 5) qsb_jac_op<NQ, WHAT, (NQ >= 5 ? 4 : 1)>(X, S, op, np, gtab, wtile, r0, c0, lane, j > jb, oj); break;
    }
  }
}

// doubles of shared memory of a one-warp CTA of qsb_k_warp_jac (tables + one W tile; every further warp adds a W tile)
QSB_HOST_DEVICE size_t qsb_warp_jac_smem_doubles(int max_ops, int max_params) {
  return QSB_WTILE_DOUBLES + (size_t)8 * (max_ops + max_params) + 2 * (size_t)(max_params + (max_params & 1)) + 3 * (size_t)(max_ops + (max_ops & 1));
}

// CTA = the WARPS warps of ONE instance (n <= 4: a single warp; n = 5: the four quadrant warps).  The instance comes from
// blockIdx alone and the warp index only selects the tile (addresses), so every branch of the sweep depends on values
// the compiler's divergence analysis can see are uniform.  With the work item derived from threadIdx.x >> 5 every branch
// counted as potentially divergent and each 32-bit shuffle of the bit exchanges was wrapped in a WARPSYNC.COLLECTIVE /
// ENDCOLLECTIVE sequence with moves through fixed registers: a third of the kernel's code, which has to stream through a
// 32 KB instruction cache.  The quadrant warps of n = 5 stay in one CTA: they run the same path in step and share
// every instruction fetch (one-warp CTAs were 22 % slower there, profiles/r1/aj_ab.log, ak_*).
template <int NQ, int WHAT>
QSB_GLOBAL __launch_bounds__(32 * QsbWarpGeom<NQ>::WARPS, (NQ == 4 ? 12 : NQ == 5 ? 2 : 1)) qsb_k_warp_jac(QsbBatchArgs a) {
  using G = QsbWarpGeom<NQ>;
  constexpr int NN = G::NN;
  QSB_DYN_SMEM(smem_raw);
  const int lane = threadIdx.x & 31, warp = G::WARPS == 1 ? 0 : (int)(threadIdx.x >> 5);
  // per CTA: gate table [max_ops + max_params][4] complex, the sin/cos table [2 max_params] and the ops of the program
  // worked on last (every sweep decodes them; from shared memory that is a short-latency load), built by all threads and
  // shared by the warps (one address for the whole CTA: the values stay uniform for the compiler); per warp: the W tile
  const size_t tables = qsb_warp_jac_smem_doubles(a.max_ops, a.max_params) - QSB_WTILE_DOUBLES;
  cplx* gtab = (cplx*)smem_raw;
  double* trig = (double*)(gtab + (size_t)4 * (a.max_ops + a.max_params));
  qsb_op* sops = (qsb_op*)(trig + 2 * (size_t)(a.max_params + (a.max_params & 1)));
  double* wtile = (double*)smem_raw + tables + (size_t)warp * QSB_WTILE_DOUBLES;
  const int tid = threadIdx.x, nth = 32 * G::WARPS;
  int staged = -1;
  const long long p_base = a.param_off[a.first];
  for (int li = qsb_cta_first_instance(a.group); li < a.count; li += gridDim.x) {
    const int inst = a.first + li;
    const int pi = a.prog_idx[inst];
    QsbDevProgram P = a.progs[pi];
    const double* x = a.params + a.param_off[inst];
    __syncthreads();  // the tables of the previous instance are no longer read
    if (pi != staged) {
      for (int e = tid; e < P.n_ops * 6; e += nth) ((int32_t*)sops)[e] = ((const int32_t*)P.ops)[e];
      staged = pi;
    }
    for (int p = tid; p < P.n_params; p += nth) {  // (cos, sin) of the scaled angle of every parameter
      double sn, cs;
      sincos(P.pscale[p] * x[p], &sn, &cs);
      trig[2 * p] = cs; trig[2 * p + 1] = sn;
    }
    __syncthreads();
    P.ops = sops;
    qsb_fill_gates(P, trig, gtab, tid, nth, true);
    __syncthreads();
    qsb_warp_jac_range<NQ, WHAT>(P, a.target, gtab, wtile, warp, lane, 0, P.n_ops,
                                 a.out_a + (size_t)li * 2 * NN, a.out_b + (size_t)(a.param_off[inst] - p_base) * 2 * NN);
  }
}
