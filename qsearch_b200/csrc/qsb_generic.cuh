// Generic gate-program interpreter: one CTA per circuit instance, N x N complex-fp64 tiles in shared memory.
// Handles every opcode (including k-qubit constants) for n <= QSB_MAX_QUBITS.  The specialised register-tiled
// kernels in qsb_fast.cuh cover the standard search gatesets; this file is the always-correct general path and
// the one the API-parity entry points (matrix / mat_jac) use.
//
// Algorithm (restating qsearch/gates.py:965-1004, ProductGate.mat_jac, on the flat program):
//   forward:   X <- g_j X for j = 0..M-1                               (X starts at I, or T^dagger for residuals)
//   backward:  for j = M-1..0:  X <- g_j^dagger X   (peel, gates.py:992: unitary layers are undone, not stored)
//                               J_k = S (dg_k (x) I) X  for the params k of op j      (gates.py:993-995)
//                               S <- S g_j                                             (gates.py:996, mirrored)
// A one-qubit gate on qubit q mixes row pairs with stride 2^(n-1-q) (left) or column pairs (right); CNOT is a
// row / column permutation.
#pragma once
#include "qsb_device.cuh"

// The template parameter NQ of everything in this file is a GEOMETRY CODE: 1..QSB_MAX_QUBITS = that many qubits
// (N = 2^NQ); QSB_GEOM_QUTRIT + m = m qutrits (N = 3^m; SingleQutritGate + constants, gates.py:411-602).  Index arithmetic
// uses / N and % N (compile-time constants: shifts and masks for the qubit geometries).
#define QSB_GEOM_QUTRIT 8
template <int NQ>
struct QsbDims {
  static constexpr bool TRIT = NQ > QSB_GEOM_QUTRIT;
  static constexpr int ND = TRIT ? NQ - QSB_GEOM_QUTRIT : NQ;  // qudits
  static constexpr int RADIX = TRIT ? 3 : 2;
  static constexpr int N = TRIT ? (ND == 1 ? 3 : ND == 2 ? 9 : 27) : (1 << ND);
  static constexpr int NN = N * N;
  static constexpr int TRIG = TRIT ? QSB_TRIG_STRIDE3 : QSB_TRIG_STRIDE;  // doubles of trig cache per op
  static constexpr int MAXNP = TRIT ? 8 : 3;                              // most parameters of one op
};
QSB_HOST_DEVICE int qsb_geom_code(int radix, int n_qudits) { return radix == 3 ? QSB_GEOM_QUTRIT + n_qudits : n_qudits; }
QSB_HOST_DEVICE size_t qsb_geom_dim(int radix, int n_qudits) {
  size_t n = 1;
  for (int i = 0; i < n_qudits; ++i) n *= (size_t)radix;
  return n;
}

template <int NQ>
struct QsbGeom {
  static constexpr int N = QsbDims<NQ>::N;
  static constexpr int NN = N * N;
  static constexpr int T = (NN / 2 < 32) ? 32 : ((NN + 1) / 2 + 31) / 32 * 32;  // threads per CTA (a multiple of 32)
};

struct QsbBatchArgs {
  const QsbDevProgram* progs;
  const int32_t* prog_idx;      // [n_instances]
  const long long* param_off;   // [n_instances + 1]
  const double* params;         // ragged
  const double* target;         // [N][N][2] (may be null for MATRIX / MATJAC)
  int first, count;
  int max_ops;
  int max_params;
  double* out_a;
  double* out_b;
  double* out_c;
  int group = 0;  // warp kernels, persistent grid = SMs x group: CTAs per SM (see qsb_cta_first_instance); 0 = plain order
};

// First instance of a CTA of a persistent grid (stride gridDim.x).  CTAs are handed to the SMs round-robin, so with the
// plain order the CTAs that share an SM work on instances that lie a whole SM count apart; this order gives the `group`
// CTAs of one SM consecutive instances, which usually belong to the same circuit: the same code path in the shared 32 KB
// instruction cache, and neighbouring output rows.  (Only the order changes: any CTA-to-SM placement gives correct results.)
QSB_DEVICE int qsb_cta_first_instance(int group) {
  if (group <= 1) return blockIdx.x;
  const int nsm = gridDim.x / group;
  return (int)(blockIdx.x % nsm) * group + (int)(blockIdx.x / nsm);
}

QSB_DEVICE int qsb_insert_zero_bit(int v, int sh) { return ((v >> sh) << (sh + 1)) | (v & ((1 << sh) - 1)); }

// ---------------------------------------------------------------------------- elementary updates
template <int NQ>
QSB_DEVICE void qsb_left_2x2(cplx* M, int q, const cplx* g) {
  constexpr int N = 1 << NQ;
  const int sh = NQ - 1 - q, s = 1 << sh;
  for (int idx = threadIdx.x; idx < N * N / 2; idx += blockDim.x) {
    int c = idx & (N - 1), r0 = qsb_insert_zero_bit(idx >> NQ, sh), r1 = r0 | s;
    cplx a = M[r0 * N + c], b = M[r1 * N + c];
    M[r0 * N + c] = cfma(g[1], b, cmul(g[0], a));
    M[r1 * N + c] = cfma(g[3], b, cmul(g[2], a));
  }
}

template <int NQ>
QSB_DEVICE void qsb_right_2x2(cplx* M, int q, const cplx* g) {
  constexpr int N = 1 << NQ;
  const int sh = NQ - 1 - q, s = 1 << sh;
  for (int idx = threadIdx.x; idx < N * N / 2; idx += blockDim.x) {
    int i = idx / (N / 2), c0 = qsb_insert_zero_bit(idx % (N / 2), sh), c1 = c0 | s;
    cplx a = M[i * N + c0], b = M[i * N + c1];
    M[i * N + c0] = cfma(b, g[2], cmul(a, g[0]));
    M[i * N + c1] = cfma(b, g[3], cmul(a, g[1]));
  }
}

template <int NQ>
QSB_DEVICE void qsb_left_cnot(cplx* M, int ctrl, int tgt) {
  constexpr int N = 1 << NQ;
  const int sh = NQ - 1 - tgt, s = 1 << sh, cb = 1 << (NQ - 1 - ctrl);
  for (int idx = threadIdx.x; idx < N * N / 2; idx += blockDim.x) {
    int c = idx & (N - 1), r0 = qsb_insert_zero_bit(idx >> NQ, sh), r1 = r0 | s;
    if (r0 & cb) { cplx a = M[r0 * N + c]; M[r0 * N + c] = M[r1 * N + c]; M[r1 * N + c] = a; }
  }
}

template <int NQ>
QSB_DEVICE void qsb_right_cnot(cplx* M, int ctrl, int tgt) {
  constexpr int N = 1 << NQ;
  const int sh = NQ - 1 - tgt, s = 1 << sh, cb = 1 << (NQ - 1 - ctrl);
  for (int idx = threadIdx.x; idx < N * N / 2; idx += blockDim.x) {
    int i = idx / (N / 2), c0 = qsb_insert_zero_bit(idx % (N / 2), sh), c1 = c0 | s;
    if (c0 & cb) { cplx a = M[i * N + c0]; M[i * N + c0] = M[i * N + c1]; M[i * N + c1] = a; }
  }
}

// k-qudit constant C (K x K, K = radix^k, row-major interleaved) on contiguous qudits q0..q0+k-1.
// left: M <- (I (x) C (x) I) M (or C^dagger); right: M <- M (I (x) C (x) I).
template <int NQ>
QSB_DEVICE void qsb_constk(cplx* M, int q0, int k, const double* C, bool dagger, bool right) {
  constexpr int N = QsbDims<NQ>::N, R = QsbDims<NQ>::RADIX;
  int K = 1, lo = 1;  // lo = radix^(qudits below the gate): stride of the gate's least significant digit
  for (int i = 0; i < k; ++i) K *= R;
  for (int i = 0; i < QsbDims<NQ>::ND - q0 - k; ++i) lo *= R;
  cplx v[1 << QSB_MAX_QUBITS];  // 32 >= 27
  for (int idx = threadIdx.x; idx < (N / K) * N; idx += blockDim.x) {
    int other = idx % N, grp = idx / N;
    int base = (grp / lo) * (lo * K) + grp % lo;
    for (int m = 0; m < K; ++m) {
      int line = base + m * lo;
      v[m] = right ? M[other * N + line] : M[line * N + other];
    }
    for (int m2 = 0; m2 < K; ++m2) {
      cplx acc = cmk(0, 0);
      for (int m = 0; m < K; ++m) {
        // left:  out[m2] = sum_m C[m2][m] v[m]   (dagger: conj(C[m][m2]))
        // right: out[m2] = sum_m v[m] C[m][m2]
        int e = right ? (m * K + m2) : (dagger ? (m * K + m2) : (m2 * K + m));
        cplx c = cmk(C[2 * e], C[2 * e + 1]);
        acc = (dagger && !right) ? cfma_conj(c, v[m], acc) : cfma(c, v[m], acc);
      }
      int line = base + m2 * lo;
      if (right) M[other * N + line] = acc; else M[line * N + other] = acc;
    }
  }
}

// Two-qubit rotation R = exp(-i x/2 P(x)P) = c I - i s P(x)P on qubits (q0, q1), P = X, Y or Z (circuits.rs:457-633):
// P(x)P is a signed permutation of the index, so R mixes index pairs (k, k ^ m0 ^ m1) [XX, YY] or is diagonal [ZZ]:
//   XX: (PP v)[k] = v[k ^ m]      YY: (PP v)[k] = -(-1)^(b0 + b1) v[k ^ m]      ZZ: (PP v)[k] = (-1)^(b0 + b1) v[k]
// dst = scale (c src - i s PP src).  With (c, s, scale) = (-s, c, 1/2) this is dR/dx = -i/2 PP R = R(x + pi) / 2; with
// s -> -s the inverse.  left: rows of M are the index; right: M <- M R, columns (PP is symmetric for XX, YY, ZZ).
// Out of place when dst != src; in place each thread owns a whole pair.
QSB_DEVICE void qsb_rpp_pair(int opcode, bool odd, cplx a, cplx b, double c, double s, cplx& oa, cplx& ob) {
  cplx pa, pb;  // (PP v)[k0], (PP v)[k1]
  if (opcode == QSB_OP_RXX) { pa = b; pb = a; }
  else if (opcode == QSB_OP_RYY) { const double f = odd ? 1.0 : -1.0; pa = cscale(b, f); pb = cscale(a, f); }
  else { const double f = odd ? -1.0 : 1.0; pa = cscale(a, f); pb = cscale(b, f); }
  // c v - i s p:  re = c v.re + s p.im,  im = c v.im - s p.re
  oa = cmk(fma(c, a.x, s * pa.y), fma(c, a.y, -s * pa.x));
  ob = cmk(fma(c, b.x, s * pb.y), fma(c, b.y, -s * pb.x));
}

template <int NQ>
QSB_DEVICE void qsb_rpp(cplx* dst, const cplx* src, int opcode, int q0, int q1, double c, double s, double scale, bool right) {
  constexpr int N = 1 << NQ;
  const int m0 = 1 << (NQ - 1 - q0), m1 = 1 << (NQ - 1 - q1), sh = NQ - 1 - q0;
  c *= scale; s *= scale;
  for (int idx = threadIdx.x; idx < N * N / 2; idx += blockDim.x) {
    const int o = idx & (N - 1), k0 = qsb_insert_zero_bit(idx >> NQ, sh), k1 = k0 ^ m0 ^ m1;  // k0 has bit q0 clear
    const int e0 = right ? o * N + k0 : k0 * N + o, e1 = right ? o * N + k1 : k1 * N + o;
    const cplx a = src[e0], b = src[e1];
    cplx oa, ob;
    qsb_rpp_pair(opcode, (k0 & m1) != 0, a, b, c, s, oa, ob);  // parity of (b0 + b1) at k0 (b0 = 0); k1 has the same parity
    dst[e0] = oa; dst[e1] = ob;
  }
}

// One-qutrit gate g (3x3 row-major) on qutrit q.  left: dst <- (I (x) g (x) I) src over row triples with stride
// 3^(n-1-q) (out of place when dst != src; in place each thread owns a whole triple); right: M <- M (I (x) g (x) I).
template <int NQ>
QSB_DEVICE void qsb_left_3x3(cplx* dst, const cplx* src, int q, const cplx* g) {
  constexpr int N = QsbDims<NQ>::N;
  int s = 1;
  for (int i = 0; i < QsbDims<NQ>::ND - 1 - q; ++i) s *= 3;
  for (int idx = threadIdx.x; idx < N * N / 3; idx += blockDim.x) {
    const int c = idx % N, tq = idx / N, r0 = (tq / s) * 3 * s + tq % s;
    const cplx a = src[r0 * N + c], b = src[(r0 + s) * N + c], d = src[(r0 + 2 * s) * N + c];
#pragma unroll
    for (int i = 0; i < 3; ++i) dst[(r0 + i * s) * N + c] = cfma(g[3 * i + 2], d, cfma(g[3 * i + 1], b, cmul(g[3 * i], a)));
  }
}

template <int NQ>
QSB_DEVICE void qsb_right_3x3(cplx* M, int q, const cplx* g) {
  constexpr int N = QsbDims<NQ>::N;
  int s = 1;
  for (int i = 0; i < QsbDims<NQ>::ND - 1 - q; ++i) s *= 3;
  for (int idx = threadIdx.x; idx < N * N / 3; idx += blockDim.x) {
    const int i = idx / (N / 3), tq = idx % (N / 3), c0 = (tq / s) * 3 * s + tq % s;
    const cplx a = M[i * N + c0], b = M[i * N + c0 + s], d = M[i * N + c0 + 2 * s];
#pragma unroll
    for (int j = 0; j < 3; ++j) M[i * N + c0 + j * s] = cfma(d, g[6 + j], cfma(b, g[3 + j], cmul(a, g[j])));
  }
}

// Apply op j to M from the left (dagger: its inverse).  All threads call; caller synchronises afterwards.
template <int NQ>
QSB_DEVICE void qsb_apply_left(cplx* M, const qsb_op& op, const double* trig, const double* consts, bool dagger) {
  if constexpr (QsbDims<NQ>::TRIT) {  // radix 3: the single-qutrit leaf and k-qutrit constants
    if (op.opcode == QSB_OP_CONSTK) { qsb_constk<NQ>(M, op.q0, op.q1, consts + op.const_off, dagger, false); return; }
    cplx g3[9];
    qsb_qutrit3x3(trig, 0, g3);
    if (dagger) qsb_dagger3x3(g3);
    qsb_left_3x3<NQ>(M, M, op.q0, g3);
    return;
  } else {
  if (op.opcode == QSB_OP_CNOT) { qsb_left_cnot<NQ>(M, op.q0, op.q1); return; }
  if (op.opcode == QSB_OP_CONSTK) { qsb_constk<NQ>(M, op.q0, op.q1, consts + op.const_off, dagger, false); return; }
  if (qsb_is_rpp(op.opcode)) { qsb_rpp<NQ>(M, M, op.opcode, op.q0, op.q1, trig[0], dagger ? -trig[1] : trig[1], 1.0, false); return; }
  cplx g[4];
  qsb_op2x2(op, trig, consts, dagger, g);
  qsb_left_2x2<NQ>(M, op.q0, g);
  }
}

template <int NQ>
QSB_DEVICE void qsb_apply_right(cplx* M, const qsb_op& op, const double* trig, const double* consts) {
  if constexpr (QsbDims<NQ>::TRIT) {
    if (op.opcode == QSB_OP_CONSTK) { qsb_constk<NQ>(M, op.q0, op.q1, consts + op.const_off, false, true); return; }
    cplx g3[9];
    qsb_qutrit3x3(trig, 0, g3);
    qsb_right_3x3<NQ>(M, op.q0, g3);
    return;
  } else {
  if (op.opcode == QSB_OP_CNOT) { qsb_right_cnot<NQ>(M, op.q0, op.q1); return; }
  if (op.opcode == QSB_OP_CONSTK) { qsb_constk<NQ>(M, op.q0, op.q1, consts + op.const_off, false, true); return; }
  if (qsb_is_rpp(op.opcode)) { qsb_rpp<NQ>(M, M, op.opcode, op.q0, op.q1, trig[0], trig[1], 1.0, true); return; }
  cplx g[4];
  qsb_op2x2(op, trig, consts, false, g);
  qsb_right_2x2<NQ>(M, op.q0, g);
  }
}

// fill the per-instance sin/cos cache (one thread per op)
QSB_DEVICE void qsb_fill_trig(const QsbDevProgram& P, const double* x, double* trig, int stride = QSB_TRIG_STRIDE) {
  for (int j = threadIdx.x; j < P.n_ops; j += blockDim.x) {
    qsb_op op = P.ops[j];
    if (qsb_num_params(op.opcode) > 0) qsb_trig(op.opcode, x + op.param_off, trig + j * stride);
  }
}

template <int NQ>
QSB_DEVICE void qsb_load_init(cplx* M, const double* target_or_null, bool dagger) {
  constexpr int N = QsbDims<NQ>::N;
  for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
    int r = e / N, c = e % N;
    if (target_or_null == nullptr) M[e] = cmk(r == c ? 1.0 : 0.0, 0.0);
    else if (dagger) { const double* t = target_or_null + 2 * (c * N + r); M[e] = cmk(t[0], -t[1]); }
    else { const double* t = target_or_null + 2 * e; M[e] = cmk(t[0], t[1]); }
  }
}

// forward pass: X <- g_{M-1} ... g_0 X
template <int NQ>
QSB_DEVICE void qsb_forward(cplx* X, const QsbDevProgram& P, const double* trig) {
  for (int j = 0; j < P.n_ops; ++j) {
    qsb_apply_left<NQ>(X, P.ops[j], trig + j * QsbDims<NQ>::TRIG, P.consts, false);
    __syncthreads();
  }
}

// deterministic block-wide sum of a complex value; result valid in every thread. red: >= 32 cplx of smem.
QSB_DEVICE cplx qsb_block_sum(cplx v, cplx* red) {
  for (int m = 16; m > 0; m >>= 1) {
    v.x += __shfl_xor_sync(0xffffffffu, v.x, m);
    v.y += __shfl_xor_sync(0xffffffffu, v.y, m);
  }
  int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  cplx s = cmk(0, 0);
  for (int i = 0; i < nw; ++i) s = cadd(s, red[i]);
  return s;
}

// ---------------------------------------------------------------------------- kernels
// MATRIX: out_a[li] = U.    DIST: out_a[li] = |1 - |Tr(U T^dagger)|/N|   (comparison.py:8-22)
template <int NQ, int WHAT>
QSB_GLOBAL __launch_bounds__(QsbGeom<NQ>::T) qsb_k_forward(QsbBatchArgs a) {
  constexpr int N = QsbDims<NQ>::N;
  QSB_DYN_SMEM(smem_raw);
  cplx* X = (cplx*)smem_raw;
  cplx* red = X + N * N;
  double* trig = (double*)(red + 32);
  for (int li = blockIdx.x; li < a.count; li += gridDim.x) {
    const int inst = a.first + li;
    const QsbDevProgram P = a.progs[a.prog_idx[inst]];
    qsb_fill_trig(P, a.params + a.param_off[inst], trig, QsbDims<NQ>::TRIG);
    qsb_load_init<NQ>(X, WHAT == QSB_RUN_DIST ? a.target : nullptr, true);
    __syncthreads();
    qsb_forward<NQ>(X, P, trig);
    if (WHAT == QSB_RUN_MATRIX) {
      double* out = a.out_a + (size_t)li * 2 * N * N;
      for (int e = threadIdx.x; e < N * N; e += blockDim.x) { out[2 * e] = X[e].x; out[2 * e + 1] = X[e].y; }
    } else {
      cplx d = cmk(0, 0);
      if (threadIdx.x < N) d = X[threadIdx.x * N + threadIdx.x];
      cplx tr = qsb_block_sum(d, red);
      if (threadIdx.x == 0) a.out_a[li] = fabs(1.0 - sqrt(tr.x * tr.x + tr.y * tr.y) / N);
    }
    __syncthreads();
  }
}

// One instance of the Jacobian sweep.  X, S, Y: N*N cplx tiles in shared memory; trig: sin/cos cache.
// MATJAC: out_u = U [N][N][2], out_j[k] = J_k [N][N][2] interleaved (T = I).
// RESID:  out_u = r = [Re(U T^dagger) - I ; Im(U T^dagger)], out_j[k] = [Re K_k ; Im K_k], K_k = J_k T^dagger.
template <int NQ, int WHAT>
QSB_DEVICE void qsb_jac_instance(cplx* X, cplx* S, cplx* Y, double* trig, const QsbDevProgram& P, const double* x,
                                 const double* target, double* out_u, double* out_j) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N, TS = QsbDims<NQ>::TRIG;
  qsb_fill_trig(P, x, trig, TS);
  qsb_load_init<NQ>(X, WHAT == QSB_RUN_RESID ? target : nullptr, true);
  qsb_load_init<NQ>(S, nullptr, false);
  __syncthreads();
  qsb_forward<NQ>(X, P, trig);
  for (int e = threadIdx.x; e < NN; e += blockDim.x) {
    if (WHAT == QSB_RUN_MATJAC) { out_u[2 * e] = X[e].x; out_u[2 * e + 1] = X[e].y; }
    else { out_u[e] = X[e].x - ((e / N == e % N) ? 1.0 : 0.0); out_u[NN + e] = X[e].y; }
  }
  __syncthreads();  // X is about to be peeled by other threads
  for (int j = P.n_ops - 1; j >= 0; --j) {
    const qsb_op op = P.ops[j];
    const double* tj = trig + j * TS;
    const int np = qsb_num_params(op.opcode);
    if (np == 0 && j == 0) break;  // nothing left to differentiate
    qsb_apply_left<NQ>(X, op, tj, P.consts, true);  // X = R_j T^dagger
    __syncthreads();
    for (int k = 0; k < np; ++k) {
      if constexpr (QsbDims<NQ>::TRIT) {  // Y = (dg (x) I) X for the 3x3 derivative, out of place
        cplx dg3[9];
        qsb_qutrit3x3(tj, 1 + k, dg3);
        qsb_left_3x3<NQ>(Y, X, op.q0, dg3);
      } else {
      cplx dg[4];
      qsb_gate2x2(op.opcode, tj, 1 + k, dg);
      if (qsb_is_rpp(op.opcode)) {  // Y = dR X = R(x + pi) X / 2, out of place
        qsb_rpp<NQ>(Y, X, op.opcode, op.q0, op.q1, -tj[1], tj[0], 0.5, false);
      } else {  // Y = (dg (x) I) X, out of place
        const int sh = NQ - 1 - op.q0, s = 1 << sh;
        for (int idx = threadIdx.x; idx < NN / 2; idx += blockDim.x) {
          int c = idx & (N - 1), r0 = qsb_insert_zero_bit(idx >> NQ, sh), r1 = r0 | s;
          cplx xa = X[r0 * N + c], xb = X[r1 * N + c];
          Y[r0 * N + c] = cfma(dg[1], xb, cmul(dg[0], xa));
          Y[r1 * N + c] = cfma(dg[3], xb, cmul(dg[2], xa));
        }
      }
      }
      __syncthreads();
      double* o = out_j + (size_t)(op.param_off + k) * 2 * NN;
      if constexpr (N % 2 == 0) {
      for (int pidx = threadIdx.x; pidx < NN / 2; pidx += blockDim.x) {  // K = S Y, two outputs per thread
        int i = pidx / (N / 2), c0 = 2 * (pidx % (N / 2));
        cplx acc0 = cmk(0, 0), acc1 = cmk(0, 0);
#pragma unroll 4
        for (int m = 0; m < N; ++m) {
          cplx sv = S[i * N + m];
          acc0 = cfma(sv, Y[m * N + c0], acc0);
          acc1 = cfma(sv, Y[m * N + c0 + 1], acc1);
        }
        int e = i * N + c0;
        if (WHAT == QSB_RUN_MATJAC) { o[2 * e] = acc0.x; o[2 * e + 1] = acc0.y; o[2 * e + 2] = acc1.x; o[2 * e + 3] = acc1.y; }
        else { o[e] = acc0.x; o[e + 1] = acc1.x; o[NN + e] = acc0.y; o[NN + e + 1] = acc1.y; }
      }
      } else {
      for (int e = threadIdx.x; e < NN; e += blockDim.x) {  // odd N (qutrits): one output per thread
        const int i = e / N, c = e % N;
        cplx acc = cmk(0, 0);
#pragma unroll 3
        for (int m = 0; m < N; ++m) acc = cfma(S[i * N + m], Y[m * N + c], acc);
        if (WHAT == QSB_RUN_MATJAC) { o[2 * e] = acc.x; o[2 * e + 1] = acc.y; }
        else { o[e] = acc.x; o[NN + e] = acc.y; }
      }
      }
      __syncthreads();
    }
    if (j > 0) { qsb_apply_right<NQ>(S, op, tj, P.consts); __syncthreads(); }
  }
  __syncthreads();
}

template <int NQ, int WHAT>
QSB_GLOBAL __launch_bounds__(QsbGeom<NQ>::T) qsb_k_jac(QsbBatchArgs a) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N;
  QSB_DYN_SMEM(smem_raw);
  cplx* X = (cplx*)smem_raw;
  cplx* S = X + NN;
  cplx* Y = S + NN;
  double* trig = (double*)(Y + NN);
  const long long p_base = a.param_off[a.first];
  for (int li = blockIdx.x; li < a.count; li += gridDim.x) {
    const int inst = a.first + li;
    const QsbDevProgram P = a.progs[a.prog_idx[inst]];
    qsb_jac_instance<NQ, WHAT>(X, S, Y, trig, P, a.params + a.param_off[inst], a.target,
                               a.out_a + (size_t)li * 2 * NN, a.out_b + (size_t)(a.param_off[inst] - p_base) * 2 * NN);
  }
}

// HS form of one instance: returns f = 1 - |S|/N and fills grad[0..P) (comparison.matrix_distance_squared_jac,
// comparison.py:37-55), S = sum(T o conj(U)), S_k = sum(T o conj(J_k)).
// Adjoint sweep: S_k = conj(<L_j, (dg_k (x) I) R_j>) with L_j = (g_{M-1}..g_{j+1})^dagger T, R_j = g_{j-1}..g_0:
// only structured one-qubit updates, no dense product.  Gradient is +inf when S == 0 (qsrs/src/utils.rs:144-146).
// X, L: N*N tiles; red: 32 cplx; wpart: QSB_HS_WPART cplx; acc: P cplx; all in shared memory.  grad may be shared or global.
#define QSB_HS_WPART 256  // [8][32]: per-warp partial sums of up to 8 parameters of one op (qubit leaves use 3)
template <int NQ>
QSB_DEVICE double qsb_hs_instance(cplx* X, cplx* L, cplx* red, cplx* wpart, cplx* acc, double* trig,
                                  const QsbDevProgram& P, const double* x, const double* target, double* grad) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N, TS = QsbDims<NQ>::TRIG, MAXNP = QsbDims<NQ>::MAXNP;
  __syncthreads();
  qsb_fill_trig(P, x, trig, TS);
  qsb_load_init<NQ>(X, nullptr, false);
  qsb_load_init<NQ>(L, target, false);
  __syncthreads();
  qsb_forward<NQ>(X, P, trig);
  cplx part = cmk(0, 0);
  for (int e = threadIdx.x; e < NN; e += blockDim.x) part = cfma_conj(L[e], X[e], part);  // <T, U>
  const cplx tr = qsb_block_sum(part, red);
  const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  for (int j = P.n_ops - 1; j >= 0; --j) {
    const qsb_op op = P.ops[j];
    const double* tj = trig + j * TS;
    const int np = qsb_num_params(op.opcode);
    if (np == 0) {
      if (j == 0) break;
      qsb_apply_left<NQ>(X, op, tj, P.consts, true);
      qsb_apply_left<NQ>(L, op, tj, P.consts, true);
      __syncthreads();
      continue;
    }
    cplx pk[MAXNP];
#pragma unroll
    for (int k = 0; k < MAXNP; ++k) pk[k] = cmk(0, 0);
    if constexpr (QsbDims<NQ>::TRIT) {
      // single-qutrit leaf: peel X, correlate L with (dg_k (x) I) X for its eight parameters, peel L; every thread owns
      // whole row triples, so no barrier is needed between the three steps
      int s = 1;
      for (int i = 0; i < QsbDims<NQ>::ND - 1 - op.q0; ++i) s *= 3;
      cplx gd[9];
      qsb_qutrit3x3(tj, 0, gd);
      qsb_dagger3x3(gd);
      for (int idx = threadIdx.x; idx < NN / 3; idx += blockDim.x) {
        const int c = idx % N, tq = idx / N, r0 = (tq / s) * 3 * s + tq % s;
        cplx xv[3], lv[3], rv[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) { xv[i] = X[(r0 + i * s) * N + c]; lv[i] = L[(r0 + i * s) * N + c]; }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          rv[i] = cfma(gd[3 * i + 2], xv[2], cfma(gd[3 * i + 1], xv[1], cmul(gd[3 * i], xv[0])));  // R_j triple
          X[(r0 + i * s) * N + c] = rv[i];
          L[(r0 + i * s) * N + c] = cfma(gd[3 * i + 2], lv[2], cfma(gd[3 * i + 1], lv[1], cmul(gd[3 * i], lv[0])));
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          cplx dg3[9];
          qsb_qutrit3x3(tj, 1 + k, dg3);
#pragma unroll
          for (int i = 0; i < 3; ++i) {
            const cplx y = cfma(dg3[3 * i + 2], rv[2], cfma(dg3[3 * i + 1], rv[1], cmul(dg3[3 * i], rv[0])));
            pk[k] = cfma_conj(lv[i], y, pk[k]);
          }
        }
      }
    } else {
    if (qsb_is_rpp(op.opcode)) {  // two-qubit rotation: peel X, correlate L with dR X = R(x + pi) X / 2, peel L; pairs (k0, k0 ^ m0 ^ m1)
      const int m0 = 1 << (NQ - 1 - op.q0), m1 = 1 << (NQ - 1 - op.q1), sh0 = NQ - 1 - op.q0;
      const double c = tj[0], sn = tj[1];
      for (int idx = threadIdx.x; idx < NN / 2; idx += blockDim.x) {
        const int o = idx & (N - 1), k0 = qsb_insert_zero_bit(idx >> NQ, sh0), k1 = k0 ^ m0 ^ m1;
        const bool odd = (k0 & m1) != 0;
        cplx ra, rb, ya, yb, la2, lb2;
        qsb_rpp_pair(op.opcode, odd, X[k0 * N + o], X[k1 * N + o], c, -sn, ra, rb);  // R_j pair
        X[k0 * N + o] = ra; X[k1 * N + o] = rb;
        qsb_rpp_pair(op.opcode, odd, ra, rb, -0.5 * sn, 0.5 * c, ya, yb);
        const cplx la = L[k0 * N + o], lb = L[k1 * N + o];
        pk[0] = cfma_conj(la, ya, pk[0]);
        pk[0] = cfma_conj(lb, yb, pk[0]);
        qsb_rpp_pair(op.opcode, odd, la, lb, c, -sn, la2, lb2);
        L[k0 * N + o] = la2; L[k1 * N + o] = lb2;
      }
    } else {
    cplx gd[4], dg[3][4];
    qsb_op2x2(op, tj, P.consts, true, gd);
    for (int k = 0; k < np; ++k) qsb_gate2x2(op.opcode, tj, 1 + k, dg[k]);
    const int sh = NQ - 1 - op.q0, s = 1 << sh;
    for (int idx = threadIdx.x; idx < NN / 2; idx += blockDim.x) {
      int c = idx & (N - 1), r0 = qsb_insert_zero_bit(idx >> NQ, sh), r1 = r0 | s;
      cplx xa = X[r0 * N + c], xb = X[r1 * N + c];
      cplx ra = cfma(gd[1], xb, cmul(gd[0], xa)), rb = cfma(gd[3], xb, cmul(gd[2], xa));  // R_j pair
      X[r0 * N + c] = ra; X[r1 * N + c] = rb;
      cplx la = L[r0 * N + c], lb = L[r1 * N + c];
      for (int k = 0; k < np; ++k) {
        cplx ya = cfma(dg[k][1], rb, cmul(dg[k][0], ra)), yb = cfma(dg[k][3], rb, cmul(dg[k][2], ra));
        pk[k] = cfma_conj(la, ya, pk[k]);
        pk[k] = cfma_conj(lb, yb, pk[k]);
      }
      L[r0 * N + c] = cfma(gd[1], lb, cmul(gd[0], la));
      L[r1 * N + c] = cfma(gd[3], lb, cmul(gd[2], la));
    }
    }
    }
#pragma unroll
    for (int k = 0; k < MAXNP; ++k) {
      if (k >= np) break;
      for (int m = 16; m > 0; m >>= 1) {
        pk[k].x += __shfl_xor_sync(0xffffffffu, pk[k].x, m);
        pk[k].y += __shfl_xor_sync(0xffffffffu, pk[k].y, m);
      }
      if ((threadIdx.x & 31) == 0) wpart[k * 32 + w] = pk[k];
    }
    __syncthreads();
    if ((int)threadIdx.x < np) {
      cplx sk = cmk(0, 0);
      for (int i = 0; i < nw; ++i) sk = cadd(sk, wpart[threadIdx.x * 32 + i]);
      acc[op.param_off + threadIdx.x] = sk;
    }
    __syncthreads();
  }
  __syncthreads();
  const double mag = sqrt(tr.x * tr.x + tr.y * tr.y);
  for (int k = threadIdx.x; k < P.n_params; k += blockDim.x) {
    // Re(S) Re(S_k) + Im(S) Im(S_k) is invariant under conjugating both, so the <T,U>-form values can be used as is
    double num = tr.x * acc[k].x + tr.y * acc[k].y;
    grad[k] = (mag == 0.0) ? __longlong_as_double(0x7ff0000000000000LL) : -num / (N * mag);
  }
  __syncthreads();
  return 1.0 - mag / N;
}

template <int NQ>
QSB_GLOBAL __launch_bounds__(QsbGeom<NQ>::T) qsb_k_hs(QsbBatchArgs a) {
  constexpr int N = QsbDims<NQ>::N, NN = N * N;
  QSB_DYN_SMEM(smem_raw);
  cplx* X = (cplx*)smem_raw;
  cplx* L = X + NN;
  cplx* red = L + NN;           // 32
  cplx* wpart = red + 32;       // [8][32]
  cplx* acc = wpart + QSB_HS_WPART;  // [max_params]
  const long long p_base = a.param_off[a.first];
  for (int li = blockIdx.x; li < a.count; li += gridDim.x) {
    const int inst = a.first + li;
    const QsbDevProgram P = a.progs[a.prog_idx[inst]];
    double* trig = (double*)(acc + P.n_params);
    double f = qsb_hs_instance<NQ>(X, L, red, wpart, acc, trig, P, a.params + a.param_off[inst], a.target,
                                   a.out_b + (a.param_off[inst] - p_base));
    if (threadIdx.x == 0) a.out_a[li] = f;
  }
}
