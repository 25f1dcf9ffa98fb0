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

template <int NQ>
struct QsbGeom {
  static constexpr int N = 1 << NQ;
  static constexpr int NN = N * N;
  static constexpr int T = (NN / 2 < 32) ? 32 : NN / 2;  // threads per CTA
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

// k-qubit constant C (K x K, row-major interleaved) on contiguous qubits q0..q0+k-1.
// left: M <- (I (x) C (x) I) M (or C^dagger); right: M <- M (I (x) C (x) I).
template <int NQ>
QSB_DEVICE void qsb_constk(cplx* M, int q0, int k, const double* C, bool dagger, bool right) {
  constexpr int N = 1 << NQ;
  const int K = 1 << k, lo_bits = NQ - q0 - k, lo_mask = (1 << lo_bits) - 1;
  cplx v[1 << QSB_MAX_QUBITS];
  for (int idx = threadIdx.x; idx < (N / K) * N; idx += blockDim.x) {
    int other = idx % N, grp = idx / N;
    int base = ((grp >> lo_bits) << (lo_bits + k)) | (grp & lo_mask);
    for (int m = 0; m < K; ++m) {
      int line = base | (m << lo_bits);
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
      int line = base | (m2 << lo_bits);
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

// Apply op j to M from the left (dagger: its inverse).  All threads call; caller synchronises afterwards.
template <int NQ>
QSB_DEVICE void qsb_apply_left(cplx* M, const qsb_op& op, const double* trig, const double* consts, bool dagger) {
  if (op.opcode == QSB_OP_CNOT) { qsb_left_cnot<NQ>(M, op.q0, op.q1); return; }
  if (op.opcode == QSB_OP_CONSTK) { qsb_constk<NQ>(M, op.q0, op.q1, consts + op.const_off, dagger, false); return; }
  if (qsb_is_rpp(op.opcode)) { qsb_rpp<NQ>(M, M, op.opcode, op.q0, op.q1, trig[0], dagger ? -trig[1] : trig[1], 1.0, false); return; }
  cplx g[4];
  qsb_op2x2(op, trig, consts, dagger, g);
  qsb_left_2x2<NQ>(M, op.q0, g);
}

template <int NQ>
QSB_DEVICE void qsb_apply_right(cplx* M, const qsb_op& op, const double* trig, const double* consts) {
  if (op.opcode == QSB_OP_CNOT) { qsb_right_cnot<NQ>(M, op.q0, op.q1); return; }
  if (op.opcode == QSB_OP_CONSTK) { qsb_constk<NQ>(M, op.q0, op.q1, consts + op.const_off, false, true); return; }
  if (qsb_is_rpp(op.opcode)) { qsb_rpp<NQ>(M, M, op.opcode, op.q0, op.q1, trig[0], trig[1], 1.0, true); return; }
  cplx g[4];
  qsb_op2x2(op, trig, consts, false, g);
  qsb_right_2x2<NQ>(M, op.q0, g);
}

// fill the per-instance sin/cos cache (one thread per op)
QSB_DEVICE void qsb_fill_trig(const QsbDevProgram& P, const double* x, double* trig) {
  for (int j = threadIdx.x; j < P.n_ops; j += blockDim.x) {
    qsb_op op = P.ops[j];
    if (qsb_num_params(op.opcode) > 0) qsb_trig(op.opcode, x + op.param_off, trig + j * QSB_TRIG_STRIDE);
  }
}

template <int NQ>
QSB_DEVICE void qsb_load_init(cplx* M, const double* target_or_null, bool dagger) {
  constexpr int N = 1 << NQ;
  for (int e = threadIdx.x; e < N * N; e += blockDim.x) {
    int r = e >> NQ, c = e & (N - 1);
    if (target_or_null == nullptr) M[e] = cmk(r == c ? 1.0 : 0.0, 0.0);
    else if (dagger) { const double* t = target_or_null + 2 * (c * N + r); M[e] = cmk(t[0], -t[1]); }
    else { const double* t = target_or_null + 2 * e; M[e] = cmk(t[0], t[1]); }
  }
}

// forward pass: X <- g_{M-1} ... g_0 X
template <int NQ>
QSB_DEVICE void qsb_forward(cplx* X, const QsbDevProgram& P, const double* trig) {
  for (int j = 0; j < P.n_ops; ++j) {
    qsb_apply_left<NQ>(X, P.ops[j], trig + j * QSB_TRIG_STRIDE, P.consts, false);
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
  constexpr int N = 1 << NQ;
  QSB_DYN_SMEM(smem_raw);
  cplx* X = (cplx*)smem_raw;
  cplx* red = X + N * N;
  double* trig = (double*)(red + 32);
  for (int li = blockIdx.x; li < a.count; li += gridDim.x) {
    const int inst = a.first + li;
    const QsbDevProgram P = a.progs[a.prog_idx[inst]];
    qsb_fill_trig(P, a.params + a.param_off[inst], trig);
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
  constexpr int N = 1 << NQ, NN = N * N;
  qsb_fill_trig(P, x, trig);
  qsb_load_init<NQ>(X, WHAT == QSB_RUN_RESID ? target : nullptr, true);
  qsb_load_init<NQ>(S, nullptr, false);
  __syncthreads();
  qsb_forward<NQ>(X, P, trig);
  for (int e = threadIdx.x; e < NN; e += blockDim.x) {
    if (WHAT == QSB_RUN_MATJAC) { out_u[2 * e] = X[e].x; out_u[2 * e + 1] = X[e].y; }
    else { out_u[e] = X[e].x - (((e >> NQ) == (e & (N - 1))) ? 1.0 : 0.0); out_u[NN + e] = X[e].y; }
  }
  __syncthreads();  // X is about to be peeled by other threads
  for (int j = P.n_ops - 1; j >= 0; --j) {
    const qsb_op op = P.ops[j];
    const double* tj = trig + j * QSB_TRIG_STRIDE;
    const int np = qsb_num_params(op.opcode);
    if (np == 0 && j == 0) break;  // nothing left to differentiate
    qsb_apply_left<NQ>(X, op, tj, P.consts, true);  // X = R_j T^dagger
    __syncthreads();
    for (int k = 0; k < np; ++k) {
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
      __syncthreads();
      double* o = out_j + (size_t)(op.param_off + k) * 2 * NN;
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
      __syncthreads();
    }
    if (j > 0) { qsb_apply_right<NQ>(S, op, tj, P.consts); __syncthreads(); }
  }
  __syncthreads();
}

template <int NQ, int WHAT>
QSB_GLOBAL __launch_bounds__(QsbGeom<NQ>::T) qsb_k_jac(QsbBatchArgs a) {
  constexpr int N = 1 << NQ, NN = N * N;
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
// X, L: N*N tiles; red: 32 cplx; wpart: 96 cplx; acc: P cplx; all in shared memory.  grad may be shared or global.
template <int NQ>
QSB_DEVICE double qsb_hs_instance(cplx* X, cplx* L, cplx* red, cplx* wpart, cplx* acc, double* trig,
                                  const QsbDevProgram& P, const double* x, const double* target, double* grad) {
  constexpr int N = 1 << NQ, NN = N * N;
  __syncthreads();
  qsb_fill_trig(P, x, trig);
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
    const double* tj = trig + j * QSB_TRIG_STRIDE;
    const int np = qsb_num_params(op.opcode);
    if (np == 0) {
      if (j == 0) break;
      qsb_apply_left<NQ>(X, op, tj, P.consts, true);
      qsb_apply_left<NQ>(L, op, tj, P.consts, true);
      __syncthreads();
      continue;
    }
    cplx pk[3] = {cmk(0, 0), cmk(0, 0), cmk(0, 0)};
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
    for (int k = 0; k < np; ++k) {
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
  constexpr int N = 1 << NQ, NN = N * N;
  QSB_DYN_SMEM(smem_raw);
  cplx* X = (cplx*)smem_raw;
  cplx* L = X + NN;
  cplx* red = L + NN;           // 32
  cplx* wpart = red + 32;       // [3][32]
  cplx* acc = wpart + 96;       // [max_params]
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
