// Hilbert-Schmidt objective and gradient on register-resident column blocks (n = 3, 4, 5; one-qubit leaves + CNOT).
//
// f = 1 - |S| / N, S = sum(T o conj(U)), S_k = sum(T o conj(J_k))   (comparison.matrix_distance_squared_jac,
// qsearch/comparison.py:37-55).  Same adjoint sweep as qsb_hs_instance (qsb_generic.cuh):
//   S_k = conj(<L_j, (dg_k (x) I) R_j>),  L_j = (g_{M-1} .. g_{j+1})^dagger T,  R_j = g_{j-1} .. g_0,
// but both R and L only ever see LEFT multiplications, so their columns are independent: one warp owns 8 columns
// of both (fragment layout of qsb_warp.cuh with one tile), N/8 warps share an instance and meet only to add up
// their partial traces.  Per parameterised op the warp forms the 2x2 pair correlation
//   C[a][b] = sum over row pairs and columns of conj(L[r_a]) R[r_b]
// once; every derivative of that op is then S_k = sum_ab dg_k[a][b] C[a][b] - three parameters for the price of one.
#pragma once
#include "qsb_warp.cuh"

// add the 8 doubles v[0..7] over the 32 lanes; every lane ends up with v[lane & 7]'s total in `out`
// (halving exchange: 4 + 2 + 1 + 1 + 1 = 9 double shuffles instead of 40)
QSB_DEVICE double qsb_reduce8(double (&v)[8], int lane) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {  // lanes with bit 4 set keep the upper half
    const bool up = (lane & 16) != 0;
    const double send = up ? v[i] : v[i + 4], keep = up ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(QSB_FULL, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const bool up = (lane & 8) != 0;
    const double send = up ? v[i] : v[i + 2], keep = up ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(QSB_FULL, send, 8);
  }
  {
    const bool up = (lane & 4) != 0;
    const double send = up ? v[0] : v[1], keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(QSB_FULL, send, 4);
  }
  v[0] += __shfl_xor_sync(QSB_FULL, v[0], 2);
  v[0] += __shfl_xor_sync(QSB_FULL, v[0], 1);
  return v[0];  // lane holds entry ((lane>>4)&1)*4 + ((lane>>3)&1)*2 + ((lane>>2)&1)
}

// pair correlation of one op on qubit q; C = {C00, C01, C10, C11} as 8 doubles (re, im interleaved), per-lane partial
template <int KS, int KB>
QSB_DEVICE void qsb_corr_thread(const QsbFrag<KS, 1>& L, const QsbFrag<KS, 1>& X, double (&c)[8]) {
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    if (ks & KB) continue;
    const double lar = L.re[ks][0], lai = L.im[ks][0], lbr = L.re[ks | KB][0], lbi = L.im[ks | KB][0];
    const double xar = X.re[ks][0], xai = X.im[ks][0], xbr = X.re[ks | KB][0], xbi = X.im[ks | KB][0];
    c[0] = fma(lar, xar, fma(lai, xai, c[0])); c[1] = fma(lar, xai, fma(-lai, xar, c[1]));  // conj(la) xa
    c[2] = fma(lar, xbr, fma(lai, xbi, c[2])); c[3] = fma(lar, xbi, fma(-lai, xbr, c[3]));  // conj(la) xb
    c[4] = fma(lbr, xar, fma(lbi, xai, c[4])); c[5] = fma(lbr, xai, fma(-lbi, xar, c[5]));  // conj(lb) xa
    c[6] = fma(lbr, xbr, fma(lbi, xbi, c[6])); c[7] = fma(lbr, xbi, fma(-lbi, xbr, c[7]));  // conj(lb) xb
  }
}

template <int KS>
QSB_DEVICE void qsb_corr_lane(const QsbFrag<KS, 1>& L, const QsbFrag<KS, 1>& X, int lbit, int lane, double (&c)[8]) {
  const int mask = 1 << lbit;
  const bool hi = (lane >> lbit) & 1;
  double sr = 0, si = 0, pr = 0, pi = 0;  // conj(l) x (same row), conj(l) x_partner
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const double lr = L.re[ks][0], li = L.im[ks][0], xr = X.re[ks][0], xi = X.im[ks][0];
    const double yr = __shfl_xor_sync(QSB_FULL, xr, mask), yi = __shfl_xor_sync(QSB_FULL, xi, mask);
    sr = fma(lr, xr, fma(li, xi, sr)); si = fma(lr, xi, fma(-li, xr, si));
    pr = fma(lr, yr, fma(li, yi, pr)); pi = fma(lr, yi, fma(-li, yr, pi));
  }
  // a lane whose bit is b holds rows r_b: same-row products go to C[b][b], partner products to C[b][1-b]
  c[0] = hi ? 0.0 : sr; c[1] = hi ? 0.0 : si; c[2] = hi ? 0.0 : pr; c[3] = hi ? 0.0 : pi;
  c[4] = hi ? pr : 0.0; c[5] = hi ? pi : 0.0; c[6] = hi ? sr : 0.0; c[7] = hi ? si : 0.0;
}

template <int NQ>
QSB_DEVICE void qsb_corr(const QsbFrag<(1 << NQ) / 4, 1>& L, const QsbFrag<(1 << NQ) / 4, 1>& X, int q, int lane,
                         double (&c)[8]) {
  constexpr int KS = (1 << NQ) / 4;
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i] = 0.0;
  const int bit = NQ - 1 - q;
  switch (bit) {
    case 0: case 1: qsb_corr_lane<KS>(L, X, bit, lane, c); break;
    case 2: qsb_corr_thread<KS, 1>(L, X, c); break;
    case 3: if (NQ >= 4) qsb_corr_thread<KS, (NQ >= 4 ? 2 : 1)>(L, X, c); break;
    default: if (NQ >= 5) qsb_corr_thread<KS, (NQ >= 5 ? 4 : 1)>(L, X, c); break;
  }
}

// One warp, the 8 columns [8 cb, 8 cb + 8) of one instance.  Writes this block's share of <T, U> to tr_part[0..1]
// (lane 0) and, for every parameterised op j, its share of C_j to cpart[(j * ncb + cb) * 8 ..] (8 doubles).
template <int NQ>
QSB_DEVICE void qsb_warp_hs_cols(const QsbDevProgram& P, const double* target, const cplx* gtab, int cb, int ncb, int lane,
                                 double* tr_part, double* cpart) {
  constexpr int N = 1 << NQ, KS = N / 4;
  const int g = lane >> 2, t = lane & 3, c = 8 * cb + g;
  QsbFrag<KS, 1> X, L;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    const int k = 4 * ks + t;
    const double2 tv = *reinterpret_cast<const double2*>(target + 2 * (k * N + c));
    L.re[ks][0] = tv.x; L.im[ks][0] = tv.y;
    X.re[ks][0] = (k == c) ? 1.0 : 0.0; X.im[ks][0] = 0.0;
  }
  for (int j = 0; j < P.n_ops; ++j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) { qsb_cnot<NQ, 1>(X, op.q0, op.q1, lane); continue; }
    cplx m[4];
    qsb_warp_gate(op, gtab, 0, m);
    qsb_mix<NQ, 1>(X, op.q0, m, lane);
  }
  {
    double sr = 0, si = 0;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {  // conj(T) U
      sr = fma(L.re[ks][0], X.re[ks][0], fma(L.im[ks][0], X.im[ks][0], sr));
      si = fma(L.re[ks][0], X.im[ks][0], fma(-L.im[ks][0], X.re[ks][0], si));
    }
    for (int m = 16; m > 0; m >>= 1) { sr += __shfl_xor_sync(QSB_FULL, sr, m); si += __shfl_xor_sync(QSB_FULL, si, m); }
    if (lane == 0) { tr_part[0] = sr; tr_part[1] = si; }
  }
  for (int j = P.n_ops - 1; j >= 0; --j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) {
      if (j == 0) break;
      qsb_cnot<NQ, 1>(X, op.q0, op.q1, lane);
      qsb_cnot<NQ, 1>(L, op.q0, op.q1, lane);
      continue;
    }
    const int np = qsb_num_params(op.opcode);
    if (np == 0 && j == 0) break;
    cplx m[4];
    qsb_warp_gate(op, gtab, 0, m);
    qsb_dagger2x2(m);
    qsb_mix<NQ, 1>(X, op.q0, m, lane);  // X = R_j
    if (np > 0) {
      double cc[8];
      qsb_corr<NQ>(L, X, op.q0, lane, cc);
      const double tot = qsb_reduce8(cc, lane);
      if ((lane & 3) == 0) cpart[((size_t)j * ncb + cb) * 8 + (((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1))] = tot;
    }
    if (j > 0) qsb_mix<NQ, 1>(L, op.q0, m, lane);  // L_{j-1} = g_j^dagger L_j
  }
}

// The threads of one instance (ncb warps), after a barrier: f and the gradient from the partial traces.
// Gradient is +inf when S == 0 (qsrs/src/utils.rs:144-146).
template <int NQ>
QSB_DEVICE double qsb_hs_finish(const QsbDevProgram& P, const cplx* gtab, const double* tr_part, const double* cpart, int ncb,
                                int tid, int nthreads, double* grad) {
  constexpr int N = 1 << NQ;
  double sr = 0, si = 0;
  for (int w = 0; w < ncb; ++w) { sr += tr_part[2 * w]; si += tr_part[2 * w + 1]; }
  const double mag = sqrt(sr * sr + si * si);
  const int n_slots = P.n_ops + P.n_params;
  for (int s = tid; s < n_slots; s += nthreads) {
    const int code = P.slotmap[s];
    if (code < 0 || (code & 3) == 0) continue;
    const int j = code >> 2;
    const qsb_op op = P.ops[j];
    double c[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = 0.0;
    for (int w = 0; w < ncb; ++w)
#pragma unroll
      for (int i = 0; i < 8; ++i) c[i] += cpart[((size_t)j * ncb + w) * 8 + i];
    const cplx* dg = gtab + 4 * s;
    double kr = 0, ki = 0;  // S_k (up to conjugation, which the gradient formula is blind to)
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      kr = fma(dg[e].x, c[2 * e], fma(-dg[e].y, c[2 * e + 1], kr));
      ki = fma(dg[e].x, c[2 * e + 1], fma(dg[e].y, c[2 * e], ki));
    }
    const double num = sr * kr + si * ki;
    grad[op.param_off + (code & 3) - 1] = (mag == 0.0) ? __longlong_as_double(0x7ff0000000000000LL) : -num / (N * mag);
  }
  return 1.0 - mag / N;
}

// ---------------------------------------------------------------------------- batch kernel (QSB_RUN_HS)
// One CTA = the N/8 warps of ONE instance: the instance comes from blockIdx alone and the warp index only selects the column
// block, so the compiler sees uniform control flow and emits plain SHFLs (with several instances per CTA every shuffle of
// the trace reductions was wrapped in WARPSYNC.COLLECTIVE / ENDCOLLECTIVE; see qsb_k_warp_jac).
template <int NQ>
struct QsbHsGeom {
  static constexpr int NCB = (1 << NQ) / 8;  // column blocks = warps per instance
  static constexpr int IPC = 1;              // instances per CTA
};

QSB_HOST_DEVICE size_t qsb_hs_slot_doubles(int nq, int max_ops, int max_params) {
  const size_t ncb = ((size_t)1 << nq) / 8;
  return 8 * (size_t)(max_ops + max_params) + 2 * (size_t)max_params + 2 * ncb + 8 * (size_t)max_ops * ncb + 2 +
         3 * (size_t)(max_ops + (max_ops & 1));  // ... + the staged ops
}

template <int NQ>
QSB_GLOBAL __launch_bounds__(32 * QsbHsGeom<NQ>::NCB) qsb_k_warp_hs(QsbBatchArgs a) {
  using G = QsbHsGeom<NQ>;
  QSB_DYN_SMEM(smem_raw);
  const int lane = threadIdx.x & 31;
  const int sub = 0, cb = G::NCB == 1 ? 0 : (int)(threadIdx.x >> 5), tid_i = threadIdx.x, nth = G::NCB * 32;
  size_t slot = qsb_hs_slot_doubles(NQ, a.max_ops, a.max_params);
  slot = (slot + 1) & ~(size_t)1;
  cplx* gtab = (cplx*)((double*)smem_raw + (size_t)sub * slot);
  double* trig = (double*)(gtab + (size_t)4 * (a.max_ops + a.max_params));
  double* trp = trig + 2 * a.max_params;
  double* cpart = trp + 2 * G::NCB;
  // the ops of the program worked on last, in shared memory: loads through the program's own (generic) pointer count as
  // divergent for the compiler, loads from shared memory at a uniform address do not
  qsb_op* sops = (qsb_op*)(cpart + 8 * (size_t)a.max_ops * G::NCB + 2);
  int staged = -1;
  const long long p_base = a.param_off[a.first];
  const int per_round = gridDim.x * G::IPC;
  for (int base = 0; base < a.count; base += per_round) {
    const int li = base + qsb_cta_first_instance(a.group);  // IPC = 1
    const bool valid = li < a.count;
    const int inst = a.first + (valid ? li : 0);
    const int pi = a.prog_idx[inst];
    QsbDevProgram P = a.progs[pi];
    if (pi != staged) {  // the previous round ended with a barrier: nobody reads the old ops any more
      for (int e = tid_i; e < P.n_ops * 6; e += nth) ((int32_t*)sops)[e] = ((const int32_t*)P.ops)[e];
      staged = pi;
    }
    P.ops = sops;
    if (valid) {
      const double* x = a.params + a.param_off[inst];
      for (int p = tid_i; p < P.n_params; p += nth) {
        double s, c;
        sincos(P.pscale[p] * x[p], &s, &c);
        trig[2 * p] = c; trig[2 * p + 1] = s;
      }
    }
    __syncthreads();
    if (valid) qsb_fill_gates(P, trig, gtab, tid_i, nth);
    __syncthreads();
    if (valid) qsb_warp_hs_cols<NQ>(P, a.target, gtab, cb, G::NCB, lane, trp + 2 * cb, cpart);
    __syncthreads();
    if (valid) {
      const double f = qsb_hs_finish<NQ>(P, gtab, trp, cpart, G::NCB, tid_i, nth, a.out_b + (a.param_off[inst] - p_base));
      if (tid_i == 0) a.out_a[li] = f;
    }
    __syncthreads();
  }
}
