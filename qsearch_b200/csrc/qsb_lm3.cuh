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
  static constexpr int WARPS = NQ >= 5 ? 8 : 4;  // warps per CTA (= per instance)
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
  cplx* gtab = (cplx*)(part + 16 + (a.max_params & 1));  // [max_ops + max_params][4], 16-byte aligned
  int* seg = (int*)(gtab + (size_t)4 * (a.max_ops + a.max_params));  // NW + 1 (<= 12 ints), then [12] = queue slot, [13] = ok flag
  qsb_op* sops = (qsb_op*)(seg + 16);       // the program's ops, staged once per instance
  double* Ls = (double*)(sops + a.max_ops); // packed Cholesky factor; during the sweep the same bytes hold V
  Ls = (double*)(((uintptr_t)Ls + 15) & ~(uintptr_t)15);
  cplx* vsm = (cplx*)Ls;
  const size_t tri = (size_t)(a.max_params + 1) * (a.max_params + 2) / 2;
  double* rinv = Ls + (tri > (size_t)2 * N * N ? tri : (size_t)2 * N * N);
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
        if (kzero) { for (int e = threadIdx.x; e < rows * rows; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
        qsb_syrk_dmma(Rg, rows, K2, A, warp, NW, lane);
        __syncthreads();
        QSB_PROF_END(1);
      }
      for (int k = threadIdx.x; k < P; k += blockDim.x) {
        g[k] = QSB_LDCG(&A[(size_t)P * rows + k]);
        const double akk = QSB_LDCG(&A[(size_t)k * rows + k]);
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
            const double akk = QSB_LDCG(&A[(size_t)k * rows + k]);
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
        { QSB_PROF_BEGIN(); chol_ok = qsb_chol_packed(A, rows, d2, mu, P, Ls, rinv, wk, okflag, a.prof); QSB_PROF_END(2); }
        if (!chol_ok) {
          mu *= nu; nu *= 2.0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
          continue;
        }
        { QSB_PROF_BEGIN(); qsb_chol_back_packed(Ls, P, wk, h); QSB_PROF_END(3); }
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
            if (kzero) { for (int e = threadIdx.x; e < rows * rows; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
            qsb_syrk_dmma(Rg, rows, K2, A, warp, NW, lane);
            __syncthreads();
            QSB_PROF_END(1);
          }
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            g[k] = QSB_LDCG(&A[(size_t)P * rows + k]);
            d2[k] = fmax(d2[k], QSB_LDCG(&A[(size_t)k * rows + k]));
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
  return base + (packed + QSB_CH_NB + 2) * 8;
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
    if (qsb_syrk_needs_zero(rows, NW)) { for (int e = threadIdx.x; e < rows * rows; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
    qsb_syrk_dmma(Rg, rows, K2, A, warp, NW, lane);
    __syncthreads();
    double* og = a.out_b + (a.param_off[inst] - p_base);
    double* oA = a.out_c + (sq_off[inst] - s_base);
    for (int k = threadIdx.x; k < P; k += blockDim.x) og[k] = QSB_LDCG(&A[(size_t)P * rows + k]);
    for (int e = threadIdx.x; e < P * P; e += blockDim.x) oA[e] = QSB_LDCG(&A[(size_t)(e / P) * rows + (e % P)]);
    if (threadIdx.x == 0) a.out_a[li] = 2.0 * F;
  }
}

static inline size_t qsb_lmsys3_smem_bytes(int nq, int max_ops, int max_params) {
  const size_t N = (size_t)1 << nq;
  return ((size_t)2 * max_params + 16 + 8 * (size_t)(max_ops + max_params) + 2 * N * N) * 8 + 16 * 4 + (size_t)24 * max_ops + 64;
}
