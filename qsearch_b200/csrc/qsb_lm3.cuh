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

// plain 16-byte load: the slab and the snapshots are written and read by ONE CTA (block-level visibility after a barrier
// holds for cached loads), so the L1 may keep the rows that several Gram tiles of the SM share
#define QSB_LD2(p) (*reinterpret_cast<const double2*>(p))

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
    } else if (op.opcode == QSB_OP_U3 || op.opcode == QSB_OP_XZXZ) {
      // The two leaves of the search gatesets have closed forms (w = sqrt(mu) conj(v), mu = 1 throughout):
      //   U3 = diag(1, e^{i phi}) Ry(theta) diag(1, e^{i lambda}):  lambda: -i g^dagger dg = P1            v = e1, gamma = 0
      //        phi: g^dagger P1 g, v = g^dagger e1 (w = row 1 of g), gamma = 0;   theta: -1/2 R^dagger Y R, R = diag(1, e^{i lambda}):
      //        v = (1, -i e^{-i lambda}) / sqrt 2, gamma = -1/2
      //   XZXZ = Rz(b) SX Rz(a) SX:  b: -1/2 g^dagger Z g = -1/2 I + g^dagger P1 g, w = row 1 of g, gamma = -1/2;
      //        a: -1/2 SX^dagger Z SX = -1/2 I + (SX^dagger e1)(..)^dagger, w = row 1 of SX = ((1-i)/2, (1+i)/2), gamma = -1/2
      const double* tg = trig + 2 * op.param_off;
      const int which = code & 3;
      const cplx z = cmk(0, 0);
      g[3] = z;
      if (op.opcode == QSB_OP_U3) {
        if (which == 1) { const double k = 0.70710678118654752440; g[0] = cmk(k, 0); g[1] = cmk(-k * tg[5], k * tg[4]); g[2] = cmk(-0.5, 0); }
        else if (which == 2) { cplx g0[4]; qsb_gate2x2(op.opcode, tg, 0, g0); g[0] = g0[2]; g[1] = g0[3]; g[2] = z; }
        else { g[0] = z; g[1] = cmk(1, 0); g[2] = z; }
      } else {
        if (which == 1) { g[0] = cmk(0.5, -0.5); g[1] = cmk(0.5, 0.5); g[2] = cmk(-0.5, 0); }
        else { cplx g0[4]; qsb_gate2x2(op.opcode, tg, 0, g0); g[0] = g0[2]; g[1] = g0[3]; g[2] = cmk(-0.5, 0); }
      }
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
    if constexpr (TW == 2) {
      // software pipeline: the 16-column chunk m0 + 16 is in flight (L2 latency) while chunk m0 feeds the tensor cores
      double2 av[2][TW][2], bv[2][TW][2];
#pragma unroll
      for (int i = 0; i < TW; ++i) {
        if (i < ni) { av[0][i][0] = QSB_LD2(pa[i]); av[0][i][1] = QSB_LD2(pa[i] + 2); }
        else { av[0][i][0] = av[0][i][1] = make_double2(0.0, 0.0); }
        bv[0][i][0] = QSB_LD2(pb[i]); bv[0][i][1] = QSB_LD2(pb[i] + 2);
      }
#pragma unroll 1
      for (int m0 = 0; m0 < klen; m0 += 32) {
#pragma unroll
        for (int ph = 0; ph < 2; ++ph) {
          const int mc = m0 + 16 * ph, mn = mc + 16;
          if (mc < klen) {
            if (mn < klen) {
#pragma unroll
              for (int i = 0; i < TW; ++i) {
                if (i < ni) { av[ph ^ 1][i][0] = QSB_LD2(pa[i] + mn); av[ph ^ 1][i][1] = QSB_LD2(pa[i] + mn + 2); }
                else { av[ph ^ 1][i][0] = av[ph ^ 1][i][1] = make_double2(0.0, 0.0); }
                bv[ph ^ 1][i][0] = QSB_LD2(pb[i] + mn); bv[ph ^ 1][i][1] = QSB_LD2(pb[i] + mn + 2);
              }
            }
#pragma unroll
            for (int s4 = 0; s4 < 4; ++s4)
#pragma unroll
              for (int i = 0; i < TW; ++i) {
                if (i >= ni) continue;  // warp-uniform
#pragma unroll
                for (int j = 0; j < TW; ++j) {
                  if (diag && j > i) continue;  // only the lower 8x8 tiles of a diagonal block
                  const double x = (s4 & 1) ? av[ph][i][s4 >> 1].y : av[ph][i][s4 >> 1].x, y = (s4 & 1) ? bv[ph][j][s4 >> 1].y : bv[ph][j][s4 >> 1].x;
                  qsb_dmma884(acc[i][j][0], acc[i][j][1], x, y);
                }
              }
          }
        }
      }
    } else {  // 32-wide blocks: the second operand buffer would not fit the register file; two chunks per trip instead
#pragma unroll 2
      for (int m0 = 0; m0 < klen; m0 += 16) {
        double2 av[TW][2], bv[TW][2];
#pragma unroll
        for (int i = 0; i < TW; ++i) {
          if (i < ni) { av[i][0] = QSB_LD2(pa[i] + m0); av[i][1] = QSB_LD2(pa[i] + m0 + 2); }
          else { av[i][0] = av[i][1] = make_double2(0.0, 0.0); }
          bv[i][0] = QSB_LD2(pb[i] + m0); bv[i][1] = QSB_LD2(pb[i] + m0 + 2);
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

// 16-wide blocks throughout (the L1 serves the rows that tiles share, so wider blocks buy nothing and cost a third of the
// register file); the contraction is cut in two when that fills the warps better (two addends per element: deterministic)
QSB_DEVICE int qsb_syrk_ksplit3(int rows, int nwarps) {
  const int nb = (rows + 15) / 16, tasks = nb * (nb + 1) / 2;
  const int r1 = (tasks + nwarps - 1) / nwarps, r2 = (2 * tasks + nwarps - 1) / nwarps;
  return (r2 + 1 < 2 * r1) ? 2 : 1;  // half-length rounds: r2 / 2 against r1, plus a little for the atomics
}
QSB_DEVICE void qsb_syrk_packed(const double* R, int rows, int K2, double* A, int warp, int nwarps, int lane) {
  qsb_syrk_packed_t<2>(R, rows, K2, A, warp, nwarps, lane, qsb_syrk_ksplit3(rows, nwarps));
}

// ---------------------------------------------------------------------------- Gram product on the TF32 tensor path
// J_r^T J_r only steers the optimiser (it is the Gauss-Newton model of the Hessian); the gradient J_r^T r, the objective and
// everything the solver returns stay fp64.  With the rows rounded to TF32 (10-bit mantissa) and fp32 accumulation the LM
// iteration counts and end points are unchanged within their seed-to-seed noise (NumPy model of this kernel's update rules on
// reachable and unreachable 3- and 4-qubit problems: 17-30 vs 17-30 iterations, 244-532 vs 281-528; BF16 rows double the
// count on reachable targets and are not used), while mma.sync.m16n8k8.tf32 runs several times faster than the FP64 DMMA
// that dominates an iteration at n = 5.  Opt-in (qsb_solver_opts.gram_tf32); the search front-end turns it on.
#ifdef QSB_EMU
static inline float qsb_to_tf32(double x) {
  float f = (float)x;
  uint32_t u; std::memcpy(&u, &f, 4);
  u = (u + 0x1000u) & 0xFFFFE000u;
  std::memcpy(&f, &u, 4);
  return f;
}
// emulated mma.sync.m16n8k8.tf32: c[16x8] += a[16x8] b[8x8]; fragments as in the PTX ISA (g = lane >> 2, t = lane & 3):
// a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4);  b0 (k = t, n = g) b1 (k = t+4, n = g);  c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1)
static inline void qsb_mma_tf32(float (&c)[4], float a0, float a1, float a2, float a3, float b0, float b1) {
  auto* st = qsb_emu::cur();
  const int w = threadIdx.x / 32, l = threadIdx.x % 32, g = l >> 2, t = l & 3;
  auto pack = [](float x, float y) { uint64_t r; float v[2] = {x, y}; std::memcpy(&r, v, 8); return r; };
  auto unpack = [](uint64_t r, int i) { float v[2]; std::memcpy(v, &r, 8); return v[i]; };
  st->warp_buf[w][l] = pack(a0, a1); st->warp_buf2[w][l] = pack(a2, a3);
  __syncwarp();
  float A0[8], A1[8];  // rows g and g + 8 of a
  for (int k = 0; k < 8; ++k) {
    const uint64_t r = (k < 4) ? st->warp_buf[w][4 * g + k] : st->warp_buf2[w][4 * g + k - 4];
    A0[k] = unpack(r, 0); A1[k] = unpack(r, 1);
  }
  __syncwarp();
  st->warp_buf[w][l] = pack(b0, b1);
  __syncwarp();
  for (int j = 0; j < 2; ++j) {
    const int n = 2 * t + j;
    float s0 = c[j], s1 = c[2 + j];
    for (int k = 0; k < 8; ++k) {
      const float b = unpack(st->warp_buf[w][4 * n + (k & 3)], k >> 2);
      s0 = std::fma(A0[k], b, s0); s1 = std::fma(A1[k], b, s1);
    }
    c[j] = s0; c[2 + j] = s1;
  }
  __syncwarp();
}
#else
__device__ __forceinline__ float qsb_to_tf32(double x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"((float)x));
  return __uint_as_float(u);
}
__device__ __forceinline__ void qsb_mma_tf32(float (&c)[4], float a0, float a1, float a2, float a3, float b0, float b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(__float_as_uint(a0)), "r"(__float_as_uint(a1)), "r"(__float_as_uint(a2)), "r"(__float_as_uint(a3)),
                 "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}
#endif

// A[TRI(k, l)] = sum_m R[k][m] R[l][m], l <= k < rows, R = TF32-rounded floats [rows][K2] (K2 a multiple of 16), result fp64
// packed.  One warp per 32 x 32 block of the lower triangle.  The contraction order is free as long as both operands agree:
// lane t takes 4 consecutive columns of a 16-column chunk as one 16-byte load and feeds them to two MMAs.
QSB_DEVICE void qsb_gram_tf32(const float* R, int rows, int K2, double* A, int warp, int nwarps, int lane) {
  const int g = lane >> 2, t = lane & 3;
  const int nb = (rows + 31) / 32, ntasks = nb * (nb + 1) / 2;
  for (int task = warp; task < ntasks; task += nwarps) {
    int bk = (int)((sqrt(8.0 * task + 1.0) - 1.0) * 0.5);
    while ((bk + 1) * (bk + 2) / 2 <= task) ++bk;
    while (bk * (bk + 1) / 2 > task) --bk;
    const int bl = task - bk * (bk + 1) / 2;
    const float* pa[2][2];
    const float* pb[4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        int r = 32 * bk + 16 * i + 8 * h + g;
        if (r > rows - 1) r = rows - 1;
        pa[i][h] = R + (size_t)r * K2 + 4 * t;
      }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int r = 32 * bl + 8 * j + g;
      if (r > rows - 1) r = rows - 1;
      pb[j] = R + (size_t)r * K2 + 4 * t;
    }
    float acc[2][4][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;
#pragma unroll 2
    for (int m0 = 0; m0 < K2; m0 += 16) {
      float4 av[2][2], bv[4];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int h = 0; h < 2; ++h) av[i][h] = *reinterpret_cast<const float4*>(pa[i][h] + m0);
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = *reinterpret_cast<const float4*>(pb[j] + m0);
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          qsb_mma_tf32(acc[i][j], av[i][0].x, av[i][1].x, av[i][0].y, av[i][1].y, bv[j].x, bv[j].y);
          qsb_mma_tf32(acc[i][j], av[i][0].z, av[i][1].z, av[i][0].w, av[i][1].w, bv[j].z, bv[j].w);
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int k = 32 * bk + 16 * i + 8 * (e >> 1) + g, l = 32 * bl + 8 * j + 2 * t + (e & 1);
          if (k < rows && l <= k) A[QSB_TRI(k, l)] = (double)acc[i][j][e];
        }
  }
}

// ---------------------------------------------------------------------------- factorisation
// Same augmented, packed, right-looking Cholesky as qsb_chol_packed (qsb_lm2.cuh): B = [A + mu diag(d2), g ; g^T, .],
// rows = P + 1, row P ends up as y with L y = g.  What changed, phase by phase (cycles at P = 62, profiles/r2):
//   copy      the matrix arrives packed (qsb_syrk_packed): linear, vectorised copy into shared memory;
//   diagonal  16x16 block by one warp in registers + shuffles (qsb_chp_diag's chain), also written DENSE (17-stride) so
//             that the panel reads it with broadcast loads and no triangular index arithmetic;
//   panel     one row per thread, column-oriented substitution (16 dependent steps, not 136);
//   trailing  16x16 DMMA tiles, 32-bit index arithmetic (the packed index of a 272-row matrix fits easily);
//   inverse   W_b = L_bb^-1 of every diagonal block, computed by an otherwise idle warp OFF the critical path; the
//             back-substitution then is one 16x16 matrix-vector product per block instead of one reduction per row.
#define QSB_CH3_WSZ 256  // doubles per inverted diagonal block
#define QSB_CH4_DLD 17
#define QSB_CH4_DSZ (QSB_CH_NB * QSB_CH4_DLD + QSB_CH_NB)  // dense block + 1 / L_jj
#define QSB_TRI32(i, j) ((((i) * ((i) + 1)) >> 1) + (j))

// diagonal block [kb, kb + nbk): lane = row, the row in registers.  Leaves L in Ls and, dense, in Dk (identity in the
// padding of a short block), 1 / L_jj in Dk[NB * DLD + j].
#ifdef QSB_EMU
static inline double qsb_rcp_pos(double a) { return 1.0 / a; }
#else
// 1 / a for a positive, normal a (the caller has excluded everything else; garbage in, garbage out, no trap): hardware seed and
// two Newton steps, the fast path of the library's division without its special-case branch
__device__ __forceinline__ double qsb_rcp_pos(double a) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  y = fma(fma(-a, y, 1.0), y, y);
  return fma(fma(-a, y, 1.0), y, y);
}
#endif

QSB_DEVICE void qsb_ch4_diag(double* Ls, int kb, int nbk, double* Dk, int* okflag, int lane) {
  constexpr int NB = QSB_CH_NB, LD = QSB_CH4_DLD;
  const int base = QSB_TRI32(kb + lane, kb);
  double d[NB];
#pragma unroll
  for (int j = 0; j < NB; ++j) d[j] = (lane < nbk && j <= lane) ? Ls[base + j] : ((j == lane) ? 1.0 : 0.0);
  // Sixteen dependent pivot steps are the critical path of the whole factorisation, so the step is arranged for a short chain:
  // the columns stay UNNORMALISED (c_ij = L_ij L_jj) while the block is eliminated, d_ik -= c_ij c_kj / p_j, which puts only
  // pivot shuffle -> reciprocal -> scale -> update on the chain; the column shuffles do not wait for the reciprocal, and the
  // sixteen square roots are taken afterwards, independently of each other.  No early exit: a non-positive pivot only clears the
  // flag and the remaining steps run on garbage that nobody reads, so the steps are straight-line code.
  double piv[NB];
  bool good = true;
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    piv[j] = __shfl_sync(QSB_FULL, d[j], j);
    good = good && (piv[j] > 0.0) && (piv[j] < 1e300);
    const double t = d[j] * qsb_rcp_pos(piv[j]);
#pragma unroll
    for (int k = j + 1; k < NB; ++k) d[k] = fma(-t, __shfl_sync(QSB_FULL, d[j], k), d[k]);
  }
#pragma unroll
  for (int j = 0; j < NB; ++j) {
    const double inv = rsqrt(piv[j]);
    if (lane == j) Dk[NB * LD + j] = inv;
    d[j] *= inv;  // lane j: p_j / sqrt(p_j) = L_jj
  }
  if (!good && lane == 0) *okflag = 0;
  if (lane < NB) {
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      if (lane < nbk && j <= lane) Ls[base + j] = d[j];
      Dk[lane * LD + j] = (j <= lane) ? d[j] : 0.0;
    }
  }
}

// rows R0 .. R0 + nrow - 1 of the panel below the block: x L_kk^T = a, one row per thread; after x_j is known it is
// folded into the columns to its right, so the dependent chain is 16 multiply + FMA pairs
QSB_DEVICE void qsb_ch4_panel(double* Ls, int R0, int kb, int nbk, int nrow, const double* Dk) {
  constexpr int NB = QSB_CH_NB, LD = QSB_CH4_DLD;
  for (int ri = threadIdx.x; ri < nrow; ri += blockDim.x) {
    double* src = Ls + QSB_TRI32(R0 + ri, kb);
    double x[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) x[j] = (j < nbk) ? src[j] : 0.0;
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const double xj = x[j] * Dk[NB * LD + j];
      x[j] = xj;
#pragma unroll
      for (int c = j + 1; c < NB; ++c) x[c] = fma(-xj, Dk[c * LD + j], x[c]);
    }
#pragma unroll
    for (int j = 0; j < NB; ++j) if (j < nbk) src[j] = x[j];
  }
}

// W = L_kk^-1 (row-major 16x16, zero above the diagonal) from the dense block; one warp, lane = column of W
QSB_DEVICE void qsb_ch4_inverse(const double* Dk, double* Wb, int lane) {
  constexpr int NB = QSB_CH_NB, LD = QSB_CH4_DLD;
  if (lane < NB) {
    double w[NB];
#pragma unroll
    for (int i = 0; i < NB; ++i) w[i] = (i == lane) ? 1.0 : 0.0;
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const double wj = w[j] * Dk[NB * LD + j];
      w[j] = wj;
#pragma unroll
      for (int i = j + 1; i < NB; ++i) w[i] = fma(-Dk[i * LD + j], wj, w[i]);
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) Wb[i * NB + lane] = w[i];
  }
}

// one 16x16 tile (bi, bj) of the trailing update C -= X X^T behind panel [kb, kb + nbk); the trailing matrix starts at R0
QSB_DEVICE void qsb_ch4_tile(double* Ls, int R0, int kb, int nbk, int nrow, int bi, int bj, int lane) {
  constexpr int NB = QSB_CH_NB;
  const int g = lane >> 2, t = lane & 3;
  double acc[2][2][2];
  int pa[2], pb[2], rowb[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int r = 16 * bi + 8 * i + g;
    int ra = r, rb = 16 * bj + 8 * i + g;
    if (ra > nrow - 1) ra = nrow - 1;
    if (rb > nrow - 1) rb = nrow - 1;
    pa[i] = QSB_TRI32(R0 + ra, kb) + t;
    pb[i] = QSB_TRI32(R0 + rb, kb) + t;
    rowb[i] = QSB_TRI32(R0 + r, R0);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int r = 16 * bi + 8 * i + g, c = 16 * bj + 8 * j + 2 * t + e;
        acc[i][j][e] = (r < nrow && c <= r) ? Ls[rowb[i] + c] : 0.0;
      }
#pragma unroll
  for (int ks = 0; ks < NB / 4; ++ks) {
    const bool in = 4 * ks + t < nbk;
    const double a0 = in ? -Ls[pa[0] + 4 * ks] : 0.0, a1 = in ? -Ls[pa[1] + 4 * ks] : 0.0;
    const double b0 = in ? Ls[pb[0] + 4 * ks] : 0.0, b1 = in ? Ls[pb[1] + 4 * ks] : 0.0;
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
        if (r < nrow && c <= r) Ls[rowb[i] + c] = acc[i][j][e];
      }
}

// Apk: packed Gram matrix of [Q ; -C] in global memory (rows = P + 1).  Dk: 2 * QSB_CH4_DSZ doubles, Winv:
// ceil(P / 16) * 256 doubles of shared memory.  Needs >= 3 warps.  Returns false (uniformly) if a pivot is not positive.
QSB_DEVICE bool qsb_chol4(const double* Apk, int rows, const double* d2, double mu, int P, double* Ls, double* Dk, double* Winv,
                          int* okflag, long long* prof) {
#define QSB_CH_T(slot) do { if (prof && threadIdx.x == 0) { const long long _n = QSB_CLOCK(); atomicAdd((unsigned long long*)&prof[slot], (unsigned long long)(_n - _tc)); _tc = _n; } } while (0)
  long long _tc = prof ? QSB_CLOCK() : 0ll;
  constexpr int NB = QSB_CH_NB;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index(), nwarps = blockDim.x >> 5;
  const int tri = rows * (rows + 1) / 2;
  __syncthreads();
  for (int e = 2 * threadIdx.x; e + 1 < tri; e += 2 * blockDim.x) *reinterpret_cast<double2*>(Ls + e) = QSB_LDCG2(Apk + e);
  if ((tri & 1) && threadIdx.x == 0) Ls[tri - 1] = QSB_LDCG(Apk + tri - 1);
  if (threadIdx.x == 0) *okflag = 1;
  __syncthreads();
  for (int k = threadIdx.x; k < P; k += blockDim.x) Ls[QSB_TRI32(k, k)] += mu * d2[k];
  __syncthreads();
  QSB_CH_T(6);
  if (warp == 0) qsb_ch4_diag(Ls, 0, (P < NB) ? P : NB, Dk, okflag, lane);
  __syncthreads();
  QSB_CH_T(7);
  for (int kb = 0; kb < P; kb += NB) {
    const int blk = kb / NB, nbk = (P - kb < NB) ? P - kb : NB;
    double* Dc = Dk + (blk & 1) * QSB_CH4_DSZ;
    if (!*okflag) { __syncthreads(); return false; }
    const int R0 = kb + nbk, nrow = rows - R0;
    qsb_ch4_panel(Ls, R0, kb, nbk, nrow, Dc);
    __syncthreads();
    QSB_CH_T(8);
    if (R0 < P) {
      const int nbt = (nrow + 15) >> 4;
      for (int bi = warp; bi < nbt; bi += nwarps) qsb_ch4_tile(Ls, R0, kb, nbk, nrow, bi, 0, lane);  // first block column
      __syncthreads();
      const int nbk1 = (P - R0 < NB) ? P - R0 : NB;
      if (warp == 0) qsb_ch4_diag(Ls, R0, nbk1, Dk + ((blk + 1) & 1) * QSB_CH4_DSZ, okflag, lane);  // next diagonal block ...
      else if (warp == 1) qsb_ch4_inverse(Dc, Winv + (size_t)blk * QSB_CH3_WSZ, lane);               // ... this block's inverse ...
      else {                                                                                         // ... under the rest of the update
        const int nw = nwarps - 2, w = warp - 2;
        int bi = 1, bj = 1;
        for (int skip = w; skip > 0; --skip) { if (++bj > bi) { ++bi; bj = 1; } }
        while (bi < nbt) {
          qsb_ch4_tile(Ls, R0, kb, nbk, nrow, bi, bj, lane);
          for (int skip = nw; skip > 0; --skip) { if (++bj > bi) { ++bi; bj = 1; } }
        }
      }
      __syncthreads();
    } else {
      if (warp == 1) qsb_ch4_inverse(Dc, Winv + (size_t)blk * QSB_CH3_WSZ, lane);
      __syncthreads();
    }
    QSB_CH_T(9);
  }
  if (!*okflag) { __syncthreads(); return false; }
  return true;
#undef QSB_CH_T
}

// h = -(L^T)^-1 y, y = row P of the packed factor: one warp, one 16-row block per step.  With W_b = L_bb^-1:
//   h_b = W_b^T z_b,   z = -y - sum over later blocks c of L_cb^T h_c  (folded into z as soon as h_c is known)
// zs: 32 doubles of shared memory.
QSB_DEVICE void qsb_ch4_back(const double* Ls, int P, const double* Winv, double* zs, double* h) {
  constexpr int NB = QSB_CH_NB;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  __syncthreads();
  if (warp == 0) {
    constexpr int NR = (QSB_CH_MAXROWS + 31) / 32;
    const double* y = Ls + QSB_TRI32(P, 0);
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
      if (li >= 0 && li < NB) {
        double hb = 0.0;
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
        if (32 * b < kb) {  // warp-uniform
          double s = z[b];
          if (i < kb) {
            int rb = QSB_TRI32(kb, i);
            for (int j = 0; j < nbk; ++j) { s = fma(-Ls[rb], zs[NB + j], s); rb += kb + j + 1; }
          }
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
template <int NQ, int KB, bool TF32 = false>
QSB_DEVICE void qsb_q_param(const QsbFrag<(1 << NQ) / 4, (1 << NQ) / 8>& X, const cplx* slot, double* row, int lane,
                            float* row32 = nullptr, const double* crow = nullptr, double* gk = nullptr) {
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
  if (TF32) {
    // TF32 Gram mode: the row is stored as TF32-rounded floats (same packed order) and its inner product with the -C row,
    // g_k = (J_r^T r)[k], is taken here from the fp64 accumulators: the gradient never sees the reduced precision
    double dot = 0.0;
#pragma unroll
    for (int tt = 0; tt < TT; ++tt) {
      const double2 c = *reinterpret_cast<const double2*>(crow + tt * 64 + 2 * lane);
      dot = fma(zd[tt][0], c.x, fma(zd[tt][1], c.y, dot));
      *reinterpret_cast<float2*>(row32 + tt * 64 + 2 * lane) = make_float2(qsb_to_tf32(zd[tt][0]), qsb_to_tf32(zd[tt][1]));
    }
    if (NOFF > 0) {
#pragma unroll
      for (int o = 0; o < NOFF; ++o) {
        const int off = TT * 64 + o * 128 + 2 * lane;
        const double2 c0 = *reinterpret_cast<const double2*>(crow + off), c1 = *reinterpret_cast<const double2*>(crow + off + 64);
        const double a0 = r2 * hr[o][0], a1 = r2 * hr[o][1], b0 = r2 * hi[o][0], b1 = r2 * hi[o][1];
        dot = fma(a0, c0.x, fma(a1, c0.y, fma(b0, c1.x, fma(b1, c1.y, dot))));
        *reinterpret_cast<float2*>(row32 + off) = make_float2(qsb_to_tf32(a0), qsb_to_tf32(a1));
        *reinterpret_cast<float2*>(row32 + off + 64) = make_float2(qsb_to_tf32(b0), qsb_to_tf32(b1));
      }
    }
    for (int m = 16; m > 0; m >>= 1) dot += __shfl_xor_sync(QSB_FULL, dot, m);
    if (lane == 0) *gk = dot;
    return;
  }
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

// ---------------------------------------------------------------------------- forward pass, one matrix row per lane
// X <- g_j X from T^dagger with the N columns of X dealt to the warps of the CTA and lane = row: a one-qubit gate is one
// __shfl_xor of the partner row and one 2x2 combination per element, a CNOT one shuffle and a select - no bit juggling,
// a few hundred cycles per op for the whole CTA.  This IS the trial evaluation of an LM step (it ends with V = U T^dagger,
// |V - I|^2 and Tr V), and on the way it leaves a snapshot of X in front of every op that has parameters
// (snap + ordinal * 2 N^2 doubles, TRANSPOSED: [col][row] complex, so the stores of a warp are contiguous).  If the step
// is accepted, the product phase below reads the snapshots: no warp ever recomputes a prefix.
#define QSB_ROWS_PART 32  // doubles of shared memory for the per-warp partial sums of qsb_cta_eval_rows (3 per active warp)
template <int NQ, int NW = 8>
struct QsbRowGeom {
  static constexpr int N = 1 << NQ;
  static constexpr int SLOTS = 32 / N;                          // columns that fit side by side in one warp (N <= 32)
  static constexpr int CW = (N / NW > SLOTS) ? N / NW : SLOTS;  // columns per warp (8 warps): n = 5: 4, n = 4: 2, n = 3: 4
  static constexpr int EPL = CW / SLOTS;                        // elements per lane
  static constexpr int AW = N / CW;                             // warps that take part (n = 3: 2 of the 8)
};

template <int NQ, int NW = 8>
QSB_DEVICE void qsb_warp_forward_rows(const QsbDevProgram& P, const double* target, const cplx* gtab, int warp, int lane,
                                      double* snap, cplx* vout, double& norm2, cplx& tr) {
  using R = QsbRowGeom<NQ, NW>;
  constexpr int N = R::N, EPL = R::EPL;
  const int r = lane & (N - 1), c0 = R::CW * warp + (lane >> NQ);
  double xr[EPL], xi[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) {  // X = T^dagger
    const int c = c0 + R::SLOTS * e;
    const double2 tv = *reinterpret_cast<const double2*>(target + 2 * (c * N + r));
    xr[e] = tv.x; xi[e] = -tv.y;
  }
  int pidx = 0;
  for (int j = 0; j < P.n_ops; ++j) {
    const qsb_op op = P.ops[j];
    if (op.opcode == QSB_OP_CNOT) {
      const int mask = 1 << (NQ - 1 - op.q1);
      const bool ctrl = (r >> (NQ - 1 - op.q0)) & 1;
#pragma unroll
      for (int e = 0; e < EPL; ++e) {
        const double pr = __shfl_xor_sync(QSB_FULL, xr[e], mask), pi = __shfl_xor_sync(QSB_FULL, xi[e], mask);
        if (ctrl) { xr[e] = pr; xi[e] = pi; }
      }
      continue;
    }
    if (qsb_num_params(op.opcode) > 0) {
      if (snap) {
        double* sp = snap + (size_t)pidx * 2 * N * N;
#pragma unroll
        for (int e = 0; e < EPL; ++e) *reinterpret_cast<double2*>(sp + 2 * ((c0 + R::SLOTS * e) * N + r)) = make_double2(xr[e], xi[e]);
      }
      ++pidx;
    }
    const cplx* m = gtab + 4 * op.reserved;
    const int bit = NQ - 1 - op.q0, mask = 1 << bit;
    const bool hi = (r >> bit) & 1;
    const cplx ca = hi ? m[3] : m[0], cb = hi ? m[2] : m[1];
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const double ar = xr[e], ai = xi[e];
      const double pr = __shfl_xor_sync(QSB_FULL, ar, mask), pi = __shfl_xor_sync(QSB_FULL, ai, mask);
      xr[e] = fma(cb.x, pr, fma(-cb.y, pi, fma(ca.x, ar, -ca.y * ai)));
      xi[e] = fma(cb.x, pi, fma(cb.y, pr, fma(ca.x, ai, ca.y * ar)));
    }
  }
  double s = 0.0;
  cplx d = cmk(0, 0);
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    const int c = c0 + R::SLOTS * e;
    const bool diag = r == c;
    const double re = xr[e] - (diag ? 1.0 : 0.0), im = xi[e];
    s = fma(re, re, s); s = fma(im, im, s);
    if (diag) { d.x += xr[e]; d.y += xi[e]; }
    if (vout) vout[r * N + c] = cmk(xr[e], xi[e]);
  }
  for (int m = 16; m > 0; m >>= 1) {
    s += __shfl_xor_sync(QSB_FULL, s, m);
    d.x += __shfl_xor_sync(QSB_FULL, d.x, m);
    d.y += __shfl_xor_sync(QSB_FULL, d.y, m);
  }
  norm2 = s; tr = d;
}

// CTA-wide: (cos, sin) table and body-frame gate table of x, the forward pass above, then F = |U T^dagger - I|_F^2 / 2 and
// Tr(U T^dagger).  snap / vsm may be null (no snapshots / V not wanted).  part: QSB_ROWS_PART doubles.  NW: warps of the CTA.
template <int NQ, int NW = 8>
QSB_DEVICE double qsb_cta_eval_rows(const QsbDevProgram& P, const double* x, const double* target, double* trig, cplx* gtab,
                                    double* part, double* snap, cplx* vsm, cplx* tr_out) {
  constexpr int AW = QsbRowGeom<NQ, NW>::AW;
  static_assert(3 * AW <= QSB_ROWS_PART, "partial-sum buffer too small");
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
  if (warp < AW) {
    double n2; cplx tr;
    qsb_warp_forward_rows<NQ, NW>(P, target, gtab, warp, lane, snap, vsm, n2, tr);
    if (lane == 0) { part[3 * warp] = n2; part[3 * warp + 1] = tr.x; part[3 * warp + 2] = tr.y; }
  }
  __syncthreads();
  double F = 0.0;
  cplx tr = cmk(0, 0);
  for (int w = 0; w < AW; ++w) { F += part[3 * w]; tr.x += part[3 * w + 1]; tr.y += part[3 * w + 2]; }
  if (tr_out) *tr_out = tr;
  return 0.5 * F;
}

// ---------------------------------------------------------------------------- product phase
// Q rows 0..P-1 and -C in row P of the slab (row-major [(P+1)][K2]) from the snapshots and V of the forward pass at the
// SAME parameter vector (gtab = its body-frame table).  The ops that have parameters are dealt round-robin to the warps;
// a warp loads the snapshot in front of its op as a DMMA B fragment with the op's qubit mapped to k-step bit 0 (the rows
// come from memory, so any row order is free), which leaves ONE code path for every qubit.  Ends with a block barrier.
template <int NQ, bool TF32 = false>
QSB_DEVICE void qsb_cta_qrows_snap(const QsbDevProgram& P, const cplx* gtab, const double* snap, const cplx* vsm, double* slab,
                                   double* gvec = nullptr) {
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, KS = G::KS, TT = G::TT, K2 = G::K2;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index(), nwarps = blockDim.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  // TF32 mode: rows are floats in the first half of the fp64 slab's bytes (row k at slab32 + k K2); the -C row keeps its fp64
  // place (row P of the fp64 layout, beyond the float rows) and is needed by every row's gradient entry, so it comes first
  double* crow = slab + (size_t)P.n_params * K2;
  float* slab32 = reinterpret_cast<float*>(slab);
  if (TF32) { qsb_q_crow<NQ>(vsm, crow); __syncthreads(); }
  int pidx = 0;
  for (int j = 0; j < P.n_ops; ++j) {
    const qsb_op op = P.ops[j];
    const int np = qsb_num_params(op.opcode);
    if (np == 0) continue;
    if (pidx % nwarps == warp) {
      const double* sp = snap + (size_t)pidx * 2 * N * N;
      const int b = NQ - 1 - op.q0;  // the gate's bit of the row index; physical bit 2 of k = 4 ks + t takes its place
      QsbFrag<KS, TT> X;
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        int k = 4 * ks + t;
        if (b != 2 && (((k >> 2) ^ (k >> b)) & 1)) k ^= 4 | (1 << b);
#pragma unroll
        for (int o = 0; o < TT; ++o) {
          const double2 v = QSB_LD2(sp + 2 * ((8 * o + g) * N + k));
          X.re[ks][o] = v.x; X.im[ks][o] = v.y;
        }
      }
      for (int k = 0; k < np; ++k) {
        const int pk = op.param_off + k;
        if (TF32) qsb_q_param<NQ, 1, true>(X, gtab + 4 * (op.reserved + 1 + k), nullptr, lane, slab32 + (size_t)pk * K2, crow, gvec + pk);
        else qsb_q_param<NQ, 1, false>(X, gtab + 4 * (op.reserved + 1 + k), slab + (size_t)pk * K2, lane);
      }
    }
    ++pidx;
  }
  if (!TF32) qsb_q_crow<NQ>(vsm, crow);
  __syncthreads();
}

// NW = 8: one CTA per SM, the latency-optimal shape for a batch that fits the GPU once.  NW = 4 (n <= 4): two CTAs per SM at
// the same 255 registers per thread; every phase of one instance is up to 2x slower, but the phases are latency-bound (one
// warp per scheduler already drives the FP64 pipe at 92 % of its peak, profiles/r2/v_dmma_occupancy.log), so two instances per
// SM overlap: the shape for batches of more than one instance per SM.
template <int NQ, bool TF32, int NW = QsbQGeom<NQ>::WARPS>
QSB_GLOBAL __launch_bounds__(NW * 32, (NW == 4 ? 2 : 1)) qsb_k_lm3(QsbSolveArgs a) {
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, K2 = G::K2;
  QSB_DYN_SMEM(smem_raw);
  cplx* red = (cplx*)smem_raw;  // 32
  double* x = (double*)(red + 32);
  double* xn = x + a.max_params;
  double* g = xn + a.max_params;
  double* h = g + a.max_params;
  double* d2 = h + a.max_params;
  double* wk = d2 + a.max_params;
  double* trig = wk + a.max_params;         // 2 * maxP
  double* part = trig + 2 * a.max_params;   // QSB_ROWS_PART
  cplx* gtab = (cplx*)(part + QSB_ROWS_PART);  // [max_ops + max_params][4]; 16-byte aligned: 8 maxP + 32 doubles precede it
  int* seg = (int*)(gtab + (size_t)4 * (a.max_ops + a.max_params));  // 16 ints: [12] = queue slot, [13] = ok flag
  qsb_op* sops = (qsb_op*)(seg + 16);       // the program's ops, staged once per instance
  // packed Cholesky factor; during the sweep the same bytes hold V.  Aligned through its OFFSET in the dynamic shared memory, not
  // through an integer cast of the pointer: the cast hides the address space from the compiler, and every access to the factor
  // (the whole Cholesky and the back-substitution) became a generic LD / ST instead of LDS / STS
  const size_t ls_off = ((size_t)((unsigned char*)(sops + a.max_ops) - (unsigned char*)smem_raw) + 15) & ~(size_t)15;
  double* Ls = (double*)((unsigned char*)smem_raw + ls_off);
  cplx* vsm = (cplx*)Ls;
  const size_t tri = (size_t)(a.max_params + 1) * (a.max_params + 2) / 2;
  double* Winv = Ls + (((tri > (size_t)2 * N * N ? tri : (size_t)2 * N * N) + 1) & ~(size_t)1);  // [ceil(maxP / 16)][256]
  double* zs = Winv + (size_t)((a.max_params + QSB_CH_NB - 1) / QSB_CH_NB) * QSB_CH3_WSZ;  // 32
  double* Dk = zs + 32;                                                                    // 2 * QSB_CH4_DSZ
  int* okflag = seg + 13;
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  double* Rg = a.scratch + (size_t)blockIdx.x * a.scratch_per_cta;
  double* A = Rg + (size_t)(a.max_params + 1) * K2;                               // packed Gram matrix
  double* snap = A + (((size_t)(a.max_params + 1) * (a.max_params + 2) / 2 + 1) & ~(size_t)1);  // [max_pops][2 N^2]
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
    const bool kzero = qsb_syrk_ksplit3(rows, NW) == 2;
    constexpr bool tf32 = TF32;
    const double* x0 = a.x0 + a.param_off[inst];
    const int max_iters = a.max_iters > 0 ? a.max_iters : 100 * (P > 0 ? P : 1);
    for (int k = threadIdx.x; k < P; k += blockDim.x) x[k] = x0[k];
    int iters = 0, status = 0;
    cplx tr;
    double F = qsb_cta_eval_rows<NQ, NW>(prog, x, a.target, trig, gtab, part, snap, vsm, &tr);  // tables, snapshots and V of x
    if (P > 0) {
      { QSB_PROF_BEGIN(); qsb_cta_qrows_snap<NQ, TF32>(prog, gtab, snap, vsm, Rg, g); QSB_PROF_END(0); }
      iters = 1;
      {
        QSB_PROF_BEGIN();
        if (tf32) {  // Gram of the P parameter rows on the TF32 path; the augmented row (g, r^T r) is fp64 and already known
          qsb_gram_tf32(reinterpret_cast<const float*>(Rg), P, K2, A, warp, NW, lane);
          for (int k = threadIdx.x; k < P; k += blockDim.x) A[QSB_TRI(P, k)] = g[k];
          if (threadIdx.x == 0) A[QSB_TRI(P, P)] = 2.0 * F;
        } else {
          if (kzero) { for (int e = threadIdx.x; e < rows * (rows + 1) / 2; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
          qsb_syrk_packed(Rg, rows, K2, A, warp, NW, lane);
        }
        __syncthreads();
        QSB_PROF_END(1);
      }
      for (int k = threadIdx.x; k < P; k += blockDim.x) {
        g[k] = QSB_LDCG(&A[QSB_TRI(P, k)]);
        const double akk = QSB_LDCG(&A[QSB_TRI(k, k)]);
        d2[k] = akk > 0.0 ? akk : 1.0;
      }
      __syncthreads();
      double mu = 1e-3, nu = a.mu_up;
      int rejects = 0, streak = 0;
      while (true) {
        if (!(F > 1e-300) || (a.stop_f > 0.0 && F <= a.stop_f)) { status = 4; break; }
        if (a.prof && threadIdx.x == 0) atomicAdd((unsigned long long*)&a.prof[10], 1ull);  // factorisation attempts
        {  // MINPACK's gtol: max cosine between r and the columns of J_r; the batch-level stop rules ride on the same reduction
          double gm = 0.0;
          const double rn = sqrt(2.0 * F);
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            const double akk = QSB_LDCG(&A[QSB_TRI(k, k)]);
            if (akk > 0.0) gm = fmax(gm, fabs(g[k]) / (sqrt(akk) * rn));
          }
          if (threadIdx.x == 0) {  // encoded as huge values: every thread sees the same maximum
            if (a.sibling_f > 0.0 && QSB_LDCG(&a.ctl[QSB_CTL_SOLVED + pidx]) != 0) gm = 1e301;
            else if (a.straggler_c > 0.0 && 2 * QSB_LDCG(&a.ctl[QSB_CTL_FINISHED]) >= a.count && QSB_LDCG(&a.ctl[QSB_CTL_MEDIAN]) > 0 &&
                     (double)iters > a.straggler_c * (double)QSB_LDCG(&a.ctl[QSB_CTL_MEDIAN]) && iters >= a.straggler_min) gm = 1e300;
          }
          gm = qsb_block_max(gm, red);
          if (gm >= 1e301) { status = 7; break; }   // another start of this circuit already reached the target
          if (gm >= 1e300) { status = 6; break; }   // straggler cap
          if (gm <= a.gtol) { status = 1; break; }
        }
        bool chol_ok;
        { QSB_PROF_BEGIN(); chol_ok = qsb_chol4(A, rows, d2, mu, P, Ls, Dk, Winv, okflag, a.prof); QSB_PROF_END(2); }
        if (!chol_ok) {
          mu *= nu; nu *= a.mu_up; streak = 0;
          if (++rejects > 60 || mu > 1e30) { status = 5; break; }
          continue;
        }
        { QSB_PROF_BEGIN(); qsb_ch4_back(Ls, P, Winv, zs, h); QSB_PROF_END(3); }
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
        { QSB_PROF_BEGIN(); Fn = qsb_cta_eval_rows<NQ, NW>(prog, xn, a.target, trig, gtab, part, snap, vsm, &tr); QSB_PROF_END(4); }  // tables, snapshots, V now belong to xn
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
          { QSB_PROF_BEGIN(); qsb_cta_qrows_snap<NQ, TF32>(prog, gtab, snap, vsm, Rg, g); QSB_PROF_END(0); }
          ++iters;
          {
            QSB_PROF_BEGIN();
            if (tf32) {
              qsb_gram_tf32(reinterpret_cast<const float*>(Rg), P, K2, A, warp, NW, lane);
              for (int k = threadIdx.x; k < P; k += blockDim.x) A[QSB_TRI(P, k)] = g[k];
              if (threadIdx.x == 0) A[QSB_TRI(P, P)] = 2.0 * F;
            } else {
              if (kzero) { for (int e = threadIdx.x; e < rows * (rows + 1) / 2; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
              qsb_syrk_packed(Rg, rows, K2, A, warp, NW, lane);
            }
            __syncthreads();
            QSB_PROF_END(1);
          }
          for (int k = threadIdx.x; k < P; k += blockDim.x) {
            g[k] = QSB_LDCG(&A[QSB_TRI(P, k)]);
            d2[k] = fmax(d2[k], QSB_LDCG(&A[QSB_TRI(k, k)]));
          }
          __syncthreads();
          const double q = 2.0 * rho - 1.0;
          double fac = fmax((a.mu_fast > 0 && streak + 1 >= a.mu_fast) ? 1.0 / 3.0 : a.mu_down, 1.0 - q * q * q);
          if (++streak <= a.mu_streak) fac = fmax(fac, 1.0);
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
    qsb_cta_eval_rows<NQ, NW>(prog, x, a.target, trig, gtab, part, nullptr, nullptr, &tr);
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
      // Every termination counts towards "half of the batch is done" (an instance stopped by the sibling rule is done as far
      // as the frontier is concerned; counting only natural finishes left the cap switched off for good in batches where most
      // instances end that way, and one uncapped start then ran its 100 P iterations: profiles/r2/zj_mu_bench.log).  The
      // reference count comes from the natural finishes among that first half, or from the first one after it if there was none.
      const int nfin = atomicAdd(&a.ctl[QSB_CTL_FINISHED], 1) + 1;
      if (status != 6 && status != 7 && (2 * nfin <= a.count + 1 || QSB_LDCG(&a.ctl[QSB_CTL_MEDIAN]) == 0))
        atomicMax(&a.ctl[QSB_CTL_MEDIAN], iters);
    }
  }
}

static inline size_t qsb_lm3_smem_bytes(int nq, int max_ops, int max_params) {
  const size_t N = (size_t)1 << nq;
  const size_t base = 32 * 16 + (size_t)8 * max_params * 8 + QSB_ROWS_PART * 8 + 8 + (size_t)64 * (max_ops + max_params) + 16 * 4 + (size_t)24 * max_ops + 16;
  const size_t packed = std::max((size_t)(max_params + 1) * (max_params + 2) / 2, 2 * N * N);
  return base + (packed + 1 + (size_t)((max_params + QSB_CH_NB - 1) / QSB_CH_NB) * QSB_CH3_WSZ + 32 + 2 * QSB_CH4_DSZ) * 8;
}

// ---------------------------------------------------------------------------- QSB_RUN_LMSYS, unitary target
// r^T r, g = J_r^T r and A = J_r^T J_r of every instance of a batch in the body-frame form above: a forward-only sweep
// and a Gram product over rows of N^2 (not 2 N^2) reals; nothing of size N^2 P leaves the CTA's L2-resident slab.
// One CTA per instance.  out_a[inst] = r^T r, out_b[param_off ..] = g, out_c[sum P_i^2 ..] = A (row-major P x P).
template <int NQ>
QSB_GLOBAL __launch_bounds__(QsbQGeom<NQ>::WARPS * 32, (NQ <= 4 ? 2 : 1)) qsb_k_lmsys3(QsbBatchArgs a, double* scratch, size_t scratch_per_cta,
                                                                    const long long* sq_off) {  // scratch per CTA: slab, packed Gram, snapshots
  using G = QsbQGeom<NQ>;
  constexpr int N = G::N, K2 = G::K2, NW = G::WARPS;
  QSB_DYN_SMEM(smem_raw);
  double* trig = (double*)smem_raw;                       // 2 * maxP
  double* part = trig + 2 * a.max_params;                 // QSB_ROWS_PART
  cplx* gtab = (cplx*)(part + QSB_ROWS_PART);
  cplx* vsm = gtab + (size_t)4 * (a.max_ops + a.max_params);  // N * N
  int* seg = (int*)(vsm + N * N);                         // 16
  qsb_op* sops = (qsb_op*)(seg + 16);
  const int lane = threadIdx.x & 31, warp = qsb_warp_index();
  double* Rg = scratch + (size_t)blockIdx.x * scratch_per_cta;
  double* A = Rg + (size_t)(a.max_params + 1) * K2;
  double* snap = A + (((size_t)(a.max_params + 1) * (a.max_params + 2) / 2 + 1) & ~(size_t)1);
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
    const double F = qsb_cta_eval_rows<NQ>(prog, a.params + a.param_off[inst], a.target, trig, gtab, part, snap, vsm, nullptr);
    qsb_cta_qrows_snap<NQ>(prog, gtab, snap, vsm, Rg);
    if (qsb_syrk_ksplit3(rows, NW) == 2) { for (int e = threadIdx.x; e < rows * (rows + 1) / 2; e += blockDim.x) A[e] = 0.0; __syncthreads(); }
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
  return ((size_t)2 * max_params + QSB_ROWS_PART + 8 * (size_t)(max_ops + max_params) + 2 * N * N) * 8 + 16 * 4 + (size_t)24 * max_ops + 64;
}
