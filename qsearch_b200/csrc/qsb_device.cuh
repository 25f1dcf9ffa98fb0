// Device-side building blocks: complex-fp64 helpers and the closed-form 2x2 leaf gates with their analytic
// parameter derivatives.  Formulas restate qsearch/gates.py (U3 :324-346, XZXZ :282-304, ZXZXZ :219-254,
// U2 :364-376, U1 :393-399) and qsearch/unitaries.py (rot_z :126-130, rot_x :132-136, rot_y :138-142) as
// closed-form entries (no 2x2 matrix products at run time); twins in qsrs/src/circuits.rs:159-759.
#pragma once
#include "../../include/qsb200.h"

#ifdef QSB_EMU
// tests/emu/cuda_emu.h is force-included by the test build (g++ -include ...)
#define QSB_DEVICE static inline
#define QSB_HOST_DEVICE static inline
#define QSB_GLOBAL static void
#else
#include <cuda_runtime.h>
#define QSB_DEVICE __device__ __forceinline__
#define QSB_HOST_DEVICE __host__ __device__ static inline
#define QSB_GLOBAL __global__ void
#define QSB_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#define QSB_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif

#define QSB_TRIG_STRIDE 6     // doubles of cached sin/cos per parameterised op (qubit programs)
#define QSB_TRIG_STRIDE3 16   // the same for radix-3 programs (SingleQutritGate: 3 angles + 5 phases)

struct QsbDevProgram {
  const qsb_op* ops;
  const double* consts;
  const double* pscale;  // [n_params]: factor applied to a parameter before its sin/cos (0.5 for half-angle gates)
  const int32_t* slotmap; // [n_ops + n_params]: gate-table slot -> (op index << 2 | which), -1 for unused slots; the
                          // slot of op j's matrix is ops[j].reserved, its derivative k sits at reserved + 1 + k
  int n_ops;
  int n_params;
};

// ---------------------------------------------------------------------------- complex helpers
typedef double2 cplx;
QSB_DEVICE cplx cmk(double x, double y) { return make_double2(x, y); }
QSB_DEVICE cplx cadd(cplx a, cplx b) { return cmk(a.x + b.x, a.y + b.y); }
QSB_DEVICE cplx cmul(cplx a, cplx b) { return cmk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
QSB_DEVICE cplx cscale(cplx a, double s) { return cmk(a.x * s, a.y * s); }
QSB_DEVICE cplx cconj(cplx a) { return cmk(a.x, -a.y); }
QSB_DEVICE cplx cmuli(cplx a) { return cmk(-a.y, a.x); }  // i * a
// acc + a*b
QSB_DEVICE cplx cfma(cplx a, cplx b, cplx acc) {
  acc.x = fma(a.x, b.x, acc.x);
  acc.x = fma(-a.y, b.y, acc.x);
  acc.y = fma(a.x, b.y, acc.y);
  acc.y = fma(a.y, b.x, acc.y);
  return acc;
}
// acc + conj(a)*b
QSB_DEVICE cplx cfma_conj(cplx a, cplx b, cplx acc) {
  acc.x = fma(a.x, b.x, acc.x);
  acc.x = fma(a.y, b.y, acc.x);
  acc.y = fma(a.x, b.y, acc.y);
  acc.y = fma(-a.y, b.x, acc.y);
  return acc;
}

QSB_DEVICE bool qsb_is_rpp(int opcode) { return opcode == QSB_OP_RXX || opcode == QSB_OP_RYY || opcode == QSB_OP_RZZ; }

QSB_DEVICE int qsb_num_params(int opcode) {
  switch (opcode) {
    case QSB_OP_RX: case QSB_OP_RY: case QSB_OP_RZ: case QSB_OP_U1: case QSB_OP_RXX: case QSB_OP_RYY: case QSB_OP_RZZ: return 1;
    case QSB_OP_U2: case QSB_OP_XZXZ: return 2;
    case QSB_OP_U3: case QSB_OP_ZXZXZ: return 3;
    case QSB_OP_QUTRIT1: return 8;
    default: return 0;
  }
}

// ---------------------------------------------------------------------------- leaf gates
// Cache the sin/cos a parameterised op needs (6 doubles), so every thread can rebuild g / dg cheaply.
QSB_DEVICE void qsb_trig(int opcode, const double* x, double* t) {
  switch (opcode) {
    case QSB_OP_RX: case QSB_OP_RY: case QSB_OP_RZ: case QSB_OP_RXX: case QSB_OP_RYY: case QSB_OP_RZZ:
      sincos(0.5 * x[0], &t[1], &t[0]);  // t0 = cos(x/2), t1 = sin(x/2)
      break;
    case QSB_OP_U1:
      sincos(x[0], &t[1], &t[0]);
      break;
    case QSB_OP_U2:
      sincos(x[0], &t[1], &t[0]);
      sincos(x[1], &t[3], &t[2]);
      break;
    case QSB_OP_U3:
      sincos(0.5 * x[0], &t[1], &t[0]);  // ct, st
      sincos(x[1], &t[3], &t[2]);        // cos phi, sin phi
      sincos(x[2], &t[5], &t[4]);        // cos lambda, sin lambda
      break;
    case QSB_OP_XZXZ:
      sincos(0.5 * x[0], &t[1], &t[0]);
      sincos(0.5 * x[1], &t[3], &t[2]);
      break;
    case QSB_OP_ZXZXZ:
      sincos(0.5 * x[0], &t[1], &t[0]);
      sincos(0.5 * x[1], &t[3], &t[2]);
      sincos(0.5 * x[2], &t[5], &t[4]);
      break;
    case QSB_OP_QUTRIT1:  // (cos, sin) of the three angles, then of the five phases
#pragma unroll
      for (int k = 0; k < 8; ++k) sincos(x[k], &t[2 * k + 1], &t[2 * k]);
      break;
    default: break;
  }
}

// which = 0: the gate g (row-major g[0]=g00, g[1]=g01, g[2]=g10, g[3]=g11); which = 1+k: dg/dx_k.
QSB_DEVICE void qsb_gate2x2(int opcode, const double* t, int which, cplx* g) {
  const cplx Z = cmk(0.0, 0.0);
  switch (opcode) {
    case QSB_OP_RX: {  // [[c, -is], [-is, c]]
      double c = t[0], s = t[1];
      if (which == 0) { g[0] = cmk(c, 0); g[1] = cmk(0, -s); g[2] = cmk(0, -s); g[3] = cmk(c, 0); }
      else { g[0] = cmk(-0.5 * s, 0); g[1] = cmk(0, -0.5 * c); g[2] = cmk(0, -0.5 * c); g[3] = cmk(-0.5 * s, 0); }
    } break;
    case QSB_OP_RY: {  // [[c, -s], [s, c]]
      double c = t[0], s = t[1];
      if (which == 0) { g[0] = cmk(c, 0); g[1] = cmk(-s, 0); g[2] = cmk(s, 0); g[3] = cmk(c, 0); }
      else { g[0] = cmk(-0.5 * s, 0); g[1] = cmk(-0.5 * c, 0); g[2] = cmk(0.5 * c, 0); g[3] = cmk(-0.5 * s, 0); }
    } break;
    case QSB_OP_RZ: {  // diag(e^{-ix/2}, e^{ix/2})
      double c = t[0], s = t[1];
      g[1] = Z; g[2] = Z;
      if (which == 0) { g[0] = cmk(c, -s); g[3] = cmk(c, s); }
      else { g[0] = cmk(-0.5 * s, -0.5 * c); g[3] = cmk(-0.5 * s, 0.5 * c); }
    } break;
    case QSB_OP_U1: {  // diag(1, e^{ix})
      double c = t[0], s = t[1];
      g[1] = Z; g[2] = Z;
      if (which == 0) { g[0] = cmk(1, 0); g[3] = cmk(c, s); }
      else { g[0] = Z; g[3] = cmk(-s, c); }
    } break;
    case QSB_OP_U2: {  // 1/sqrt2 [[1, -e1], [e0, e0 e1]]
      const double k = 0.70710678118654752440;
      cplx e0 = cmk(t[0] * k, t[1] * k), e1 = cmk(t[2], t[3]);
      cplx e01 = cmul(e0, e1);
      cplx e1k = cscale(e1, k);
      if (which == 0) { g[0] = cmk(k, 0); g[1] = cscale(e1k, -1.0); g[2] = e0; g[3] = e01; }
      else if (which == 1) { g[0] = Z; g[1] = Z; g[2] = cmuli(e0); g[3] = cmuli(e01); }
      else { g[0] = Z; g[1] = cscale(cmuli(e1k), -1.0); g[2] = Z; g[3] = cmuli(e01); }
    } break;
    case QSB_OP_U3: {  // [[ct, -st el], [st ep, ct ep el]]
      double ct = t[0], st = t[1];
      cplx ep = cmk(t[2], t[3]), el = cmk(t[4], t[5]);
      cplx epl = cmul(ep, el);
      if (which == 0) { g[0] = cmk(ct, 0); g[1] = cscale(el, -st); g[2] = cscale(ep, st); g[3] = cscale(epl, ct); }
      else if (which == 1) {
        g[0] = cmk(-0.5 * st, 0); g[1] = cscale(el, -0.5 * ct); g[2] = cscale(ep, 0.5 * ct); g[3] = cscale(epl, -0.5 * st);
      } else if (which == 2) { g[0] = Z; g[1] = Z; g[2] = cmuli(cscale(ep, st)); g[3] = cmuli(cscale(epl, ct)); }
      else { g[0] = Z; g[1] = cscale(cmuli(el), -st); g[2] = Z; g[3] = cmuli(cscale(epl, ct)); }
    } break;
    case QSB_OP_XZXZ: {  // Rz(b) [[s0, c0], [c0, -s0]],  s0 = sin(a/2), c0 = cos(a/2)
      double c0 = t[0], s0 = t[1];
      cplx em = cmk(t[2], -t[3]), epz = cmk(t[2], t[3]);  // e^{-ib/2}, e^{+ib/2}
      if (which == 1) { double c = c0; c0 = -0.5 * s0; s0 = 0.5 * c; }  // d/da: s0 -> c0/2, c0 -> -s0/2
      g[0] = cscale(em, s0); g[1] = cscale(em, c0); g[2] = cscale(epz, c0); g[3] = cscale(epz, -s0);
      if (which == 2) {  // d/db: row0 *= -i/2, row1 *= +i/2
        g[0] = cscale(cmuli(g[0]), -0.5); g[1] = cscale(cmuli(g[1]), -0.5);
        g[2] = cscale(cmuli(g[2]), 0.5); g[3] = cscale(cmuli(g[3]), 0.5);
      }
    } break;
    case QSB_OP_ZXZXZ: {  // Rz(c) [[s1, c1], [c1, -s1]] Rz(a)
      double c1 = t[2], s1 = t[3];
      if (which == 2) { double c = c1; c1 = -0.5 * s1; s1 = 0.5 * c; }
      cplx am = cmk(t[0], -t[1]), ap = cmk(t[0], t[1]);  // e^{-ia/2}, e^{+ia/2}
      cplx cm = cmk(t[4], -t[5]), cp = cmk(t[4], t[5]);  // e^{-ic/2}, e^{+ic/2}
      g[0] = cscale(cmul(cm, am), s1); g[1] = cscale(cmul(cm, ap), c1);
      g[2] = cscale(cmul(cp, am), c1); g[3] = cscale(cmul(cp, ap), -s1);
      if (which == 1) {  // d/da: col0 *= -i/2, col1 *= +i/2
        g[0] = cscale(cmuli(g[0]), -0.5); g[2] = cscale(cmuli(g[2]), -0.5);
        g[1] = cscale(cmuli(g[1]), 0.5); g[3] = cscale(cmuli(g[3]), 0.5);
      } else if (which == 3) {  // d/dc: row0 *= -i/2, row1 *= +i/2
        g[0] = cscale(cmuli(g[0]), -0.5); g[1] = cscale(cmuli(g[1]), -0.5);
        g[2] = cscale(cmuli(g[2]), 0.5); g[3] = cscale(cmuli(g[3]), 0.5);
      }
    } break;
    default:
      g[0] = cmk(1, 0); g[1] = Z; g[2] = Z; g[3] = cmk(1, 0);
      break;
  }
}

// 2x2 of a one-qubit op (parameterised from the trig cache, constant from the pool); dagger -> conjugate transpose
QSB_DEVICE void qsb_op2x2(const qsb_op& op, const double* trig, const double* consts, bool dagger, cplx* g) {
  if (op.opcode == QSB_OP_CONST1) {
    const double* c = consts + op.const_off;
    g[0] = cmk(c[0], c[1]); g[1] = cmk(c[2], c[3]); g[2] = cmk(c[4], c[5]); g[3] = cmk(c[6], c[7]);
  } else {
    qsb_gate2x2(op.opcode, trig, 0, g);
  }
  if (dagger) {
    cplx t = g[1];
    g[0] = cconj(g[0]); g[1] = cconj(g[2]); g[2] = cconj(t); g[3] = cconj(g[3]);
  }
}

// ---------------------------------------------------------------------------- the single-qutrit leaf
// SingleQutritGate (qsearch/gates.py:411-523; the parameterisation of Phys. Rev. D 38, 1994 quoted in utils.py:122-150):
// every entry of the 3x3 is a sum of at most two monomials  sign * f1(th1) f2(th2) f3(th3) * exp(i (e . phi)),
// f in {1, sin, cos}, e in {-1, 0, 1}^5.  The eight parameter derivatives follow from the monomials by calculus
// (d sin = cos, d cos = -sin, d/dphi_b = i e_b), so one table serves the gate and its Jacobian (gates.py:441-517 writes
// the nine matrices out entry by entry; the tests compare against fixtures generated from that code).
// t: (cos, sin) of v[0..7] as written by qsb_trig.  which = 0: the gate, 1..3: d/dtheta, 4..8: d/dphi.
#define QSB_Q3_SIN 1
#define QSB_Q3_COS 2
QSB_DEVICE void qsb_q3_term(cplx* g, int which, const double* t, int entry, double sign, int f1, int f2, int f3,
                            int e1, int e2, int e3, int e4, int e5) {
  const int f[3] = {f1, f2, f3}, e[5] = {e1, e2, e3, e4, e5};
  double r = sign;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const double c = t[2 * a], s = t[2 * a + 1];
    if (which == 1 + a) {
      if (f[a] == 0) return;
      r *= (f[a] == QSB_Q3_SIN) ? c : -s;
    } else if (f[a] != 0) {
      r *= (f[a] == QSB_Q3_SIN) ? s : c;
    }
  }
  cplx ph = cmk(r, 0.0);
#pragma unroll
  for (int b = 0; b < 5; ++b)
    if (e[b] != 0) ph = cmul(ph, cmk(t[6 + 2 * b], e[b] * t[7 + 2 * b]));
  if (which >= 4) {
    const int eb = e[which - 4];
    if (eb == 0) return;
    ph = cscale(cmuli(ph), (double)eb);
  }
  g[entry] = cadd(g[entry], ph);
}

// g[9] row-major
QSB_DEVICE void qsb_qutrit3x3(const double* t, int which, cplx* g) {
  const int S = QSB_Q3_SIN, C = QSB_Q3_COS;
#pragma unroll
  for (int i = 0; i < 9; ++i) g[i] = cmk(0.0, 0.0);
  //                       entry sign th1 th2 th3   phi1..phi5
  qsb_q3_term(g, which, t, 0, +1.0, C, C, 0, +1, 0, 0, 0, 0);
  qsb_q3_term(g, which, t, 1, +1.0, S, 0, 0, 0, 0, +1, 0, 0);
  qsb_q3_term(g, which, t, 2, +1.0, C, S, 0, 0, 0, 0, +1, 0);
  qsb_q3_term(g, which, t, 3, +1.0, 0, S, S, 0, 0, 0, -1, -1);
  qsb_q3_term(g, which, t, 3, -1.0, S, C, C, +1, +1, -1, 0, 0);
  qsb_q3_term(g, which, t, 4, +1.0, C, 0, C, 0, +1, 0, 0, 0);
  qsb_q3_term(g, which, t, 5, -1.0, 0, C, S, -1, 0, 0, 0, -1);
  qsb_q3_term(g, which, t, 5, -1.0, S, S, C, 0, +1, -1, +1, 0);
  qsb_q3_term(g, which, t, 6, -1.0, S, C, S, +1, 0, -1, 0, +1);
  qsb_q3_term(g, which, t, 6, -1.0, 0, S, C, 0, -1, 0, -1, 0);
  qsb_q3_term(g, which, t, 7, +1.0, C, 0, S, 0, 0, 0, 0, +1);
  qsb_q3_term(g, which, t, 8, +1.0, 0, C, C, -1, -1, 0, 0, 0);
  qsb_q3_term(g, which, t, 8, -1.0, S, S, S, 0, 0, -1, +1, +1);
}

QSB_DEVICE void qsb_dagger3x3(cplx* g) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    g[4 * i] = cconj(g[4 * i]);
#pragma unroll
    for (int j = i + 1; j < 3; ++j) { const cplx a = g[3 * i + j]; g[3 * i + j] = cconj(g[3 * j + i]); g[3 * j + i] = cconj(a); }
  }
}
