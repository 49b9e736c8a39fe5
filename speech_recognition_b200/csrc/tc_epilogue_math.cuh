// tc_epilogue_math.cuh - arithmetic of the tensor-core kernel's epilogue: the 5-point DFT over the
// residues and the powers |X|^2, on packed f32x2 instructions.  Compiles for the device (inline PTX) and,
// without __CUDACC__, for the host (plain C emulation of the same operations), so that the generated
// epilogue (tc_tables_generated.inc) can be verified against the oracle on a machine without a GPU.
#pragma once

#include <cmath>

namespace b200mel {

// ---------------------------------------------------------------------------------------------
// epilogue arithmetic: 5-point DFT over the residues, returning only the powers |X|^2.
// ---------------------------------------------------------------------------------------------
#define TC_C1 0.30901699437494742410f
#define TC_C2 (-0.80901699437494742410f)
#define TC_S1 0.95105651629515357212f
#define TC_S2 0.58778525229247312917f

// Packed f32x2 arithmetic (sm_100: one instruction, two FP32 lanes): a complex value is one
// (re, im) register pair, straight out of two adjacent TMEM columns.
#if defined(__CUDACC__)
#define TC_MATH_FN __device__ __forceinline__
typedef unsigned long long tc_f2;
TC_MATH_FN tc_f2 tc_pk(float a, float b) {
  tc_f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
TC_MATH_FN void tc_upk(tc_f2 v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
TC_MATH_FN tc_f2 tc_add2(tc_f2 a, tc_f2 b) {
  tc_f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
TC_MATH_FN tc_f2 tc_mul2(tc_f2 a, tc_f2 b) {
  tc_f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
TC_MATH_FN tc_f2 tc_fma2(tc_f2 a, tc_f2 b, tc_f2 c) {
  tc_f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
#else
// Host emulation of the packed instructions (same rounding: one IEEE operation per lane), so that the
// generated epilogue can be compiled with a host compiler and checked against the oracle without a GPU
// (tests/test_tc_epilogue_host.py).
#define TC_MATH_FN static inline
struct tc_f2 { float lo, hi; };
TC_MATH_FN tc_f2 tc_pk(float a, float b) { return tc_f2{a, b}; }
TC_MATH_FN void tc_upk(tc_f2 v, float& a, float& b) { a = v.lo; b = v.hi; }
TC_MATH_FN tc_f2 tc_add2(tc_f2 a, tc_f2 b) { return tc_f2{a.lo + b.lo, a.hi + b.hi}; }
TC_MATH_FN tc_f2 tc_mul2(tc_f2 a, tc_f2 b) { return tc_f2{a.lo * b.lo, a.hi * b.hi}; }
TC_MATH_FN tc_f2 tc_fma2(tc_f2 a, tc_f2 b, tc_f2 c) { return tc_f2{fmaf(a.lo, b.lo, c.lo), fmaf(a.hi, b.hi, c.hi)}; }
#endif

// complex inputs x[r] = (re, im) pairs; outputs p[k2] = |sum_r W5^{r k2} x_r|^2.
// |p + q|^2 and |p - q|^2 with q = -i t share E = |p|^2 + |t|^2 and differ by +-2 (p_r t_i - p_i t_r).
TC_MATH_FN void tc_radix5_power2(const tc_f2* x, float& p0, float& p1, float& p2, float& p3, float& p4) {
  const tc_f2 m1 = tc_pk(-1.0f, -1.0f);
  const tc_f2 c1 = tc_pk(TC_C1, TC_C1), c2 = tc_pk(TC_C2, TC_C2), s1 = tc_pk(TC_S1, TC_S1), s2 = tc_pk(TC_S2, TC_S2);
  const tc_f2 ns1 = tc_pk(-TC_S1, -TC_S1);
  const tc_f2 a1 = tc_add2(x[1], x[4]), b1 = tc_fma2(x[4], m1, x[1]);
  const tc_f2 a2 = tc_add2(x[2], x[3]), b2 = tc_fma2(x[3], m1, x[2]);
  const tc_f2 x0 = tc_add2(x[0], tc_add2(a1, a2));
  const tc_f2 q1 = tc_fma2(c2, a2, tc_fma2(c1, a1, x[0]));
  const tc_f2 q2 = tc_fma2(c1, a2, tc_fma2(c2, a1, x[0]));
  const tc_f2 t1 = tc_fma2(s2, b2, tc_mul2(s1, b1));
  const tc_f2 t2 = tc_fma2(ns1, b2, tc_mul2(s2, b1));
  float ar, ai;
  tc_upk(tc_mul2(x0, x0), ar, ai);
  p0 = ar + ai;
  float e1a, e1b, e2a, e2b, q1r, q1i, q2r, q2i, t1r, t1i, t2r, t2i;
  tc_upk(tc_fma2(t1, t1, tc_mul2(q1, q1)), e1a, e1b);
  tc_upk(tc_fma2(t2, t2, tc_mul2(q2, q2)), e2a, e2b);
  tc_upk(q1, q1r, q1i);
  tc_upk(t1, t1r, t1i);
  tc_upk(q2, q2r, q2i);
  tc_upk(t2, t2r, t2i);
  const float e1 = e1a + e1b, e2 = e2a + e2b;
  const float x1 = fmaf(q1r, t1i, -(q1i * t1r)), x2 = fmaf(q2r, t2i, -(q2i * t2r));
  p1 = fmaf(2.0f, x1, e1);
  p4 = fmaf(-2.0f, x1, e1);
  p2 = fmaf(2.0f, x2, e2);
  p3 = fmaf(-2.0f, x2, e2);
}

// Planar pair: d[r] = (Re a, Re b, Im a, Im b) of residue r for two neighbouring k1 (a, b) - four
// consecutive accumulator columns = two aligned register pairs.  Every instruction of the radix-5
// butterfly and of the powers works on both k1 at once (f32x2), and nothing is unpacked before the
// end: 46 instructions per pair of k1.  Outputs P[2 k2 + lane] = |sum_r W5^{r k2} Y_r[k1_lane]|^2.
// Same formulas as tc_radix5_power2 (t = s b, X_{1,4} = q1 -+ i t1, X_{2,3} = q2 -+ i t2); the
// products with -t_re come from negated constants, so no negation instruction exists.
TC_MATH_FN void tc_radix5_power_pair(const float (&d)[5][4], float* P) {
  tc_f2 xr[5], xi[5];
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    xr[r] = tc_pk(d[r][0], d[r][1]);
    xi[r] = tc_pk(d[r][2], d[r][3]);
  }
  const tc_f2 m1 = tc_pk(-1.0f, -1.0f), two = tc_pk(2.0f, 2.0f), mtwo = tc_pk(-2.0f, -2.0f);
  const tc_f2 c1 = tc_pk(TC_C1, TC_C1), c2 = tc_pk(TC_C2, TC_C2), s1 = tc_pk(TC_S1, TC_S1), s2 = tc_pk(TC_S2, TC_S2);
  const tc_f2 ns1 = tc_pk(-TC_S1, -TC_S1), ns2 = tc_pk(-TC_S2, -TC_S2);
  const tc_f2 a1r = tc_add2(xr[1], xr[4]), a1i = tc_add2(xi[1], xi[4]);
  const tc_f2 b1r = tc_fma2(xr[4], m1, xr[1]), b1i = tc_fma2(xi[4], m1, xi[1]);
  const tc_f2 a2r = tc_add2(xr[2], xr[3]), a2i = tc_add2(xi[2], xi[3]);
  const tc_f2 b2r = tc_fma2(xr[3], m1, xr[2]), b2i = tc_fma2(xi[3], m1, xi[2]);
  const tc_f2 x0r = tc_add2(xr[0], tc_add2(a1r, a2r)), x0i = tc_add2(xi[0], tc_add2(a1i, a2i));
  const tc_f2 q1r = tc_fma2(c2, a2r, tc_fma2(c1, a1r, xr[0])), q1i = tc_fma2(c2, a2i, tc_fma2(c1, a1i, xi[0]));
  const tc_f2 q2r = tc_fma2(c1, a2r, tc_fma2(c2, a1r, xr[0])), q2i = tc_fma2(c1, a2i, tc_fma2(c2, a1i, xi[0]));
  const tc_f2 nt1r = tc_fma2(ns2, b2r, tc_mul2(ns1, b1r)), t1i = tc_fma2(s2, b2i, tc_mul2(s1, b1i));   // -t1.re, t1.im
  const tc_f2 nt2r = tc_fma2(s1, b2r, tc_mul2(ns2, b1r)), t2i = tc_fma2(ns1, b2i, tc_mul2(s2, b1i));   // -t2.re, t2.im
  const tc_f2 p0 = tc_fma2(x0i, x0i, tc_mul2(x0r, x0r));
  const tc_f2 e1 = tc_fma2(t1i, t1i, tc_fma2(nt1r, nt1r, tc_fma2(q1i, q1i, tc_mul2(q1r, q1r))));
  const tc_f2 e2 = tc_fma2(t2i, t2i, tc_fma2(nt2r, nt2r, tc_fma2(q2i, q2i, tc_mul2(q2r, q2r))));
  const tc_f2 x1 = tc_fma2(q1r, t1i, tc_mul2(q1i, nt1r));   // q1.re t1.im - q1.im t1.re
  const tc_f2 x2 = tc_fma2(q2r, t2i, tc_mul2(q2i, nt2r));
  tc_upk(p0, P[0], P[1]);
  tc_upk(tc_fma2(two, x1, e1), P[2], P[3]);
  tc_upk(tc_fma2(two, x2, e2), P[4], P[5]);
  tc_upk(tc_fma2(mtwo, x2, e2), P[6], P[7]);
  tc_upk(tc_fma2(mtwo, x1, e1), P[8], P[9]);
}

// complex inputs (xr[r], xi[r]); outputs p[k2] = |sum_r W5^{r k2} x_r|^2
TC_MATH_FN void tc_radix5_power(const float* xr, const float* xi, float& p0, float& p1, float& p2,
                                                float& p3, float& p4) {
  const float a1r = xr[1] + xr[4], a1i = xi[1] + xi[4], b1r = xr[1] - xr[4], b1i = xi[1] - xi[4];
  const float a2r = xr[2] + xr[3], a2i = xi[2] + xi[3], b2r = xr[2] - xr[3], b2i = xi[2] - xi[3];
  const float x0r = xr[0] + a1r + a2r, x0i = xi[0] + a1i + a2i;
  const float p1r = fmaf(TC_C2, a2r, fmaf(TC_C1, a1r, xr[0])), p1i = fmaf(TC_C2, a2i, fmaf(TC_C1, a1i, xi[0]));
  const float p2r = fmaf(TC_C1, a2r, fmaf(TC_C2, a1r, xr[0])), p2i = fmaf(TC_C1, a2i, fmaf(TC_C2, a1i, xi[0]));
  // q = -i * (s b): (re, im) -> (im, -re)
  const float t1r = fmaf(TC_S2, b2r, TC_S1 * b1r), t1i = fmaf(TC_S2, b2i, TC_S1 * b1i);
  const float t2r = fmaf(-TC_S1, b2r, TC_S2 * b1r), t2i = fmaf(-TC_S1, b2i, TC_S2 * b1i);
  const float q1r = t1i, q1i = -t1r, q2r = t2i, q2i = -t2r;
  const float y1r = p1r + q1r, y1i = p1i + q1i, y4r = p1r - q1r, y4i = p1i - q1i;
  const float y2r = p2r + q2r, y2i = p2i + q2i, y3r = p2r - q2r, y3i = p2i - q2i;
  p0 = fmaf(x0r, x0r, x0i * x0i);
  p1 = fmaf(y1r, y1r, y1i * y1i);
  p2 = fmaf(y2r, y2r, y2i * y2i);
  p3 = fmaf(y3r, y3r, y3i * y3i);
  p4 = fmaf(y4r, y4r, y4i * y4i);
}

// real inputs (k1 = 0): outputs k2 = 0, 1, 2 only (3, 4 are their conjugates)
TC_MATH_FN void tc_radix5_power_real(const float* x, float& p0, float& p1, float& p2) {
  const float a1 = x[1] + x[4], b1 = x[1] - x[4], a2 = x[2] + x[3], b2 = x[2] - x[3];
  const float x0 = x[0] + a1 + a2;
  const float r1 = fmaf(TC_C2, a2, fmaf(TC_C1, a1, x[0])), i1 = fmaf(TC_S2, b2, TC_S1 * b1);
  const float r2 = fmaf(TC_C1, a2, fmaf(TC_C2, a1, x[0])), i2 = fmaf(-TC_S1, b2, TC_S2 * b1);
  p0 = x0 * x0;
  p1 = fmaf(r1, r1, i1 * i1);
  p2 = fmaf(r2, r2, i2 * i2);
}

// Quad 0 of a residue slot: d[r] = (Re Y_r[1], Im Y_r[1], Y_r[0], R_r) with Y_r[32] = R_r e^{-i pi r / 5}.
// P[0..4] = k1 = 1 (k2 = 0..4), P[5..7] = k1 = 0 (k2 = 0..2), P[8..10] = k1 = 32 (k2 = 0..2; the other two
// outputs of both real-input transforms are mirror images).
TC_MATH_FN void tc_radix5_power_special(const float (&d)[5][4], float* P) {
  tc_f2 xc[5];
  float x0[5], xr[5], xi[5];
  const float cr[5] = {1.0f, 0.80901699437494742410f, 0.30901699437494742410f, -0.30901699437494742410f,
                       -0.80901699437494742410f};
  const float sr[5] = {0.0f, 0.58778525229247312917f, 0.95105651629515357212f, 0.95105651629515357212f,
                       0.58778525229247312917f};
#pragma unroll
  for (int r = 0; r < 5; ++r) {
    xc[r] = tc_pk(d[r][0], d[r][1]);
    x0[r] = d[r][2];
    xr[r] = d[r][3] * cr[r];
    xi[r] = -d[r][3] * sr[r];
  }
  tc_radix5_power2(xc, P[0], P[1], P[2], P[3], P[4]);
  tc_radix5_power_real(x0, P[5], P[6], P[7]);
  float p3, p4;
  tc_radix5_power(xr, xi, P[8], P[9], P[10], p3, p4);
}

}  // namespace b200mel
