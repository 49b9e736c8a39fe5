// tc_kernel_device.cuh - device code of the tensor-core front-end kernel (see tc_kernel.cuh).
#pragma once

namespace b200mel {

struct TcParams {
  const uint8_t* b_image;  // 10 x 8 KB
  float log_eps;
  int32_t tiles_per_utt;
  int32_t total_tiles;
  int32_t tout;            // output frames per tile = 128 - halo
  int32_t halo;            // 2 with delta features, else 0
};

// ---------------------------------------------------------------------------------------------
// producer: 40 consecutive int16 samples -> 8 consecutive elements of each of the 5 decimated
// streams, as exact f16 (hi, lo) pairs.
// ---------------------------------------------------------------------------------------------
__device__ __noinline__ uint4 tc_load8_slow(const short* __restrict__ base, int64_t s, int32_t len) {
  // unaligned utterance start, or a chunk that crosses [0, len): element-wise, zero outside
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t s0 = s + 2 * i, s1 = s0 + 1;
    uint32_t lo = (s0 >= 0 && s0 < len) ? static_cast<uint16_t>(__ldg(base + s0)) : 0u;
    uint32_t hi = (s1 >= 0 && s1 < len) ? static_cast<uint16_t>(__ldg(base + s1)) : 0u;
    w[i] = lo | (hi << 16);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}
// x = 256 h + l (h = signed high byte, l = unsigned low byte).  With the f16 "magic" encodings
// 0x64uu = 1024 + uu:  hq = (1024 + (h + 128)) - 1156 = h - 4,  lq = 1024 + l, so that
//   hi = fma(hq, 256, lq) = rn_f16(x)         (one rounding)
//   lo = fma(hq, 256, -hi) + lq = x - hi      (both steps exact: small integers)
// `d` holds two samples already XOR-ed with 0x8000.
__device__ __forceinline__ void tc_split_pair(uint32_t d, uint32_t& hi_out, uint32_t& lo_out) {
  const uint32_t hp = tc::prmt(d, 0x64646464u, 0x4341u);
  const uint32_t lp = tc::prmt(d, 0x64646464u, 0x4240u);
  const __half2 c1156 = __half2half2(__ushort_as_half(0x6484));  // 1156.0
  const __half2 c256 = __half2half2(__ushort_as_half(0x5C00));   // 256.0
  const __half2 hq = __hsub2(*reinterpret_cast<const __half2*>(&hp), c1156);
  const __half2 lq = *reinterpret_cast<const __half2*>(&lp);
  const __half2 hi = __hfma2(hq, c256, lq);
  const __half2 e = __hfma2(hq, c256, __hneg2(hi));
  const __half2 lo = __hadd2(e, lq);
  hi_out = *reinterpret_cast<const uint32_t*>(&hi);
  lo_out = *reinterpret_cast<const uint32_t*>(&lo);
}

// Swizzled byte offset of element block g (8 elements = 16 bytes) inside a stream: SWIZZLE_64B.
__device__ __forceinline__ uint32_t tc_stream_offset(int g) {
  const uint32_t o = 16u * static_cast<uint32_t>(g);
  return o ^ (((o >> 7) & 3u) << 4);
}

// One (group, residue): w[20] = the group's 40 samples, already XOR-ed with 0x8000, two per word.
// Stream R takes samples 5 e + R (e = 0..7): two 16-byte stores, (hi, lo) terms.
template <int R>
__device__ __forceinline__ void tc_produce_res(const uint32_t (&w)[20], uint8_t* sA, uint32_t ph) {
  uint32_t hi[4], lo[4];
#pragma unroll
  for (int pi = 0; pi < 4; ++pi) {
    const int a = 10 * pi + R, b = a + 5;  // sample indices of stream elements 2 pi, 2 pi + 1
    const uint32_t sel = (2 * (a & 1)) | ((2 * (a & 1) + 1) << 4) | ((4 + 2 * (b & 1)) << 8) | ((5 + 2 * (b & 1)) << 12);
    const uint32_t d = tc::prmt(w[a >> 1], w[b >> 1], sel);
    tc_split_pair(d, hi[pi], lo[pi]);
  }
#if defined(TC_ABLATE_PSTS)
  if (ph != 0xFFFFFFFFu) return;  // (experiment: conversion work without the shared-memory stores)
#endif
  *reinterpret_cast<uint4*>(sA + (0 * 5 + R) * kTcStreamBytes + ph) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
  *reinterpret_cast<uint4*>(sA + (1 * 5 + R) * kTcStreamBytes + ph) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
}

__device__ __forceinline__ void tc_xor_words(const uint4 (&raw)[5], uint32_t (&w)[20]) {
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    w[4 * q + 0] = raw[q].x ^ 0x80008000u;
    w[4 * q + 1] = raw[q].y ^ 0x80008000u;
    w[4 * q + 2] = raw[q].z ^ 0x80008000u;
    w[4 * q + 3] = raw[q].w ^ 0x80008000u;
  }
}

// One group, all residues (slow path only).  Out of line.
__device__ __noinline__ void tc_produce_group(uint4 r0, uint4 r1, uint4 r2, uint4 r3, uint4 r4, uint8_t* sA, int g) {
  const uint4 raw[5] = {r0, r1, r2, r3, r4};
  uint32_t w[20];
  tc_xor_words(raw, w);
  const uint32_t ph = tc_stream_offset(g);
  tc_produce_res<0>(w, sA, ph);
  tc_produce_res<1>(w, sA, ph);
  tc_produce_res<2>(w, sA, ph);
  tc_produce_res<3>(w, sA, ph);
  tc_produce_res<4>(w, sA, ph);
}

// ---------------------------------------------------------------------------------------------
// epilogue arithmetic: 5-point DFT over the residues, returning only the powers |X|^2.
// ---------------------------------------------------------------------------------------------
#define TC_C1 0.30901699437494742410f
#define TC_C2 (-0.80901699437494742410f)
#define TC_S1 0.95105651629515357212f
#define TC_S2 0.58778525229247312917f

// Packed f32x2 arithmetic (sm_100: one instruction, two FP32 lanes): a complex value is one
// (re, im) register pair, straight out of two adjacent TMEM columns.
typedef unsigned long long tc_f2;
__device__ __forceinline__ tc_f2 tc_pk(float a, float b) {
  tc_f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void tc_upk(tc_f2 v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ tc_f2 tc_add2(tc_f2 a, tc_f2 b) {
  tc_f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ tc_f2 tc_mul2(tc_f2 a, tc_f2 b) {
  tc_f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ tc_f2 tc_fma2(tc_f2 a, tc_f2 b, tc_f2 c) {
  tc_f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

// complex inputs x[r] = (re, im) pairs; outputs p[k2] = |sum_r W5^{r k2} x_r|^2.
// |p + q|^2 and |p - q|^2 with q = -i t share E = |p|^2 + |t|^2 and differ by +-2 (p_r t_i - p_i t_r).
__device__ __forceinline__ void tc_radix5_power2(const tc_f2* x, float& p0, float& p1, float& p2, float& p3, float& p4) {
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

// complex inputs (xr[r], xi[r]); outputs p[k2] = |sum_r W5^{r k2} x_r|^2
__device__ __forceinline__ void tc_radix5_power(const float* xr, const float* xi, float& p0, float& p1, float& p2,
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
__device__ __forceinline__ void tc_radix5_power_real(const float* x, float& p0, float& p1, float& p2) {
  const float a1 = x[1] + x[4], b1 = x[1] - x[4], a2 = x[2] + x[3], b2 = x[2] - x[3];
  const float x0 = x[0] + a1 + a2;
  const float r1 = fmaf(TC_C2, a2, fmaf(TC_C1, a1, x[0])), i1 = fmaf(TC_S2, b2, TC_S1 * b1);
  const float r2 = fmaf(TC_C1, a2, fmaf(TC_C2, a1, x[0])), i2 = fmaf(-TC_S1, b2, TC_S2 * b1);
  p0 = x0 * x0;
  p1 = fmaf(r1, r1, i1 * i1);
  p2 = fmaf(r2, r2, i2 * i2);
}

}  // namespace b200mel

#include "tc_kernel_main.cuh"
