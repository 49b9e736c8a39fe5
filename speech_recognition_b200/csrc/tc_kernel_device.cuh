// tc_kernel_device.cuh - device code of the tensor-core front-end kernel (see tc_kernel.cuh).
#pragma once

#include "tc_epilogue_math.cuh"

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

// float32 entry (data.py:98-99, 104-105, 116: the reference's audio is int16 / 32768 in float32): eight samples
// (two 128-bit loads) -> the same four words of packed 16-bit samples the int16 entry loads, plus the
// exactness test.  m = fma(x, 32768, 1.5 * 2^23) is an integer-valued float whose low 16 bits ARE the
// two's-complement sample when -32768 <= x * 32768 <= 32767 - its bit pattern then lies in
// [0x4B3F8000, 0x4B407FFF] (NaN and infinities do not) - and fma(x, 32768, 1.5 * 2^23 - m) is exactly +0.0 iff
// x * 32768 was that integer.  Nothing is tested per sample: the residuals' bits and the patterns' offsets
// are OR-ed into two words the thread looks at once, at the end of the kernel.  Three packed f32x2
// instructions, two integer adds, two LOP3 and one PRMT per pair of samples.
struct TcF32Check {
  uint32_t resid = 0;   // OR of the residuals' bit patterns: 0 = every product was an integer
  uint32_t range = 0;   // OR of (pattern - 0x4B3F8000): < 0x10000 = every sample was inside the int16 range
  __device__ __forceinline__ bool bad() const { return resid != 0 || range >= 0x10000u; }
};
__device__ __forceinline__ uint32_t tc_f32_pair(float x0, float x1, TcF32Check& chk) {
  const tc_f2 k = tc_pk(32768.0f, 32768.0f), magic = tc_pk(12582912.0f, 12582912.0f), m1 = tc_pk(-1.0f, -1.0f);
  const tc_f2 x = tc_pk(x0, x1);
  const tc_f2 m = tc_fma2(x, k, magic);
  const tc_f2 nback = tc_fma2(m, m1, magic);      // -(m - magic), exact
  const tc_f2 resid = tc_fma2(x, k, nback);       // x * 32768 - (m - magic), one rounding of the exact value
  float ma, mb, ra, rb;
  tc_upk(m, ma, mb);
  tc_upk(resid, ra, rb);
  const uint32_t u0 = __float_as_uint(ma), u1 = __float_as_uint(mb);
  chk.resid |= __float_as_uint(ra) | __float_as_uint(rb);
  chk.range |= (u0 - 0x4B3F8000u) | (u1 - 0x4B3F8000u);
  return tc::prmt(u0, u1, 0x5410u);
}
__device__ __forceinline__ void tc_f32_block(const uint4 a, const uint4 b, uint32_t* w4, TcF32Check& chk) {
  w4[0] = tc_f32_pair(__uint_as_float(a.x), __uint_as_float(a.y), chk);
  w4[1] = tc_f32_pair(__uint_as_float(a.z), __uint_as_float(a.w), chk);
  w4[2] = tc_f32_pair(__uint_as_float(b.x), __uint_as_float(b.y), chk);
  w4[3] = tc_f32_pair(__uint_as_float(b.z), __uint_as_float(b.w), chk);
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

}  // namespace b200mel

#include "tc_epilogue_math.cuh"
#include "tc_kernel_main.cuh"
