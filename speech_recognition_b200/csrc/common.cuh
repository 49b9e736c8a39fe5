// Device-side shared definitions: kernel parameter block, Philox4x32-10, SpecAugment draws.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace b200mel {

constexpr int kMaxMasks = 16;
constexpr int kMaxRadices = 16;

// Everything a kernel needs, passed by value (fits the 4 KB parameter space comfortably).
struct KernelParams {
  // input
  const void* pcm;
  const int64_t* offsets;  // nullable
  const int32_t* lengths;
  int64_t row_stride;
  int32_t pcm_is_f32;
  int32_t batch;
  // output
  float* out;
  int64_t t_max;
  int32_t* n_frames;  // nullable
  int32_t* status;    // nullable
  // framing / FFT
  int32_t Lw, H, N, K, Lw_eff, Nc, even;
  int32_t n_radices;
  int32_t radices[kMaxRadices];
  const float* window;
  const float2* twiddle;
  const float2* post;
  // feature
  int32_t feature_type;
  int32_t M, n_mfcc, F, C;
  float epsilon;
  const int32_t* mel_start;
  const int32_t* mel_count;
  const int32_t* mel_off;
  const float* mel_w;
  const float* dct;
  // SpecAugment
  int32_t n_masks_f, n_masks_t;
  int32_t sa_F, sa_T;
  float sa_p;
  const int32_t* mask_table;  // nullable: draw from Philox
  uint64_t seed, first_utt;
  // tiling (generic kernel)
  int32_t frames_per_tile, halo, warps;
  int32_t smem_per_warp_floats;  // FFT ping-pong + scratch, in floats
  // float32 PCM on the tensor-core path: both kernels are launched, each runs only when the device flag
  // `*gate` (0 = every sample was an exact int16 / 32768) equals its gate_value; null = always run
  const int32_t* gate;
  int32_t gate_value;
  // float32 PCM read by the tensor-core kernel itself (pcm_is_f32 = 1 there): the kernel ORs 1 into this word
  // when a sample it used was not an exact int16 / 32768; the generic kernel behind it is gated on the word
  int32_t* f32_flag;
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011).  Counter-based: no hidden state, so SpecAugment masks are a
// pure function of (seed, global utterance id, draw index) at any GPU count / sharding.
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline uint32_t mulhi32(uint32_t a, uint32_t b) {
  return static_cast<uint32_t>((static_cast<uint64_t>(a) * static_cast<uint64_t>(b)) >> 32);
}

__host__ __device__ inline uint32_t philox4x32_10_first(uint32_t c0, uint32_t c1, uint32_t c2,
                                                        uint32_t c3, uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = mulhi32(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = mulhi32(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return c0;
}

__host__ __device__ inline uint32_t specaug_raw_draw(uint64_t seed, uint64_t utt, uint32_t idx) {
  return philox4x32_10_first(idx, 0u, static_cast<uint32_t>(utt), static_cast<uint32_t>(utt >> 32),
                             static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
}

// One utterance's mask table in the reference's draw order (data.py:282-301):
// (f, f_0) x m_F then (t, t_0) x m_T; entries are (start, width).  tf.random.uniform((), lo, hi,
// int32) = lo + u32 % (hi - lo) and raises when hi <= lo: that case yields an empty mask and
// returns a non-zero status.
__host__ __device__ inline int specaug_draw_table(int32_t num_time, int32_t v, int32_t F, int32_t m_F,
                                                  int32_t T, float p, int32_t m_T, uint64_t seed,
                                                  uint64_t utt, int32_t* table) {
  int bad = 0;
  uint32_t idx = 0;
  int o = 0;
  for (int i = 0; i < m_F; ++i) {
    int32_t f = static_cast<int32_t>(specaug_raw_draw(seed, utt, idx++) % static_cast<uint32_t>(F));
    uint32_t u = specaug_raw_draw(seed, utt, idx++);
    int32_t range = v - f;
    int32_t f0 = 0;
    if (range <= 0) { bad = 1; f = 0; } else { f0 = static_cast<int32_t>(u % static_cast<uint32_t>(range)); }
    table[o++] = f0;
    table[o++] = f;
  }
  if (m_T > 0) {
    int32_t applied = 0;
    int32_t cap = static_cast<int32_t>(static_cast<float>(num_time) * p);
    for (int i = 0; i < m_T; ++i) {
      int32_t t = static_cast<int32_t>(specaug_raw_draw(seed, utt, idx++) % static_cast<uint32_t>(T));
      int32_t room = cap - applied;
      t = t < room ? t : room;
      applied += t;
      uint32_t u = specaug_raw_draw(seed, utt, idx++);
      int32_t range = num_time - t;
      int32_t t0 = 0;
      if (range <= 0) { bad = 1; t = 0; } else { t0 = static_cast<int32_t>(u % static_cast<uint32_t>(range)); }
      table[o++] = t0;
      table[o++] = t;
    }
  }
  return bad;
}

// 0/1 multiplicative mask value at (frame t, channel j): zero inside any [start, start+width).
__device__ __forceinline__ float mask_value(const int32_t* tbl, int n_f, int n_t, int t, int j) {
  float m = 1.0f;
  for (int i = 0; i < n_f; ++i) {
    int s = tbl[2 * i], w = tbl[2 * i + 1];
    if (j >= s && j < s + w) m = 0.0f;
  }
  for (int i = 0; i < n_t; ++i) {
    int s = tbl[2 * (n_f + i)], w = tbl[2 * (n_f + i) + 1];
    if (t >= s && t < s + w) m = 0.0f;
  }
  return m;
}

__host__ __device__ inline int32_t frames_of(int32_t n_samples, int32_t Lw, int32_t H) {
  return n_samples < Lw ? 0 : 1 + (n_samples - Lw) / H;
}

}  // namespace b200mel
