// Kernels either side of the front-end: 16-bit PCM / WAV payload
// bytes -> samples (load_audio_file, data.py:94-117) and the models' conv-reduced padding mask
// (models/las.py:205-217, models/deepspeech2.py:68-78).  Streaming, HBM-bound byte / compare work.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200mel {

// 32-bit little-endian word k of the byte range [base + start, base + start + n_bytes), counted from the
// 4-byte-aligned address at or below its first byte.  Bytes outside the range read as zero: that is the
// "\x00" the reference appends to an odd-length .pcm file (data.py:102-103), and it keeps every access
// inside the caller's byte range.
__device__ __forceinline__ uint32_t ingest_word(const uint8_t* __restrict__ base, int64_t lo, int64_t hi, int64_t w) {
  // lo / hi: first / one-past-last valid byte index; w: byte index of the aligned word (multiple of 4
  // relative to the address, i.e. (base + w) % 4 == 0)
  if (w >= lo && w + 4 <= hi) return __ldg(reinterpret_cast<const uint32_t*>(base + w));
  uint32_t v = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t b = w + i;
    if (b >= lo && b < hi) v |= static_cast<uint32_t>(__ldg(base + b)) << (8 * i);
  }
  return v;
}

// Output words i = 0..3 of a sample group from the eight words of its two aligned 16-byte blocks: the payload
// starts Q words + sh bits into the first block.
template <int Q>
__device__ __forceinline__ void ingest_pick(const uint32_t (&x)[8], int sh, uint32_t (&w)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) w[i] = __funnelshift_r(x[Q + i], x[Q + i + 1], sh);
}

// Mono: int16 out (the tensor-core kernel's native input) or float32 = int16 / 32768 (data.py:99,105).
// grid = (blocks, B); a thread produces 8 consecutive samples (one 128-bit store for int16) from five
// aligned 32-bit words funnel-shifted by the payload's byte misalignment (a canonical WAV header is 44
// bytes, so payloads are rarely 16-byte aligned).
template <typename OutT>
__global__ void __launch_bounds__(256)
ingest_mono_kernel(const uint8_t* __restrict__ bytes, const int64_t* __restrict__ byte_offsets,
                   const int64_t* __restrict__ byte_lengths, OutT* __restrict__ out,
                   const int64_t* __restrict__ out_offsets, int64_t row_stride, int32_t* __restrict__ lengths_out) {
  const int b = blockIdx.y;
  const int64_t lo = byte_offsets[b];
  const int64_t nbytes = byte_lengths[b] < 0 ? 0 : byte_lengths[b];
  const int64_t hi = lo + nbytes;
  int64_t n = (nbytes + 1) >> 1;  // odd length: one more sample whose high byte is the appended zero
  OutT* dst;
  int64_t cap;  // elements this utterance may write
  if (out_offsets) {
    dst = out + out_offsets[b];
    cap = n;
  } else {
    dst = out + static_cast<int64_t>(b) * row_stride;
    if (n > row_stride) n = row_stride;
    cap = row_stride;  // dense rows are zero-filled behind the samples
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) lengths_out[b] = static_cast<int32_t>(n);
  const uintptr_t addr = reinterpret_cast<uintptr_t>(bytes) + static_cast<uintptr_t>(lo);
  const int mis = static_cast<int>(addr & 3);
  const int64_t w0 = lo - mis;  // aligned byte index of the first word
  const int sh = mis * 8;
  const int mis16 = static_cast<int>(addr & 15);
  const int64_t base16 = lo - mis16;  // 16-byte aligned block holding the first payload byte
  const int q16 = mis16 >> 2;
  const bool vec_ok = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
  for (int64_t g = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; g * 8 < cap;
       g += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t s0 = g * 8;
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    const int64_t blk = base16 + s0 * 2;  // byte index of the aligned 16-byte block the group starts in
    if (s0 < n && blk >= lo && blk + 32 <= hi) {
      // interior group: two aligned 128-bit loads (the second is the next thread's first: an L1 hit) instead
      // of five 32-bit ones
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(bytes + blk));
      const uint4 c = __ldg(reinterpret_cast<const uint4*>(bytes + blk + 16));
      const uint32_t x[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
      switch (q16) {
        case 0: ingest_pick<0>(x, sh, w); break;
        case 1: ingest_pick<1>(x, sh, w); break;
        case 2: ingest_pick<2>(x, sh, w); break;
        default: ingest_pick<3>(x, sh, w); break;
      }
    } else if (s0 < n) {
      uint32_t r[5];
#pragma unroll
      for (int i = 0; i < 5; ++i) r[i] = ingest_word(bytes, lo, hi, w0 + s0 * 2 + 4 * i);
#pragma unroll
      for (int i = 0; i < 4; ++i) w[i] = __funnelshift_r(r[i], r[i + 1], sh);
    }
    // samples at or behind n (dense truncation at row_stride, or the zero tail) are zero
    if (sizeof(OutT) == 2) {
      uint32_t o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint32_t v = w[i];
        if (s0 + 2 * i + 1 >= n) v &= 0xFFFFu;
        if (s0 + 2 * i >= n) v = 0u;
        o[i] = v;
      }
      if (vec_ok && s0 + 8 <= cap) {
        *reinterpret_cast<uint4*>(reinterpret_cast<int16_t*>(dst) + s0) = make_uint4(o[0], o[1], o[2], o[3]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (s0 + i < cap)
            reinterpret_cast<int16_t*>(dst)[s0 + i] = static_cast<int16_t>((o[i >> 1] >> (16 * (i & 1))) & 0xFFFFu);
      }
    } else {
      float f[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        int16_t s = static_cast<int16_t>((w[i >> 1] >> (16 * (i & 1))) & 0xFFFFu);
        f[i] = (s0 + i < n) ? static_cast<float>(s) * (1.0f / 32768.0f) : 0.0f;
      }
      if (vec_ok && s0 + 8 <= cap) {
        float4* q = reinterpret_cast<float4*>(reinterpret_cast<float*>(dst) + s0);
        q[0] = make_float4(f[0], f[1], f[2], f[3]);
        q[1] = make_float4(f[4], f[5], f[6], f[7]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (s0 + i < cap) reinterpret_cast<float*>(dst)[s0 + i] = f[i];
      }
    }
  }
}

// Multi-channel: float32 out = reduce_mean over channels of int16 / 32768 (data.py:99, 116): the sum of
// C exact terms is exact in float32 for C <= 256, then one rounded division, as the reference's mean.
__global__ void __launch_bounds__(256)
ingest_multi_kernel(const uint8_t* __restrict__ bytes, const int64_t* __restrict__ byte_offsets,
                    const int64_t* __restrict__ byte_lengths, const int32_t* __restrict__ channels,
                    float* __restrict__ out, const int64_t* __restrict__ out_offsets, int64_t row_stride,
                    int32_t* __restrict__ lengths_out) {
  const int b = blockIdx.y;
  const int64_t lo = byte_offsets[b];
  const int64_t nbytes = byte_lengths[b] < 0 ? 0 : byte_lengths[b];
  const int64_t hi = lo + nbytes;
  const int C = channels[b] < 1 ? 1 : channels[b];
  int64_t n = ((nbytes + 1) >> 1) / C;  // whole sample frames only
  float* dst;
  int64_t cap;
  if (out_offsets) {
    dst = out + out_offsets[b];
    cap = n;
  } else {
    dst = out + static_cast<int64_t>(b) * row_stride;
    if (n > row_stride) n = row_stride;
    cap = row_stride;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) lengths_out[b] = static_cast<int32_t>(n);
  const float fc = static_cast<float>(C);
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < cap;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float acc = 0.0f;
    if (i < n) {
      int64_t p = lo + i * C * 2;
      for (int c = 0; c < C; ++c, p += 2) {
        uint32_t l = __ldg(bytes + p);
        uint32_t h = (p + 1 < hi) ? __ldg(bytes + p + 1) : 0u;
        acc += static_cast<float>(static_cast<int16_t>(l | (h << 8))) * (1.0f / 32768.0f);
      }
      acc = __fdiv_rn(acc, fc);
    }
    dst[i] = acc;
  }
}

// _audio_mask (models/las.py:205-217, models/deepspeech2.py:68-78): frame t is "real" when any of its
// F*C values differs from the pad value 0 (so -0.0 and +0.0 both count as padding, NaN as real);
// group consecutive frames by the product of the convolution time strides and OR them.
// A warp takes a PAIR of consecutive entries of one utterance - 2 * group frames, contiguous in memory - as one
// stream of 128-bit loads (five independent loads per lane at F*C = 80, group = 4), over a flat grid-stride
// loop of all pairs (one resident wave).
__global__ void __launch_bounds__(256)
audio_mask_kernel(const float* __restrict__ feat, int batch, int64_t t_max, int64_t row_elems, int group, int64_t reduced,
                  uint8_t* __restrict__ mask) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const int64_t n = static_cast<int64_t>(group) * row_elems;  // floats per entry
  const int64_t pairs = (reduced + 1) >> 1;
  const int64_t total = static_cast<int64_t>(batch) * pairs;
  for (int64_t e = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; e < total; e += warps) {
    const int64_t b = e / pairs, r = 2 * (e - b * pairs);
    const bool two = r + 1 < reduced;
    const float* p = feat + (b * t_max + r * group) * row_elems;
    const int64_t len = two ? 2 * n : n;
    bool any0 = false, any1 = false;
    if (((reinterpret_cast<uintptr_t>(p) & 15) == 0) && (n & 3) == 0) {
      const float4* q = reinterpret_cast<const float4*>(p);
      const int64_t n4 = n >> 2, len4 = len >> 2;
      int64_t i = lane;
      for (; i + 64 < len4; i += 96) {
        const float4 u = __ldg(q + i), v = __ldg(q + i + 32), w = __ldg(q + i + 64);
        const bool zu = (u.x != 0.0f) | (u.y != 0.0f) | (u.z != 0.0f) | (u.w != 0.0f);
        const bool zv = (v.x != 0.0f) | (v.y != 0.0f) | (v.z != 0.0f) | (v.w != 0.0f);
        const bool zw = (w.x != 0.0f) | (w.y != 0.0f) | (w.z != 0.0f) | (w.w != 0.0f);
        any0 |= (zu & (i < n4)) | (zv & (i + 32 < n4)) | (zw & (i + 64 < n4));
        any1 |= (zu & (i >= n4)) | (zv & (i + 32 >= n4)) | (zw & (i + 64 >= n4));
      }
      for (; i < len4; i += 32) {
        const float4 u = __ldg(q + i);
        const bool zu = (u.x != 0.0f) | (u.y != 0.0f) | (u.z != 0.0f) | (u.w != 0.0f);
        any0 |= zu & (i < n4);
        any1 |= zu & (i >= n4);
      }
    } else {
      for (int64_t i = lane; i < len; i += 32) {
        const bool z = __ldg(p + i) != 0.0f;
        any0 |= z & (i < n);
        any1 |= z & (i >= n);
      }
    }
    const unsigned m0 = __ballot_sync(0xFFFFFFFFu, any0), m1 = __ballot_sync(0xFFFFFFFFu, any1);
    if (lane == 0) mask[b * reduced + r] = m0 ? 1 : 0;
    if (lane == 1 && two) mask[b * reduced + r + 1] = m1 ? 1 : 0;
  }
}

}  // namespace b200mel
