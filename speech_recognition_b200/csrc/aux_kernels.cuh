// Stand-alone stages for callers that compose the reference's per-example maps one by one
// (run/train.py:88-116) instead of the fused forward: SpecAugment draw / apply, delta.
#pragma once

#include "common.cuh"

namespace b200mel {

// spec_augment draws (data.py:282-301), one thread per utterance.
__global__ void specaug_draw_kernel(const int32_t* __restrict__ num_time, int batch, int v, int F, int m_F,
                                    int T, float p, int m_T, uint64_t seed, uint64_t first_utt,
                                    int32_t* __restrict__ table, int32_t* status) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  int32_t local[2 * kMaxMasks];
  int bad = specaug_draw_table(num_time[b], v, F, m_F, T, p, m_T, seed, first_utt + static_cast<uint64_t>(b), local);
  int n = 2 * (m_F + m_T);
  for (int i = 0; i < n; ++i) table[static_cast<size_t>(b) * n + i] = local[i];
  if (bad && status) atomicOr(status, 1);
}

// Same draws from the sample counts (num_time = frames_of(lengths[b])): the pre-pass of the fused
// tensor-core kernel when the caller passes no mask table.
__global__ void specaug_draw_from_lengths_kernel(const int32_t* __restrict__ lengths, int batch, int Lw, int H, int v,
                                                 int F, int m_F, int T, float p, int m_T, uint64_t seed,
                                                 uint64_t first_utt, int32_t* __restrict__ table, int32_t* status) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  int32_t local[2 * kMaxMasks];
  int bad = specaug_draw_table(frames_of(lengths[b], Lw, H), v, F, m_F, T, p, m_T, seed,
                               first_utt + static_cast<uint64_t>(b), local);
  int n = 2 * (m_F + m_T);
  for (int i = 0; i < n; ++i) table[static_cast<size_t>(b) * n + i] = local[i];
  if (bad && status) atomicOr(status, 1);
}

// audio *= mask (data.py:286-288, 299-301) over [B, t_max, v, C].
__global__ void specaug_apply_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t t_max, int v,
                                     int C, const int32_t* __restrict__ table, int m_F, int m_T) {
  __shared__ int32_t s_table[2 * kMaxMasks];
  const int b = blockIdx.y;
  const int n = 2 * (m_F + m_T);
  if (threadIdx.x < n) s_table[threadIdx.x] = table[static_cast<size_t>(b) * n + threadIdx.x];
  __syncthreads();
  const int64_t per = t_max * v * C;
  const int64_t base = static_cast<int64_t>(b) * per;
  for (int64_t o = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; o < per;
       o += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int64_t e = o / C;
    int j = static_cast<int>(e % v);
    int t = static_cast<int>(e / v);
    out[base + o] = in[base + o] * mask_value(s_table, m_F, m_T, t, j);
  }
}

// delta_accelerate (data.py:310-328): [B, t_max, v, 1] -> [B, t_max, v, 3].
__global__ void delta_kernel(const float* __restrict__ in, float* __restrict__ out,
                             const int32_t* __restrict__ n_frames, int64_t t_max, int v) {
  const int b = blockIdx.y;
  const int64_t per = t_max * v;
  const float* x = in + static_cast<int64_t>(b) * per;
  float* y = out + static_cast<int64_t>(b) * per * 3;
  const int64_t nf = n_frames ? n_frames[b] : t_max;
  for (int64_t o = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; o < per * 3;
       o += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int c = static_cast<int>(o % 3);
    int64_t e = o / 3;
    int64_t t = e / v;
    float val = 0.0f;
    if (t < nf) {
      float x0 = x[e];
      float x1 = t >= 1 ? x[e - v] : 0.0f;
      float x2 = t >= 2 ? x[e - 2 * v] : 0.0f;
      float d0 = x0 - x1, d1 = x1 - x2;
      val = c == 0 ? x0 : (c == 1 ? d0 : d0 - d1);
    }
    y[o] = val;
  }
}

}  // namespace b200mel
