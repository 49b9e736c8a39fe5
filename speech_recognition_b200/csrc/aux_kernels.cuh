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

// ---------------------------------------------------------------------------------------------
// Second stage of the tensor-core MFCC path (make_mfcc, data.py:192-241, for the shipped kspon /
// clovacall parameters): log-mel [B, t_max, M] (written by tc_frontend_kernel<1>) ->
// DCT-II * rsqrt(2M), first n_mfcc coefficients (data.py:235) -> SpecAugment on the coefficients
// (run/train.py:100-110: v = num_mfcc) -> delta / delta-delta (data.py:310-328) -> [B, t_max, n_mfcc, C],
// padding frames +0.0.  One block = 64 frames of one utterance (+ 2 frames of halo for the deltas).
// Register blocking: a thread owns 4 coefficients (c, c + Q, c + 2Q, c + 3Q; Q = ceil(n_mfcc / 4)) of 4
// frames: one 128-bit read of a log-mel row feeds 16 FMAs (a 128-bit shared-memory read by a warp
// costs four wavefronts even when it is a broadcast, so 1 coefficient per thread was LSU-bound at
// 1.1 wavefronts per FMA instruction; this is 0.5 with all six warps of the block at work).
// ---------------------------------------------------------------------------------------------
constexpr int kMfccFrames = 64, kMfccGroups = 9, kMfccMaxM = 80, kMfccMaxC = 80, kMfccThreads = 192;
inline size_t mfcc_post_smem_bytes(int M, int n_mfcc) {
  return sizeof(float) * (static_cast<size_t>(8 * kMfccGroups) * M + static_cast<size_t>(M) * n_mfcc + (n_mfcc > 40 ? 16 * 192 : 0));
}
__global__ void __launch_bounds__(kMfccThreads)
    mfcc_post_kernel(const float* __restrict__ lm, float* __restrict__ out, const int32_t* __restrict__ lengths,
                     int64_t t_max, int Lw, int H, int M, int n_mfcc, int C, const float* __restrict__ dct,
                     int n_masks_f, int n_masks_t, int sa_F, int sa_T, float sa_p,
                     const int32_t* __restrict__ mask_table, uint64_t seed, uint64_t first_utt, int32_t* status,
                     const int32_t* __restrict__ gate, int gate_value) {
  extern __shared__ __align__(16) float mfcc_smem[];
  if (gate && (*gate != 0) != (gate_value != 0)) return;  // (float32 entry: the generic kernel takes the batch)
  const int halo = C == 3 ? 2 : 0, rows = kMfccFrames + halo;
  float* sx = mfcc_smem;                       // [72][M]   log-mel rows (rows >= `rows` are zero)
  float* sd = sx + 8 * kMfccGroups * M;        // [M][n_mfcc]
  float* sm = sx;                              // [72][n_mfcc] coefficients: reuses sx after the DCT
  __shared__ int32_t s_table[2 * kMaxMasks];
  const int b = blockIdx.y;
  const int64_t t0 = static_cast<int64_t>(blockIdx.x) * kMfccFrames;
  if (t0 >= t_max) return;
  const int32_t n_fr_all = frames_of(lengths[b], Lw, H);
  const int64_t n_fr = n_fr_all < t_max ? n_fr_all : t_max;
  const int n_masks = n_masks_f + n_masks_t;
  if (threadIdx.x == 0 && n_masks > 0) {
    if (mask_table) {
      for (int i = 0; i < 2 * n_masks; ++i) s_table[i] = mask_table[static_cast<size_t>(b) * 2 * n_masks + i];
    } else {
      const int bad = specaug_draw_table(n_fr_all, n_mfcc, sa_F, n_masks_f, sa_T, sa_p, n_masks_t, seed,
                                         first_utt + static_cast<uint64_t>(b), s_table);
      if (bad && blockIdx.x == 0 && status) atomicOr(status, 1);
    }
  }
  for (int i = threadIdx.x; i < M * n_mfcc; i += blockDim.x) sd[i] = dct[i];
  const float* lmb = lm + static_cast<int64_t>(b) * t_max * M;
  {  // 72 rows x M/4 float4, (row, quad) advanced without divisions
    const int m4 = M >> 2, step_r = blockDim.x / m4, step_q = blockDim.x - step_r * m4;
    int r = threadIdx.x / m4, q = threadIdx.x - r * m4;
    while (r < 8 * kMfccGroups) {
      const int64_t t = t0 - halo + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < rows && t >= 0 && t < n_fr) v = *reinterpret_cast<const float4*>(lmb + t * M + 4 * q);
      *reinterpret_cast<float4*>(sx + r * M + 4 * q) = v;
      r += step_r;
      q += step_q;
      if (q >= m4) {
        q -= m4;
        ++r;
      }
    }
  }
  __syncthreads();
  const int Q = (n_mfcc + 3) >> 2;             // coefficient quads per row group
  const int g = threadIdx.x / Q, cq = threadIdx.x - g * Q;
  const bool worker = g < 2 * kMfccGroups;     // 18 groups of 4 rows (Q <= 10 keeps them inside 192 threads)
  float acc[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[r][i] = 0.0f;
  for (int pass = 0; pass < (Q > 10 ? 2 : 1); ++pass) {  // (n_mfcc > 40: the row groups in two passes)
    const int gg = Q > 10 ? 2 * (threadIdx.x / Q) + pass : g;
    const bool work = Q > 10 ? (threadIdx.x / Q) < kMfccGroups : worker;
    if (work) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[r][i] = 0.0f;
      int cc[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) cc[i] = cq + i * Q < n_mfcc ? cq + i * Q : n_mfcc - 1;  // (clamped duplicates are dropped below)
      for (int m = 0; m < M; m += 4) {
        float d[4][4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
          for (int i = 0; i < 4; ++i) d[k][i] = sd[(m + k) * n_mfcc + cc[i]];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float4 x = *reinterpret_cast<const float4*>(sx + (4 * gg + r) * M + m);
#pragma unroll
          for (int i = 0; i < 4; ++i) {  // same order as the generic kernel: m ascending, one FMA each
            acc[r][i] = fmaf(x.x, d[0][i], acc[r][i]);
            acc[r][i] = fmaf(x.y, d[1][i], acc[r][i]);
            acc[r][i] = fmaf(x.z, d[2][i], acc[r][i]);
            acc[r][i] = fmaf(x.w, d[3][i], acc[r][i]);
          }
        }
      }
    }
    if (Q > 10 && pass == 0) {
      // two-pass shape: park the first pass's coefficients in the scratch behind the DCT matrix
      float* park = sd + M * n_mfcc;
      if (work)
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int i = 0; i < 4; ++i) park[(threadIdx.x * 4 + r) * 4 + i] = acc[r][i];
    }
  }
  __syncthreads();  // every read of sx is done: the coefficients overwrite it
  for (int pass = 0; pass < (Q > 10 ? 2 : 1); ++pass) {
    const int gg = Q > 10 ? 2 * (threadIdx.x / Q) + pass : g;
    const bool work = Q > 10 ? (threadIdx.x / Q) < kMfccGroups : worker;
    if (work) {
      const float* park = sd + M * n_mfcc;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int row = 4 * gg + r;
        const int64_t t = t0 - halo + row;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int c = cq + i * Q;
          if (c < n_mfcc) {
            const float a = (Q > 10 && pass == 0) ? park[(threadIdx.x * 4 + r) * 4 + i] : acc[r][i];
            float v = (t >= 0 && t < n_fr) ? a : 0.0f;
            if (n_masks > 0 && t >= 0 && t < n_fr) v *= mask_value(s_table, n_masks_f, n_masks_t, static_cast<int>(t), c);
            sm[row * n_mfcc + c] = v;
          }
        }
      }
    }
  }
  __syncthreads();
  float* ob = out + (static_cast<int64_t>(b) * t_max + t0) * n_mfcc * C;
  // thread = (coefficient, quarter of the rows): consecutive rows, so x[t-1], x[t-2] slide in registers
  // and one shared-memory read per row gives all C channels
  const int per_row = n_mfcc * C;
  const int n_rows = t_max - t0 < kMfccFrames ? static_cast<int>(t_max - t0) : kMfccFrames;
  for (int u = threadIdx.x; u < 4 * n_mfcc; u += blockDim.x) {
    const int ph = u / n_mfcc, c = u - ph * n_mfcc;
    const float* col = sm + halo * n_mfcc + c;
    const int r0 = 16 * ph, r1 = r0 + 16 < n_rows ? r0 + 16 : n_rows;
    float x1 = 0.0f, x2 = 0.0f;
    if (C == 3) {
      x1 = col[(r0 - 1) * n_mfcc];
      x2 = col[(r0 - 2) * n_mfcc];
    }
    for (int r = r0; r < r1; ++r) {
      const bool live = t0 + r < n_fr;
      const float x0 = col[r * n_mfcc];
      float* o = ob + r * per_row + c * C;
      if (C == 3) {
        const float d0 = x0 - x1, d1 = x1 - x2;
        o[0] = live ? x0 : 0.0f;
        o[1] = live ? d0 : 0.0f;
        o[2] = live ? d0 - d1 : 0.0f;
        x2 = x1;
        x1 = x0;
      } else {
        o[0] = live ? x0 : 0.0f;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// float32 PCM entry of the tensor-core path.  The reference's audio is int16 / 32768 in float32
// (load_audio_file, data.py:98-99, 104-105, 116): such samples convert back to int16 exactly and the
// tensor-core kernel (whose operand split is exact for 16-bit samples) can take them.  Dense [B,
// row_stride] input; only samples below lengths[b] are looked at.  *flag is set when a sample is not
// an exact int16 / 32768: the tensor-core launch then does nothing and the generic kernel runs.
// ---------------------------------------------------------------------------------------------
__global__ void pcm_f32_to_i16_kernel(const float* __restrict__ in, const int32_t* __restrict__ lengths, int64_t row_stride,
                                      int16_t* __restrict__ out, int32_t* flag) {
  const int b = blockIdx.y;
  const int64_t len = lengths[b];
  const float* x = in + static_cast<int64_t>(b) * row_stride;
  int16_t* y = out + static_cast<int64_t>(b) * row_stride;
  bool bad = false;
  const bool vec = (row_stride & 7) == 0 && (reinterpret_cast<uintptr_t>(in) & 31) == 0;
  const int64_t n8 = vec ? len >> 3 : 0;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    if (*reinterpret_cast<volatile int32_t*>(flag) != 0) return;  // already decided: stop converting
    const float4 a = __ldg(reinterpret_cast<const float4*>(x) + 2 * i), c = __ldg(reinterpret_cast<const float4*>(x) + 2 * i + 1);
    const float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float s0 = v[2 * k] * 32768.0f, s1 = v[2 * k + 1] * 32768.0f;
      const float r0 = rintf(s0), r1 = rintf(s1);
      bad = bad || !(r0 == s0 && r1 == s1 && fabsf(s0) <= 32768.0f && fabsf(s1) <= 32768.0f && s0 != 32768.0f && s1 != 32768.0f);
      w[k] = (static_cast<uint32_t>(static_cast<int32_t>(r0)) & 0xffffu) | (static_cast<uint32_t>(static_cast<int32_t>(r1)) << 16);
    }
    *reinterpret_cast<uint4*>(y + 8 * i) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  for (int64_t i = 8 * n8 + static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < len;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float s0 = __ldg(x + i) * 32768.0f, r0 = rintf(s0);
    bad = bad || !(r0 == s0 && fabsf(s0) <= 32768.0f && s0 != 32768.0f);
    y[i] = static_cast<int16_t>(static_cast<int32_t>(r0));
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

}  // namespace b200mel
