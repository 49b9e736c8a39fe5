// Host-side plan: every constant table the kernels need, built once per handle.
//
// Follows the TensorFlow semantics the reference relies on (TF is not vendored in the reference
// tree; call sites: speech_recognition/data.py:135,172,177,221,226,235):
//   * periodic Hann window                       (tf.signal.hann_window, window_ops.py)
//   * crop / zero-pad of frames to fft_length     (tf.signal.rfft)
//   * HTK mel filterbank built in float32 in TF's op order (mel_ops.py)
//   * DCT-II * rsqrt(2M)                          (mfcc_ops.py / dct_ops.py)
#pragma once

#include <cmath>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/b200mel.h"

namespace b200mel {

struct HostPlan {
  b200mel_config cfg;
  int N = 0;        // resolved fft_length
  int K = 0;        // N/2+1 spectrogram bins
  int Lw_eff = 0;   // min(frame_length, N): samples of each frame that reach the FFT
  int Nc = 0;       // complex FFT length: N/2 (even N, packed real transform) or N (odd N)
  int even = 0;
  int F = 0;        // frequency_dim of the output
  int C = 0;        // feature_dim (1 or 3)
  int n_masks_f = 0, n_masks_t = 0;
  std::vector<int> radices;           // product == Nc
  std::vector<float> window;          // [Lw_eff]
  std::vector<float> twiddle;         // [Nc][2]  W_Nc^j = exp(-2 pi i j / Nc)
  std::vector<float> post;            // [Nc+1][2] W_N^k (even N only)
  std::vector<float> mel_dense;       // [K][M] float32, TF op order
  std::vector<int> mel_start, mel_count, mel_off;  // band form per mel bin
  std::vector<float> mel_w;           // concatenated band weights
  std::vector<float> dct;             // [M][n_mfcc_eff]
  int n_mfcc_eff = 0;
};

inline int enclosing_pow2(int n) {
  int p = 1;
  while (p < n) p <<= 1;
  return p;
}

// tf.linspace in float32: start + delta * i with the exact `stop` as last element.
inline std::vector<float> linspace_f32(float start, float stop, int num) {
  std::vector<float> v(num);
  if (num == 1) {
    v[0] = start;
    return v;
  }
  float delta = (stop - start) / static_cast<float>(num - 1);
  v[0] = start;
  for (int i = 1; i < num - 1; ++i) v[i] = start + delta * static_cast<float>(i);
  v[num - 1] = stop;
  return v;
}

inline float hertz_to_mel_f32(float f) { return 1127.0f * logf(1.0f + f / 700.0f); }

// Returns "" on success, else the error text (TF raises ValueError for the same conditions).
inline std::string build_mel(HostPlan& p) {
  const b200mel_config& c = p.cfg;
  const int M = c.num_mel_bins;
  if (M <= 0) return "num_mel_bins must be positive";
  if (c.lower_edge_hertz < 0.0f) return "lower_edge_hertz must be non-negative";
  if (c.lower_edge_hertz >= c.upper_edge_hertz) return "lower_edge_hertz >= upper_edge_hertz";
  if (c.sample_rate <= 0) return "sample_rate must be positive";
  if (c.upper_edge_hertz > static_cast<float>(c.sample_rate) / 2.0f)
    return "upper_edge_hertz must not be larger than the Nyquist frequency (sample_rate / 2)";
  const int K = p.K;
  float nyquist = static_cast<float>(c.sample_rate) / 2.0f;
  std::vector<float> lin = linspace_f32(0.0f, nyquist, K);
  std::vector<float> edges =
      linspace_f32(hertz_to_mel_f32(c.lower_edge_hertz), hertz_to_mel_f32(c.upper_edge_hertz), M + 2);
  p.mel_dense.assign(static_cast<size_t>(K) * M, 0.0f);
  for (int k = 1; k < K; ++k) {  // DC row stays zero (bands_to_zero = 1)
    float mel = hertz_to_mel_f32(lin[k]);
    for (int m = 0; m < M; ++m) {
      float lo = edges[m], ce = edges[m + 1], hi = edges[m + 2];
      float ls = (mel - lo) / (ce - lo);
      float us = (hi - mel) / (hi - ce);
      float w = fmaxf(0.0f, fminf(ls, us));
      p.mel_dense[static_cast<size_t>(k) * M + m] = w;
    }
  }
  p.mel_start.assign(M, 0);
  p.mel_count.assign(M, 0);
  p.mel_off.assign(M, 0);
  p.mel_w.clear();
  for (int m = 0; m < M; ++m) {
    int first = -1, last = -1;
    for (int k = 0; k < K; ++k)
      if (p.mel_dense[static_cast<size_t>(k) * M + m] != 0.0f) {
        if (first < 0) first = k;
        last = k;
      }
    p.mel_off[m] = static_cast<int>(p.mel_w.size());
    if (first >= 0) {
      p.mel_start[m] = first;
      p.mel_count[m] = last - first + 1;
      for (int k = first; k <= last; ++k) p.mel_w.push_back(p.mel_dense[static_cast<size_t>(k) * M + m]);
    }
  }
  if (p.mel_w.empty()) p.mel_w.push_back(0.0f);
  return "";
}

inline std::string build_plan(const b200mel_config& cfg, HostPlan& p) {
  p.cfg = cfg;
  const int ft = cfg.audio_feature_type;
  if (ft < 0 || ft > 2) return "audio_feature_type must be spectrogram, log-mel-spectrogram or mfcc";
  if (cfg.frame_length <= 0) return "frame_length must be positive";
  if (cfg.frame_step <= 0) return "frame_step must be positive";
  p.N = cfg.fft_length > 0 ? cfg.fft_length : enclosing_pow2(cfg.frame_length);
  if (p.N > 16384) return "fft_length > 16384 is not supported";
  p.K = p.N / 2 + 1;
  p.Lw_eff = cfg.frame_length < p.N ? cfg.frame_length : p.N;
  p.even = (p.N % 2 == 0) ? 1 : 0;
  p.Nc = p.even ? p.N / 2 : p.N;
  p.C = cfg.use_delta_accelerate ? 3 : 1;

  // Periodic Hann (window_ops._raised_cosine_window): n = Lw + even - 1, Lw == 1 -> [1].
  const int Lw = cfg.frame_length;
  p.window.resize(p.Lw_eff);
  if (Lw == 1) {
    p.window[0] = 1.0f;
  } else {
    double n = static_cast<double>(Lw + (1 - Lw % 2) - 1);
    for (int i = 0; i < p.Lw_eff; ++i)
      p.window[i] = static_cast<float>(0.5 - 0.5 * cos(2.0 * M_PI * static_cast<double>(i) / n));
  }

  // radix schedule: 4s, then 2, 3, 5, then whatever primes remain (naive butterfly)
  p.radices.clear();
  int rem = p.Nc;
  while (rem % 4 == 0) { p.radices.push_back(4); rem /= 4; }
  while (rem % 2 == 0) { p.radices.push_back(2); rem /= 2; }
  while (rem % 3 == 0) { p.radices.push_back(3); rem /= 3; }
  while (rem % 5 == 0) { p.radices.push_back(5); rem /= 5; }
  for (int f = 7; rem > 1;) {
    if (rem % f == 0) { p.radices.push_back(f); rem /= f; }
    else f += 2;
  }
  if (p.radices.size() > 16) return "fft_length has too many prime factors";

  p.twiddle.resize(static_cast<size_t>(p.Nc) * 2);
  for (int j = 0; j < p.Nc; ++j) {
    double a = -2.0 * M_PI * static_cast<double>(j) / static_cast<double>(p.Nc);
    p.twiddle[2 * j] = static_cast<float>(cos(a));
    p.twiddle[2 * j + 1] = static_cast<float>(sin(a));
  }
  p.post.assign(static_cast<size_t>(p.Nc + 1) * 2, 0.0f);
  if (p.even) {
    for (int k = 0; k <= p.Nc; ++k) {
      double a = -2.0 * M_PI * static_cast<double>(k) / static_cast<double>(p.N);
      p.post[2 * k] = static_cast<float>(cos(a));
      p.post[2 * k + 1] = static_cast<float>(sin(a));
    }
  }

  p.F = p.K;
  p.n_mfcc_eff = 0;
  if (ft == B200MEL_FEATURE_LOG_MEL || ft == B200MEL_FEATURE_MFCC) {
    // (DataConfig.__post_init__'s truthiness asserts live in the Python mirror; the feature
    // functions themselves only fail where TF's mel_ops validation fails.)
    std::string e = build_mel(p);
    if (!e.empty()) return e;
    p.F = cfg.num_mel_bins;
    if (ft == B200MEL_FEATURE_MFCC) {
      if (cfg.num_mfcc <= 0) return "\"num_mfcc\" is required";
      const int M = cfg.num_mel_bins;
      p.n_mfcc_eff = cfg.num_mfcc < M ? cfg.num_mfcc : M;  // TF's [:, :num_mfcc] slice saturates
      p.dct.resize(static_cast<size_t>(M) * p.n_mfcc_eff);
      double scale = 2.0 / sqrt(2.0 * M);
      for (int m = 0; m < M; ++m)
        for (int c = 0; c < p.n_mfcc_eff; ++c)
          p.dct[static_cast<size_t>(m) * p.n_mfcc_eff + c] =
              static_cast<float>(scale * cos(M_PI * c * (2.0 * m + 1.0) / (2.0 * M)));
      p.F = p.n_mfcc_eff;
    }
  }
  if (p.mel_w.empty()) p.mel_w.push_back(0.0f);
  if (p.mel_start.empty()) { p.mel_start.push_back(0); p.mel_count.push_back(0); p.mel_off.push_back(0); }
  if (p.dct.empty()) p.dct.push_back(0.0f);

  p.n_masks_f = p.n_masks_t = 0;
  if (cfg.spec_augment_enable) {
    // spec_augment enable flags (data.py:267-269): all([F, m_F]) / all([T, p, m_T])
    if (cfg.sa_F < 0 || cfg.sa_m_F < 0 || cfg.sa_T < 0 || cfg.sa_m_T < 0)
      return "SpecAugment parameters must be non-negative";
    if (cfg.sa_F != 0 && cfg.sa_m_F != 0) p.n_masks_f = cfg.sa_m_F;
    if (cfg.sa_T != 0 && cfg.sa_p != 0.0f && cfg.sa_m_T != 0) p.n_masks_t = cfg.sa_m_T;
    if (p.n_masks_f + p.n_masks_t > 16) return "more than 16 SpecAugment masks are not supported";
  }
  return "";
}

}  // namespace b200mel
