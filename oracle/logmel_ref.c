/*
 * logmel_ref.c - plain-C CPU restatement of the reference front-end.  TEST INFRASTRUCTURE ONLY.
 *
 * Used (1) as a second, independent oracle next to oracle/logmel_oracle.py and (2) as the timed CPU
 * stand-in ("port") for the reference's TensorFlow CPU path, which cannot be installed here
 * (tensorflow is an unpinned, un-vendored dependency of the reference: requirements.txt:1).
 * Nothing in the product (speech_recognition_b200/) links or calls this file.
 *
 * Restates, in the reference's op order (speech_recognition/data.py):
 *   stft (data.py:172): frame (no padding) -> periodic Hann -> crop/zero-pad to N -> rFFT
 *   abs then square (data.py:173, 182)
 *   dense matmul with the float32 HTK mel matrix (data.py:177-182; TF builds it in float32)
 *   log(x + eps) (data.py:183), optional DCT-II * rsqrt(2M) (data.py:235),
 *   optional delta / delta-delta (data.py:319-324).
 * float32 arithmetic throughout, like TF.  Parallel over utterances (the reference parallelises
 * over examples with num_parallel_calls=AUTOTUNE, run/train.py:88).
 *
 * PARITY PIN STATUS: see oracle/logmel_oracle.py (silence golden + shape contract pinned by the
 * reference's fixtures; FFT/window/mel values pinned against independent implementations).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float re, im; } cpx;

typedef struct {
  int N, Nc, even, K, Lw, Lw_eff, H, M, n_mfcc, feature_type, use_delta;
  float eps;
  int n_fac, fac[32];
  float* window; /* [Lw_eff] */
  cpx* tw;       /* [Nc] W_Nc^j */
  cpx* post;     /* [Nc+1] W_N^k */
  float* mel;    /* [K][M] */
  float* dct;    /* [M][n_mfcc] */
} ref_plan;

static void factorize(int n, int* fac, int* n_fac) {
  int c = 0;
  while (n % 4 == 0) { fac[c++] = 4; n /= 4; }
  while (n % 2 == 0) { fac[c++] = 2; n /= 2; }
  for (int f = 3; n > 1;) {
    if (n % f == 0) { fac[c++] = f; n /= f; } else f += 2;
  }
  *n_fac = c;
}

/* Recursive decimation-in-time mixed-radix FFT: out[0..n) = DFT(in[0], in[stride], ...). */
static void fft_rec(const cpx* in, cpx* out, int n, int stride, const int* fac, const cpx* tw, int Nc, cpx* tmp) {
  if (n == 1) { out[0] = in[0]; return; }
  const int r = fac[0], m = n / r;
  for (int q = 0; q < r; ++q) fft_rec(in + q * stride, out + q * m, m, stride * r, fac + 1, tw, Nc, tmp);
  const int tstep = Nc / n; /* W_n^x = tw[x * tstep] */
  for (int k = 0; k < m; ++k) {
    cpx v[64];
    cpx* t = r <= 64 ? v : tmp;
    for (int q = 0; q < r; ++q) {
      cpx a = out[q * m + k], w = tw[((long)q * k * tstep) % Nc];
      t[q].re = a.re * w.re - a.im * w.im;
      t[q].im = a.re * w.im + a.im * w.re;
    }
    if (r == 2) {
      out[k].re = t[0].re + t[1].re; out[k].im = t[0].im + t[1].im;
      out[k + m].re = t[0].re - t[1].re; out[k + m].im = t[0].im - t[1].im;
    } else if (r == 4) {
      float ar = t[0].re + t[2].re, ai = t[0].im + t[2].im, br = t[0].re - t[2].re, bi = t[0].im - t[2].im;
      float cr = t[1].re + t[3].re, ci = t[1].im + t[3].im, dr = t[1].im - t[3].im, di = -(t[1].re - t[3].re);
      out[k].re = ar + cr; out[k].im = ai + ci;
      out[k + m].re = br + dr; out[k + m].im = bi + di;
      out[k + 2 * m].re = ar - cr; out[k + 2 * m].im = ai - ci;
      out[k + 3 * m].re = br - dr; out[k + 3 * m].im = bi - di;
    } else {
      const int rstep = Nc / r;
      for (int qo = 0; qo < r; ++qo) {
        float sr = 0.f, si = 0.f;
        for (int qi = 0; qi < r; ++qi) {
          cpx w = tw[((long)((qi * qo) % r) * rstep) % Nc];
          sr += t[qi].re * w.re - t[qi].im * w.im;
          si += t[qi].re * w.im + t[qi].im * w.re;
        }
        out[k + qo * m].re = sr; out[k + qo * m].im = si;
      }
    }
  }
}

static float hz_to_mel(float f) { return 1127.0f * logf(1.0f + f / 700.0f); }

static void linspace_f32(float start, float stop, int num, float* v) {
  if (num == 1) { v[0] = start; return; }
  float delta = (stop - start) / (float)(num - 1);
  v[0] = start;
  for (int i = 1; i < num - 1; ++i) v[i] = start + delta * (float)i;
  v[num - 1] = stop;
}

void* logmel_ref_plan_create(int feature_type, int sample_rate, int frame_length, int frame_step, int fft_length,
                             int num_mel_bins, int num_mfcc, float lower_edge_hertz, float upper_edge_hertz,
                             float epsilon, int use_delta) {
  ref_plan* p = (ref_plan*)calloc(1, sizeof(ref_plan));
  int N = fft_length;
  if (N <= 0) { N = 1; while (N < frame_length) N <<= 1; }
  p->N = N; p->K = N / 2 + 1; p->Lw = frame_length; p->H = frame_step;
  p->Lw_eff = frame_length < N ? frame_length : N;
  p->even = (N % 2 == 0); p->Nc = p->even ? N / 2 : N;
  p->M = num_mel_bins; p->feature_type = feature_type; p->use_delta = use_delta; p->eps = epsilon;
  p->n_mfcc = num_mfcc < num_mel_bins ? num_mfcc : num_mel_bins;
  factorize(p->Nc, p->fac, &p->n_fac);
  p->window = (float*)malloc(sizeof(float) * p->Lw_eff);
  if (frame_length == 1) p->window[0] = 1.0f;
  else {
    double n = (double)(frame_length + (1 - frame_length % 2) - 1);
    for (int i = 0; i < p->Lw_eff; ++i) p->window[i] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * i / n));
  }
  p->tw = (cpx*)malloc(sizeof(cpx) * p->Nc);
  for (int j = 0; j < p->Nc; ++j) {
    double a = -2.0 * M_PI * j / p->Nc;
    p->tw[j].re = (float)cos(a); p->tw[j].im = (float)sin(a);
  }
  p->post = (cpx*)malloc(sizeof(cpx) * (p->Nc + 1));
  for (int k = 0; k <= p->Nc; ++k) {
    double a = -2.0 * M_PI * k / N;
    p->post[k].re = (float)cos(a); p->post[k].im = (float)sin(a);
  }
  if (feature_type != 0) {
    const int K = p->K, M = p->M;
    p->mel = (float*)calloc((size_t)K * M, sizeof(float));
    float* lin = (float*)malloc(sizeof(float) * K);
    float* edges = (float*)malloc(sizeof(float) * (M + 2));
    linspace_f32(0.0f, (float)sample_rate / 2.0f, K, lin);
    linspace_f32(hz_to_mel(lower_edge_hertz), hz_to_mel(upper_edge_hertz), M + 2, edges);
    for (int k = 1; k < K; ++k) {
      float mel = hz_to_mel(lin[k]);
      for (int m = 0; m < M; ++m) {
        float ls = (mel - edges[m]) / (edges[m + 1] - edges[m]);
        float us = (edges[m + 2] - mel) / (edges[m + 2] - edges[m + 1]);
        float w = fminf(ls, us);
        p->mel[(size_t)k * M + m] = w > 0.0f ? w : 0.0f;
      }
    }
    free(lin); free(edges);
    if (feature_type == 2) {
      p->dct = (float*)malloc(sizeof(float) * M * p->n_mfcc);
      double scale = 2.0 / sqrt(2.0 * M);
      for (int m = 0; m < M; ++m)
        for (int c = 0; c < p->n_mfcc; ++c)
          p->dct[m * p->n_mfcc + c] = (float)(scale * cos(M_PI * c * (2.0 * m + 1.0) / (2.0 * M)));
    }
  }
  return p;
}

void logmel_ref_plan_destroy(void* vp) {
  ref_plan* p = (ref_plan*)vp;
  if (!p) return;
  free(p->window); free(p->tw); free(p->post); free(p->mel); free(p->dct); free(p);
}

int logmel_ref_feature_dim(const void* vp) {
  const ref_plan* p = (const ref_plan*)vp;
  return p->feature_type == 0 ? p->K : (p->feature_type == 1 ? p->M : p->n_mfcc);
}

int logmel_ref_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* One utterance: x float32 [len] -> feat [t_out][F] (C = 1); returns the frame count. */
static int one_utterance(const ref_plan* p, const float* x, int len, float* feat, long t_cap, cpx* a, cpx* b,
                         float* power, float* logmel, cpx* tmp) {
  const int T = len < p->Lw ? 0 : 1 + (len - p->Lw) / p->H;
  const int F = logmel_ref_feature_dim(p);
  for (long t = 0; t < T && t < t_cap; ++t) {
    const float* fr = x + t * p->H;
    if (p->even) {
      for (int j = 0; j < p->Nc; ++j) {
        int n0 = 2 * j, n1 = 2 * j + 1;
        a[j].re = n0 < p->Lw_eff ? fr[n0] * p->window[n0] : 0.0f;
        a[j].im = n1 < p->Lw_eff ? fr[n1] * p->window[n1] : 0.0f;
      }
    } else {
      for (int j = 0; j < p->Nc; ++j) { a[j].re = j < p->Lw_eff ? fr[j] * p->window[j] : 0.0f; a[j].im = 0.0f; }
    }
    fft_rec(a, b, p->Nc, 1, p->fac, p->tw, p->Nc, tmp);
    for (int k = 0; k < p->K; ++k) {
      float xr, xi;
      if (p->even) {
        cpx A = b[k == p->Nc ? 0 : k], B = b[k == 0 ? 0 : p->Nc - k];
        float er = 0.5f * (A.re + B.re), ei = 0.5f * (A.im - B.im);
        float dr = 0.5f * (A.re - B.re), di = 0.5f * (A.im + B.im);
        float or_ = di, oi = -dr; /* -i * d */
        xr = er + p->post[k].re * or_ - p->post[k].im * oi;
        xi = ei + p->post[k].re * oi + p->post[k].im * or_;
      } else { xr = b[k].re; xi = b[k].im; }
      float mag = sqrtf(xr * xr + xi * xi);          /* tf.abs */
      power[k] = p->feature_type == 0 ? mag : mag * mag; /* tf.square */
    }
    float* row = feat + t * F;
    if (p->feature_type == 0) { memcpy(row, power, sizeof(float) * p->K); continue; }
    float* dst = p->feature_type == 2 ? logmel : row;
    for (int m = 0; m < p->M; ++m) dst[m] = 0.0f;
    for (int k = 0; k < p->K; ++k) {                  /* dense matmul, like tf.matmul */
      const float pk = power[k];
      const float* w = p->mel + (size_t)k * p->M;
      for (int m = 0; m < p->M; ++m) dst[m] += pk * w[m];
    }
    for (int m = 0; m < p->M; ++m) dst[m] = logf(dst[m] + p->eps);
    if (p->feature_type == 2) {
      for (int c = 0; c < p->n_mfcc; ++c) row[c] = 0.0f;
      for (int m = 0; m < p->M; ++m) {
        const float lm = logmel[m];
        const float* d = p->dct + m * p->n_mfcc;
        for (int c = 0; c < p->n_mfcc; ++c) row[c] += lm * d[c];
      }
    }
  }
  return T;
}

/*
 * Batched run.  pcm: int16 (pcm_is_f32 = 0) or float32.  Utterance b starts at offsets[b]
 * (or b * row_stride when offsets == NULL).  out: float32 [B][t_max][F][C], zero padded.
 * n_threads <= 0: all OpenMP threads.  Returns the number of threads used.
 */
int logmel_ref_batch(const void* vp, const void* pcm, int pcm_is_f32, const int64_t* offsets, const int32_t* lengths,
                     int batch, int64_t row_stride, float* out, int64_t t_max, int32_t* n_frames, int n_threads) {
  const ref_plan* p = (const ref_plan*)vp;
  const int F = logmel_ref_feature_dim(p), C = p->use_delta ? 3 : 1;
  int used = 1;
#ifdef _OPENMP
  if (n_threads <= 0) n_threads = omp_get_max_threads();
  used = n_threads;
#pragma omp parallel num_threads(n_threads)
#endif
  {
    cpx* a = (cpx*)malloc(sizeof(cpx) * p->Nc);
    cpx* b = (cpx*)malloc(sizeof(cpx) * p->Nc);
    cpx* tmp = (cpx*)malloc(sizeof(cpx) * (p->Nc + 1));
    float* power = (float*)malloc(sizeof(float) * p->K);
    float* logmel = (float*)malloc(sizeof(float) * (p->M > 0 ? p->M : 1));
    float* xf = NULL; size_t xf_cap = 0;
    float* feat = NULL; size_t feat_cap = 0;
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
    for (int u = 0; u < batch; ++u) {
      const int len = lengths[u];
      const int64_t base = offsets ? offsets[u] : (int64_t)u * row_stride;
      const float* x;
      if (pcm_is_f32) x = (const float*)pcm + base;
      else {
        if ((size_t)len > xf_cap) { free(xf); xf_cap = (size_t)len; xf = (float*)malloc(sizeof(float) * (xf_cap + 1)); }
        const int16_t* s = (const int16_t*)pcm + base;
        for (int i = 0; i < len; ++i) xf[i] = (float)s[i] / 32768.0f; /* data.py:98,104 */
        x = xf;
      }
      float* o = out + (size_t)u * t_max * F * C;
      memset(o, 0, sizeof(float) * (size_t)t_max * F * C);
      int T;
      if (C == 1) {
        T = one_utterance(p, x, len, o, t_max, a, b, power, logmel, tmp);
      } else {
        int Tall = len < p->Lw ? 0 : 1 + (len - p->Lw) / p->H;
        long tt = Tall < t_max ? Tall : t_max;
        if ((size_t)tt * F > feat_cap) { free(feat); feat_cap = (size_t)tt * F; feat = (float*)malloc(sizeof(float) * (feat_cap + 1)); }
        T = one_utterance(p, x, len, feat, t_max, a, b, power, logmel, tmp);
        for (long t = 0; t < tt; ++t)
          for (int j = 0; j < F; ++j) {
            float x0 = feat[t * F + j], x1 = t >= 1 ? feat[(t - 1) * F + j] : 0.0f, x2 = t >= 2 ? feat[(t - 2) * F + j] : 0.0f;
            float d0 = x0 - x1, d1 = x1 - x2;
            float* q = o + ((size_t)t * F + j) * 3;
            q[0] = x0; q[1] = d0; q[2] = d0 - d1;
          }
      }
      if (n_frames) n_frames[u] = T;
    }
    free(a); free(b); free(tmp); free(power); free(logmel); free(xf); free(feat);
  }
  return used;
}
