// Host check of the tensor-core kernel's GENERATED epilogue (tc_tables_generated.inc) and of its packed
// radix-5 arithmetic (tc_epilogue_math.cuh), compiled with a host compiler: test infrastructure, no GPU.
//
// For every input frame (320 int16 samples) the harness forms the accumulator columns the tcgen05 GEMMs
// produce - Y_r[col] = sum_j s[5j+r] * 2^10 w[5j+r] f_col(5j+r), column order kTcColKind / kTcColK1 - in
// double precision, then runs the three epilogue groups' code in chain order A -> B -> C exactly as the
// kernel does (quad loop, mel FFMAs, phase 1, hand-over zone, phase 2) and writes the 80 staged values.
// tests/test_tc_epilogue_host.py compares them with the numpy oracle.
//
//   g++ -O1 -std=c++17 -o tc_epilogue_host tc_epilogue_host.cc
//   ./tc_epilogue_host frames.i16 out.f32 [mask_start mask_width]
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../speech_recognition_b200/csrc/plan.h"
#include "../../speech_recognition_b200/csrc/tc_tables_generated.inc"
#include "../../speech_recognition_b200/csrc/tc_epilogue_math.cuh"

using namespace b200mel;

#define TC_PSCALE 0x1p-50f
struct float4 { float x, y, z, w; };

template <int M>
static inline float tc_masked(float v, uint32_t b0, uint32_t b1, uint32_t b2) {
  const uint32_t w = M < 32 ? b0 : (M < 64 ? b1 : b2);
  return (w & (1u << (M & 31))) ? v * 0.0f : v;
}
#define TC_FIN(M_, A) tc_masked<(M_)>(logf(A), mb0, mb1, mb2)   // (eps is the accumulators' initial value)
#define TC_ST4(G, A, B, C_, D) do { out_row[4 * (G)] = (A); out_row[4 * (G) + 1] = (B); out_row[4 * (G) + 2] = (C_); out_row[4 * (G) + 3] = (D); } while (0)
#define TC_PARK4(S, A, B, C_, D) do { park_row[4 * (S)] = (A); park_row[4 * (S) + 1] = (B); park_row[4 * (S) + 2] = (C_); park_row[4 * (S) + 3] = (D); } while (0)
#define TC_UNPARK4(S) float4{park_row[4 * (S)], park_row[4 * (S) + 1], park_row[4 * (S) + 2], park_row[4 * (S) + 3]}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  b200mel_config cfg;
  std::memset(&cfg, 0, sizeof(cfg));
  cfg.struct_size = sizeof(cfg);
  cfg.audio_feature_type = B200MEL_FEATURE_LOG_MEL;
  cfg.sample_rate = 16000;
  cfg.frame_length = 320;
  cfg.frame_step = 160;
  cfg.fft_length = 320;
  cfg.num_mel_bins = 80;
  cfg.lower_edge_hertz = 80.0f;
  cfg.upper_edge_hertz = 7600.0f;
  cfg.epsilon = 1e-12f;
  HostPlan p;
  if (!build_plan(cfg, p).empty()) return 3;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 4;
  std::vector<int16_t> pcm;
  int16_t buf[320];
  while (fread(buf, 2, 320, f) == 320) pcm.insert(pcm.end(), buf, buf + 320);
  fclose(f);
  const size_t n_frames = pcm.size() / 320;
  uint32_t mb0 = 0, mb1 = 0, mb2 = 0;
  if (argc >= 5) {  // one frequency mask [start, start + width)
    const int s = atoi(argv[3]), w = atoi(argv[4]);
    for (int m = s; m < s + w && m < 96; ++m) (m < 32 ? mb0 : (m < 64 ? mb1 : mb2)) |= 1u << (m & 31);
  }
  const float eps = cfg.epsilon;
  std::vector<float> out(n_frames * 80);
  for (size_t t = 0; t < n_frames; ++t) {
    const int16_t* s = pcm.data() + 320 * t;
    float acc[5][64];
    for (int r = 0; r < 5; ++r)
      for (int n = 0; n < 64; ++n) {
        double y = 0.0;
        for (int j = 0; j < 64; ++j) {
          const int i = 5 * j + r;
          double fc;
          const int kind = kTcColKind[n], k1 = kTcColK1[n];
          if (kind == 2) fc = 1.0;
          else if (kind == 3) fc = (j & 1) ? -1.0 : 1.0;
          else {
            const double a = -2.0 * M_PI * static_cast<double>((i * k1) % 320) / 320.0;
            fc = kind == 1 ? sin(a) : cos(a);
          }
          y += static_cast<double>(s[i]) * 1024.0 * static_cast<double>(p.window[i]) * fc;
        }
        acc[r][n] = static_cast<float>(y);
      }
    float out_row[84], park_row[TC_PARK_PITCH];
    for (float& v : out_row) v = NAN;   // every staged value must be written
    for (float& v : park_row) v = NAN;
#define GROUP_LOOP(G, Q0, Q1)                                   \
  float P[11], d[5][4];                                         \
  for (int q = (Q0); q < (Q1); ++q) {                           \
    for (int r = 0; r < 5; ++r)                                 \
      for (int c = 0; c < 4; ++c) d[r][c] = acc[r][4 * q + c];  \
    if ((G) == 0 && q == 0) tc_radix5_power_special(d, P);      \
    else tc_radix5_power_pair(d, P);                            \
    TC_G##G##_SWITCH(q, P);                                     \
  }
    {
      TC_G0_DECL
      GROUP_LOOP(0, 0, TC_SPLIT_A)
      TC_G0_PHASE1;
    }
    {
      TC_G1_DECL
      GROUP_LOOP(1, TC_SPLIT_A, TC_SPLIT_B)
      TC_G1_PHASE1;
      { TC_G1_PHASE2; }
    }
    {
      TC_G2_DECL
      GROUP_LOOP(2, TC_SPLIT_B, 16)
      TC_G2_PHASE1;
      { TC_G2_PHASE2; }
    }
    std::memcpy(out.data() + 80 * t, out_row, 80 * sizeof(float));
  }
  f = fopen(argv[2], "wb");
  if (!f) return 5;
  fwrite(out.data(), sizeof(float), out.size(), f);
  fclose(f);
  return 0;
}
