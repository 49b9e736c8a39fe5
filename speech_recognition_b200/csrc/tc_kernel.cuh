// tc_kernel.cuh - tensor-core (tcgen05) front-end kernel for the shipped parameter set
// (frame_length 320, frame_step 160, fft_length 320, 80 mel bins: libri / kspon / clovacall YAMLs).
//
// Why tensor cores: ncu on the generic kernel shows the path is issue-bound on the CUDA cores at
// 3 % DRAM throughput (profiles/r1a_generic_kernel.md); a 320-point FFT + mel + log is ~15 flop/B,
// above the FP32 ridge, so the DFT has to leave the FP32 pipe to approach the HBM roofline.
//
// Math.  n = 5 j + r (r = 0..4, j = 0..63), k = k1 + 64 k2:
//     X[k1 + 64 k2] = sum_r W5^{r k2} * Y_r[k1],   Y_r[k1] = sum_j s[5 j + r] * (w[5 j + r] W320^{(5 j + r) k1})
//   * Y_r (five 64-point "windowed, pre-twiddled" DFTs of the decimated sample streams) is a GEMM
//     [128 frames x 64] x [64 x 64] per residue on tcgen05.mma, accumulators in TMEM.
//       - A operand: the int16 PCM itself, split EXACTLY into two f16 terms (hi = rn_f16(x),
//         lo = x - hi), de-interleaved by 5 into contiguous streams y_r[e] = s[5 e + r] in shared
//         memory.  Frame t needs y_r[32 t .. 32 t + 63]: consecutive frames overlap by half a row,
//         so the A tile is a *Hankel view* of the stream - a K-major SWIZZLE_64B descriptor whose
//         row pitch (64 B) is half the row length (128 B).  Nothing is materialised twice.
//       - B operand: window x twiddle matrix in two f16 terms (22 bits), K-major SWIZZLE_128B.
//       - three passes per residue (lo*Bhi, hi*Blo, hi*Bhi) accumulate into one f32 accumulator:
//         float32-class accuracy (the dropped lo*Blo term is < 2^-22 relative per sample).
//   * the radix-5 stage over r, power, sparse mel (FFMA immediates), log and the SpecAugment /
//     delta / padding epilogue run on the CUDA cores, one thread per frame, straight from TMEM.
//
// Roles (one persistent CTA per SM, tiles of 128 frames): warp 0 MMA issuer, warps 1-7 producers (PCM ->
// exact f16 hi/lo streams), warps 8.. two or three epilogue groups (thread = frame = TMEM lane, each
// group reduces its share of the accumulator columns).  TMEM is used as a ring of eight
// 64-column accumulator slots (a tile needs five); shared memory holds B (80 KB), one A stage (85 KB),
// the output staging tile (42 KB) and the epilogue groups' hand-over zone (10 KB).  DESIGN.md section
// 4.1 has the pipeline and the measurements.
#pragma once

#include <cuda.h>

#include <cstring>
#include <vector>

#include "common.cuh"
#include "plan.h"
#include "tc_ptx.cuh"
#ifndef TC_TABLES_INC
#define TC_TABLES_INC "tc_tables_generated.inc"   // (variant builds: another column split of the epilogue groups)
#endif
#include TC_TABLES_INC

namespace b200mel {

constexpr int kTcTileM = 128;              // frames per tile
constexpr int kTcStreamBytes = 8704;       // 136 rows x 64 B (129 needed), multiple of 512
constexpr int kTcBBytes = 8192;            // one 64 x 64 f16 matrix
constexpr int kTcSmemB = 0;
constexpr int kTcSmemA = 10 * kTcBBytes;                       // 81920
constexpr int kTcSmemOut = kTcSmemA + 10 * kTcStreamBytes;     // 168960
constexpr int kTcOutPitch = 84;                                // floats per staged row (16-B aligned, conflict-free)
// hand-over zone of the three epilogue groups: TC_PARK_PITCH floats per row (partial sums of the mel bins
// two groups feed, finished values on their way to the group that stores their float4 group)
constexpr int kTcSmemPark = kTcSmemOut + kTcTileM * kTcOutPitch * 4;  // 211968
constexpr int kTcSmemCtl = kTcSmemPark + kTcTileM * TC_PARK_PITCH * 4;
constexpr int kTcSmemMask = kTcSmemCtl + 256;                  // SpecAugment tables of the CTA's tiles
constexpr int kTcMaskTiles = 64;                               // ... at most this many tiles per CTA
constexpr int kTcMaskEntries = 16;                             // ... and 8 masks (start, width) per utterance
constexpr int kTcSmemWalk = kTcSmemMask + kTcMaskTiles * kTcMaskEntries * 4;  // one work cursor per warp (16 x 32 B)
constexpr int kTcSmemList = kTcSmemWalk + 512;                 // the CTA's tiles: (utterance, tile, length, block offset)
constexpr int kTcListTiles = 256;                              // ... at most this many per CTA and launch
constexpr int kTcSmemTotal = kTcSmemList + kTcListTiles * 16 + 1024;  // + alignment slack
static_assert(kTcSmemTotal <= 232448, "shared memory budget (227 KB per CTA)");
// Warps: 0 = MMA issuer, 1-7 = producers (224 threads), then TC_NGROUPS epilogue groups of 4 warps
// (warp % 4 = TMEM lane quadrant): A, (B,) C.  TC_NGROUPS comes with the generated tables (two groups when
// the middle group's quad range is empty).
//   3 groups: 20 warps.  (Not 21: registers are allocated to a CTA in units of four warps.)  The launch
//             allocates 96 registers per thread (640 x 96 = 60 K of the 64 K); setmaxnreg moves registers
//             between the roles within 8 x 32 x P + 12 x 32 x E <= 61 440.
//   2 groups: 16 warps, 128 registers per thread at launch; 8 x 32 x 96 + 8 x 32 x 160 = 65 536.  The
//             epilogue threads are fewer and fatter: the registers hold the last TC_TAIL_K quads of a
//             tile, so that the TMEM slots go back to the MMA issuer before the arithmetic on them.
#if TC_NGROUPS == 3
#ifndef TC_REGS_E
#define TC_REGS_E 88
#endif
#ifndef TC_REGS_P
#define TC_REGS_P 104
#endif
constexpr int kTcLaunchRegs = 96;
#else
#ifndef TC_REGS_E
#define TC_REGS_E 160
#endif
#ifndef TC_REGS_P
#define TC_REGS_P 96
#endif
constexpr int kTcLaunchRegs = 128;
#endif
#ifndef TC_TAIL_K
#define TC_TAIL_K 1    // quads of a tile held in registers when the TMEM slots are released
#endif
constexpr int kTcEpilogueThreads = 128 * TC_NGROUPS;
constexpr int kTcThreads = 256 + kTcEpilogueThreads;
constexpr int kTcMmaWarp = 0;
constexpr int kTcProducerThreads = 224;
constexpr int kTcRegsEpilogue = TC_REGS_E, kTcRegsProducer = TC_REGS_P;
static_assert(8 * 32 * kTcRegsProducer + 4 * TC_NGROUPS * 32 * kTcRegsEpilogue <= kTcThreads * kTcLaunchRegs, "register pool");
constexpr int kTcBScaleLog2 = 10;          // B is stored as 2^10 * w * twiddle
// power = |Y|^2 * 2^(-2*10) (B scale) * 2^(-30) (int16 -> [-1, 1) twice)
#define TC_PSCALE 0x1p-50f

struct TcPlan {
  uint8_t* d_b = nullptr;   // 10 x 8 KB swizzled shared-memory image of the B matrices
  float log_eps = 0.0f;     // logf(epsilon): the exact value for all-zero mel energies (silence)
};

// Host: does this plan equal the parameter set the epilogue was generated for?
inline bool tc_plan_matches(const HostPlan& p) {
  const b200mel_config& c = p.cfg;
  // (MFCC: the log-mel stage on the tensor cores, DCT + masks + deltas in mfcc_post_kernel)
  if (c.audio_feature_type != B200MEL_FEATURE_LOG_MEL && c.audio_feature_type != B200MEL_FEATURE_MFCC) return false;
  if (c.frame_length != 320 || c.frame_step != 160 || p.N != 320 || c.num_mel_bins != TC_MEL_BINS) return false;
  if (static_cast<int>(p.mel_w.size()) != TC_MEL_NNZ) return false;
  for (int m = 0; m < TC_MEL_BINS; ++m)
    if (p.mel_start[m] != kTcMelStart[m] || p.mel_count[m] != kTcMelCount[m]) return false;
  for (int i = 0; i < TC_MEL_NNZ; ++i) {
    unsigned u;
    std::memcpy(&u, &p.mel_w[i], 4);
    if (u != kTcMelWBits[i]) return false;
  }
  return true;
}

// Host: build the B operand (window x twiddle, two f16 terms) directly in its swizzled shared-memory
// layout: matrix (term, r) at (term*5 + r) * 8 KB; row n (output column) = 128 B of K = 64 f16;
// SWIZZLE_128B: 16-byte chunk index ^= (row & 7).
//   Column order = kTcColKind / kTcColK1 (generated, see gen_tc_tables.cc): quad 0 = Re, Im of k1 = 1,
//   k1 = 0 (real), the k1 = 32 term sum_j (-1)^j s w (real; its phase e^{-i pi r/5} is applied in the
//   epilogue); quad q = Re k1 = 2q, Re k1 = 2q + 1, Im k1 = 2q, Im k1 = 2q + 1 ("planar pairs").
inline bool tc_plan_build(const HostPlan& p, TcPlan& tc, int max_smem_optin) {
  if (!tc_plan_matches(p) || max_smem_optin < kTcSmemTotal) return false;
  std::vector<__half> img(10 * kTcBBytes / 2);
  const double scale = static_cast<double>(1 << kTcBScaleLog2);
  for (int r = 0; r < 5; ++r)
    for (int n = 0; n < 64; ++n)
      for (int j = 0; j < 64; ++j) {
        const int s = 5 * j + r;
        const double w = static_cast<double>(p.window[s]);
        double f;
        const int kind = kTcColKind[n], k1 = kTcColK1[n];
        if (kind == 2) f = 1.0;
        else if (kind == 3) f = (j & 1) ? -1.0 : 1.0;
        else {
          const double a = -2.0 * M_PI * static_cast<double>((s * k1) % 320) / 320.0;
          f = kind == 1 ? sin(a) : cos(a);
        }
        const double v = scale * w * f;
        const __half hi = __float2half_rn(static_cast<float>(v));
        const __half lo = __float2half_rn(static_cast<float>(v - static_cast<double>(__half2float(hi))));
        const uint32_t o = 128 * n + 2 * j;
        const uint32_t ph = o ^ (((o >> 7) & 7) << 4);
        img[((0 * 5 + r) * kTcBBytes + ph) / 2] = hi;
        img[((1 * 5 + r) * kTcBBytes + ph) / 2] = lo;
      }
  if (cudaMalloc(reinterpret_cast<void**>(&tc.d_b), 10 * kTcBBytes) != cudaSuccess) return false;
  if (cudaMemcpy(tc.d_b, img.data(), 10 * kTcBBytes, cudaMemcpyHostToDevice) != cudaSuccess) return false;
  tc.log_eps = logf(p.cfg.epsilon);
  return true;
}

inline void tc_plan_free(TcPlan& tc) {
  cudaFree(tc.d_b);
  tc.d_b = nullptr;
}

}  // namespace b200mel

#include "tc_kernel_device.cuh"
