// libb200mel.so - C ABI (include/b200mel.h) over the sm_100a kernels.  No CPU fallback.
#include <atomic>
#include <cstdio>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/b200mel.h"
#include "aux_kernels.cuh"
#include "common.cuh"
#include "generic_kernel.cuh"
#include "host_io.h"
#include "ingest_kernels.cuh"
#include "plan.h"
#ifdef B200MEL_WITH_TC
#include "tc_kernel.cuh"
#endif

using namespace b200mel;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}

int cuda_fail(cudaError_t e, const char* what) {
  return fail(B200MEL_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

#define CU_CHECK(expr)                                   \
  do {                                                   \
    cudaError_t _e = (expr);                             \
    if (_e != cudaSuccess) return cuda_fail(_e, #expr);  \
  } while (0)

// Raise a kernel's dynamic shared-memory cap to the device maximum (opt-in limit minus the kernel's
// static shared memory).  The attribute belongs to the (function, device) pair, not to a handle: setting
// it to one handle's own tile size would lower the cap under every earlier handle on the device.
template <typename K>
cudaError_t raise_smem_cap(K kernel, int max_smem_optin) {
  cudaFuncAttributes fa;
  cudaError_t e = cudaFuncGetAttributes(&fa, kernel);
  if (e != cudaSuccess) return e;
  const int cap = max_smem_optin - static_cast<int>(fa.sharedSizeBytes);
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, cap > 0 ? cap : 0);
}

template <typename T>
cudaError_t upload(const std::vector<T>& v, T** dptr) {
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(dptr), v.size() * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*dptr, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
}

// ---- MFCC second stage on the tensor cores (aux_kernels.cuh: mfcc_post_mma_kernel<NT, C>) ----
struct MfccMmaArgs {
  const float* lm; float* out; const int32_t* lengths; int64_t t_max; int Lw, H; const uint32_t* dct_image;
  int n_masks_f, n_masks_t, sa_F, sa_T; float sa_p; const int32_t* mask_table; uint64_t seed, first_utt;
  int32_t* status; const int32_t* gate; int gate_value; int batch, sm_count;
};
// 0 = set the shared-memory cap (handle creation), 1 = launch
template <int NT, int C>
cudaError_t mfcc_mma_do(int what, int max_smem_optin, const MfccMmaArgs& a, cudaStream_t st) {
  auto kernel = mfcc_post_mma_kernel<NT, C>;
  if (what == 0) return raise_smem_cap(kernel, max_smem_optin);
  const int new_rows = kMfccMmaRows - (C == 3 ? 2 : 0);
  const int64_t tiles = (a.t_max + new_rows - 1) / new_rows * a.batch;
  if (tiles <= 0) return cudaSuccess;
  // one resident wave: three blocks per SM, each a contiguous run of the batch's tiles
  // (walked from the end: the rows the first stage wrote last are still in L2 - r2t: 137.5 vs 138.5 us)
  static const int rev = [] { const char* e = std::getenv("B200MEL_MFCC_REVERSE"); return e ? std::atoi(e) : 1; }();
  const int64_t slots = static_cast<int64_t>(kMfccMmaBlocksPerSm) * a.sm_count;
  dim3 grid(static_cast<unsigned>(tiles < slots ? tiles : slots));
  kernel<<<grid, kMfccMmaThreads, mfcc_post_mma_smem_bytes(8 * NT), st>>>(
      a.lm, a.out, a.lengths, a.batch, a.t_max, a.Lw, a.H, a.dct_image, rev, a.n_masks_f, a.n_masks_t,
      a.sa_F, a.sa_T, a.sa_p, a.mask_table, a.seed, a.first_utt, a.status, a.gate, a.gate_value);
  return cudaGetLastError();
}
template <int C>
cudaError_t mfcc_mma_nt(int nt, int what, int max_smem_optin, const MfccMmaArgs& a, cudaStream_t st) {
  switch (nt) {
    case 1: return mfcc_mma_do<1, C>(what, max_smem_optin, a, st);
    case 2: return mfcc_mma_do<2, C>(what, max_smem_optin, a, st);
    case 3: return mfcc_mma_do<3, C>(what, max_smem_optin, a, st);
    case 4: return mfcc_mma_do<4, C>(what, max_smem_optin, a, st);
    case 5: return mfcc_mma_do<5, C>(what, max_smem_optin, a, st);
    case 6: return mfcc_mma_do<6, C>(what, max_smem_optin, a, st);
    case 7: return mfcc_mma_do<7, C>(what, max_smem_optin, a, st);
    case 8: return mfcc_mma_do<8, C>(what, max_smem_optin, a, st);
    case 9: return mfcc_mma_do<9, C>(what, max_smem_optin, a, st);
    case 10: return mfcc_mma_do<10, C>(what, max_smem_optin, a, st);
  }
  return cudaErrorInvalidValue;
}
inline cudaError_t mfcc_mma(int n_mfcc, int C, int what, int max_smem_optin, const MfccMmaArgs& a, cudaStream_t st) {
  return C == 3 ? mfcc_mma_nt<3>(n_mfcc / 8, what, max_smem_optin, a, st) : mfcc_mma_nt<1>(n_mfcc / 8, what, max_smem_optin, a, st);
}
inline bool mfcc_mma_shape_ok(int M, int n_mfcc) { return M == kMfccMmaM && n_mfcc >= 8 && n_mfcc <= 80 && n_mfcc % 8 == 0; }

// Stream-ordered scratch that must be released on every path out of a function (error returns included).
struct AsyncScratch {
  void* p = nullptr;
  cudaStream_t st = nullptr;
  ~AsyncScratch() {
    if (p) cudaFreeAsync(p, st);
  }
};

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace

struct b200mel_handle {
  HostPlan plan;
  int device = 0;
  int sm_count = 148;
  int max_smem_optin = 0;
  // device tables
  float* d_window = nullptr;
  float* d_twiddle = nullptr;
  float* d_post = nullptr;
  int* d_mel_start = nullptr;
  int* d_mel_count = nullptr;
  int* d_mel_off = nullptr;
  float* d_mel_w = nullptr;
  float* d_dct = nullptr;
  float* d_noalloc_logmel = nullptr;  // device address of g_noalloc_logmel on this handle's device (b200mel_forward_mfcc_noalloc)
  uint32_t* d_dct_mma = nullptr;  // TF32 (hi, lo) image [2][M][pitch] of the DCT matrix (mfcc_post_mma_kernel)
  // generic kernel tiling
  int g_warps = 8, g_ft = 32, g_halo = 0, g_per_warp_floats = 0;
  size_t g_smem = 0;
#ifdef B200MEL_WITH_TC
  TcPlan tc;
#endif
  bool tc_ok = false;
  cudaMemPool_t pool = nullptr;  // MFCC on the tensor-core path: stream-ordered scratch for the log-mel stage
  std::atomic<int64_t> launches{0};
  // e2e staging (b200mel_forward_host / _submit / _wait): two slots, so that the uploads of one step
  // overlap the downloads of the previous one on the full-duplex link
  std::mutex host_mu;
  struct HostSlot {
    void* s_pcm = nullptr;
    size_t s_pcm_bytes = 0;
    float* s_out = nullptr;
    size_t s_out_bytes = 0;
    int32_t* s_small = nullptr;  // lengths | n_frames | status
    int64_t* s_offsets = nullptr;
    size_t s_small_cap = 0;
    // pinned host mirrors of the small arrays: copies from / to the caller's (usually pageable) arrays
    // would make the asynchronous calls block
    int32_t* p_small = nullptr;  // lengths | n_frames
    int64_t* p_offsets = nullptr;
    int32_t* user_n_frames = nullptr;  // where _wait delivers the frame counts
    int32_t user_batch = 0;
    cudaEvent_t ev_in[64];
    cudaEvent_t ev_done_k = nullptr, ev_done_out = nullptr;
    bool pending = false;
  } slots[2];
  cudaStream_t streams[3] = {nullptr, nullptr, nullptr};
  bool ev_init = false;
};

extern "C" {

int b200mel_version(void) { return B200MEL_VERSION; }

const char* b200mel_last_error(void) { return g_last_error.c_str(); }

static bool noalloc_resolve(int device, float** out_ptr);  // (b200mel_forward_mfcc_noalloc, below)

int b200mel_create(const b200mel_config* config, int device, b200mel_handle** out_handle) {
  if (!config || !out_handle) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null argument");
  if (config->struct_size != static_cast<int32_t>(sizeof(b200mel_config)))
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "b200mel_config.struct_size mismatch");
  *out_handle = nullptr;
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
    cudaGetLastError();
    return fail(B200MEL_ERR_NO_DEVICE, "no CUDA device available: libb200mel has no CPU path");
  }
  if (device < 0 || device >= n_dev) return fail(B200MEL_ERR_INVALID_ARGUMENT, "device index out of range");

  b200mel_handle* h = new (std::nothrow) b200mel_handle();
  if (!h) return fail(B200MEL_ERR_CUDA, "out of host memory");
  std::string err = build_plan(*config, h->plan);
  if (!err.empty()) {
    delete h;
    return fail(B200MEL_ERR_INVALID_ARGUMENT, err);
  }
  h->device = device;
  DeviceGuard guard(device);
  if (!guard.ok) {
    delete h;
    return fail(B200MEL_ERR_CUDA, "cudaSetDevice failed");
  }
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) {
    delete h;
    return cuda_fail(e, "cudaGetDeviceProperties");
  }
  h->sm_count = prop.multiProcessorCount;
  h->max_smem_optin = static_cast<int>(prop.sharedMemPerBlockOptin);

  HostPlan& p = h->plan;
  e = upload(p.window, &h->d_window);
  if (e == cudaSuccess) e = upload(p.twiddle, &h->d_twiddle);
  if (e == cudaSuccess) e = upload(p.post, &h->d_post);
  if (e == cudaSuccess) e = upload(p.mel_start, &h->d_mel_start);
  if (e == cudaSuccess) e = upload(p.mel_count, &h->d_mel_count);
  if (e == cudaSuccess) e = upload(p.mel_off, &h->d_mel_off);
  if (e == cudaSuccess) e = upload(p.mel_w, &h->d_mel_w);
  if (e == cudaSuccess) e = upload(p.dct, &h->d_dct);
  if (e != cudaSuccess) {
    b200mel_destroy(h);
    return cuda_fail(e, "uploading plan tables");
  }

  // generic kernel tiling: per-warp FFT ping-pong (2 * Nc complex) + mfcc scratch (M), and a
  // [(FT + halo), F] feature tile; shrink until it fits the opt-in shared memory.
  h->g_halo = p.C == 3 ? 2 : 0;
  h->g_per_warp_floats = 4 * p.Nc + ((p.cfg.audio_feature_type == B200MEL_FEATURE_MFCC) ? p.cfg.num_mel_bins : 0);
  h->g_per_warp_floats = (h->g_per_warp_floats + 3) & ~3;
  int warps = 8, ft = 64;
  const size_t budget = static_cast<size_t>(h->max_smem_optin > 0 ? h->max_smem_optin : 48 * 1024) - 2048;
  auto smem_need = [&](int w, int f) {
    return (static_cast<size_t>(w) * h->g_per_warp_floats + static_cast<size_t>(f + h->g_halo) * p.F) * sizeof(float);
  };
  while (ft > 8 && static_cast<size_t>(ft + h->g_halo) * p.F * sizeof(float) > 48 * 1024) ft /= 2;
  while (warps > 1 && smem_need(warps, ft) > budget) warps /= 2;
  while (ft > 1 && smem_need(warps, ft) > budget) ft /= 2;
  if (smem_need(warps, ft) > budget) {
    b200mel_destroy(h);
    return fail(B200MEL_ERR_UNSUPPORTED, "configuration does not fit in shared memory");
  }
  h->g_warps = warps;
  h->g_ft = ft;
  h->g_smem = smem_need(warps, ft);
  e = raise_smem_cap(generic_frontend_kernel, h->max_smem_optin);
  if (e != cudaSuccess) {
    b200mel_destroy(h);
    return cuda_fail(e, "cudaFuncSetAttribute(generic_frontend_kernel)");
  }
#ifdef B200MEL_WITH_TC
  h->tc_ok = tc_plan_build(p, h->tc, h->max_smem_optin);
  if (h->tc_ok) {
    // the epilogue's fast log must reproduce logf(epsilon) exactly for silence (see tc_finish)
    float* d_probe = nullptr;
    float probe = 0.0f;
    if (cudaMalloc(reinterpret_cast<void**>(&d_probe), sizeof(float)) == cudaSuccess) {
      tc_eps_probe_kernel<<<1, 1>>>(p.cfg.epsilon, d_probe);
      if (cudaMemcpy(&probe, d_probe, sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess) h->tc_ok = false;
      cudaFree(d_probe);
    } else {
      h->tc_ok = false;
    }
    cudaGetLastError();
    if (std::memcmp(&probe, &h->tc.log_eps, sizeof(float)) != 0) h->tc_ok = false;
  }
  if (h->tc_ok) {
    // per-device function attribute: set it while this handle's device is current
    auto cap = [](auto kernel) {
      return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemTotal) != cudaSuccess;
    };
    if (cap(tc_frontend_kernel<1, false, false>) || cap(tc_frontend_kernel<1, true, false>) ||
        cap(tc_frontend_kernel<3, false, false>) || cap(tc_frontend_kernel<3, true, false>) ||
        cap(tc_frontend_kernel<1, false, true>) || cap(tc_frontend_kernel<1, true, true>) ||
        cap(tc_frontend_kernel<3, false, true>) || cap(tc_frontend_kernel<3, true, true>)) {
      cudaGetLastError();
      h->tc_ok = false;
    }
  }
  if (h->tc_ok) {
    // scratch of the two-stage paths (MFCC: the log-mel stage; float32 PCM: the int16 copy) comes from a
    // pool that keeps it between calls (the device's default pool releases at every synchronisation)
    cudaMemPoolProps props;
    std::memset(&props, 0, sizeof(props));
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    uint64_t keep = ~0ull;
    if (cudaMemPoolCreate(&h->pool, &props) != cudaSuccess ||
        cudaMemPoolSetAttribute(h->pool, cudaMemPoolAttrReleaseThreshold, &keep) != cudaSuccess) {
      cudaGetLastError();
      h->tc_ok = false;
    }
  }
  if (h->tc_ok && p.cfg.audio_feature_type == B200MEL_FEATURE_MFCC) {
    const size_t smem = mfcc_post_smem_bytes(p.cfg.num_mel_bins, p.n_mfcc_eff);
    if (p.cfg.num_mel_bins > kMfccMaxM || p.n_mfcc_eff > kMfccMaxC || smem + 2048 > static_cast<size_t>(h->max_smem_optin) ||
        raise_smem_cap(mfcc_post_kernel, h->max_smem_optin) != cudaSuccess) {
      cudaGetLastError();
      h->tc_ok = false;
    }
  }
  if (h->tc_ok && p.cfg.audio_feature_type == B200MEL_FEATURE_MFCC && mfcc_mma_shape_ok(p.cfg.num_mel_bins, p.n_mfcc_eff)) {
    // the DCT matrix split into f16 (hi, lo) parts, in the shared-memory layout of mfcc_post_mma_kernel: rows
    // folded by the mirror symmetry (m < 40; parity 0 = even coefficients, 1 = odd), entry (parity, k-step j,
    // tq, n) = (b0 hi, b1 hi, b0 lo, b1 lo) for coefficient 2 n + parity, b0 = fragment columns (2 tq,
    // 2 tq + 1), b1 = (2 tq + 8, 2 tq + 9), each a packed pair of halves (low half = the even column)
    const int M = p.cfg.num_mel_bins, nm = p.n_mfcc_eff, PN = mfcc_mma_image_pitch(8 * mfcc_mma_ntiles(nm));
    std::vector<uint32_t> img(mfcc_mma_image_words(nm), 0u);
    auto split = [&](int k, int c, uint16_t& hi, uint16_t& lo) {
      const float d = (k < M / 2 && c < nm) ? p.dct[static_cast<size_t>(k) * nm + c] : 0.0f;
      const __half h = __float2half_rn(d);
      const __half l = __float2half_rn(d - __half2float(h));
      std::memcpy(&hi, &h, 2);
      std::memcpy(&lo, &l, 2);
    };
    for (int par = 0; par < 2; ++par)
      for (int j = 0; j < kMfccMmaKSteps; ++j)
        for (int tq = 0; tq < 4; ++tq)
          for (int n = 0; n < 8 * mfcc_mma_ntiles(nm); ++n) {
            uint32_t* e = &img[(static_cast<size_t>((par * kMfccMmaKSteps + j) * 4 + tq) * PN + n) * 4];
            for (int part = 0; part < 2; ++part) {   // b0, b1
              uint16_t h0, l0, h1, l1;
              split(mfcc_mma_image_k(j, 2 * tq + 8 * part), 2 * n + par, h0, l0);
              split(mfcc_mma_image_k(j, 2 * tq + 8 * part + 1), 2 * n + par, h1, l1);
              e[part] = static_cast<uint32_t>(h0) | (static_cast<uint32_t>(h1) << 16);
              e[2 + part] = static_cast<uint32_t>(l0) | (static_cast<uint32_t>(l1) << 16);
            }
          }
    MfccMmaArgs none{};
    if (upload(img, &h->d_dct_mma) != cudaSuccess ||
        mfcc_mma(nm, 1, 0, h->max_smem_optin, none, nullptr) != cudaSuccess ||
        mfcc_mma(nm, 3, 0, h->max_smem_optin, none, nullptr) != cudaSuccess) {
      cudaGetLastError();
      cudaFree(h->d_dct_mma);
      h->d_dct_mma = nullptr;  // (the scalar second stage stays available)
    }
  }
  if (h->tc_ok && p.cfg.audio_feature_type == B200MEL_FEATURE_MFCC) {
    // set-up of b200mel_forward_mfcc_noalloc: this device is current here (DeviceGuard above)
    if (!noalloc_resolve(device, &h->d_noalloc_logmel)) h->d_noalloc_logmel = nullptr;
  }
#endif
  *out_handle = h;
  return B200MEL_OK;
}

int b200mel_destroy(b200mel_handle* h) {
  if (!h) return B200MEL_OK;
  DeviceGuard guard(h->device);
  cudaFree(h->d_window);
  cudaFree(h->d_twiddle);
  cudaFree(h->d_post);
  cudaFree(h->d_mel_start);
  cudaFree(h->d_mel_count);
  cudaFree(h->d_mel_off);
  cudaFree(h->d_mel_w);
  cudaFree(h->d_dct);
  cudaFree(h->d_dct_mma);
  for (auto& sl : h->slots) {
    cudaFree(sl.s_pcm);
    cudaFree(sl.s_out);
    cudaFree(sl.s_small);
    cudaFree(sl.s_offsets);
    cudaFreeHost(sl.p_small);
    cudaFreeHost(sl.p_offsets);
  }
#ifdef B200MEL_WITH_TC
  tc_plan_free(h->tc);
#endif
  if (h->pool) cudaMemPoolDestroy(h->pool);
  for (auto& s : h->streams)
    if (s) cudaStreamDestroy(s);
  if (h->ev_init)
    for (auto& sl : h->slots) {
      for (auto& ev : sl.ev_in) cudaEventDestroy(ev);
      cudaEventDestroy(sl.ev_done_k);
      cudaEventDestroy(sl.ev_done_out);
    }
  delete h;
  return B200MEL_OK;
}

int b200mel_num_frames(const b200mel_handle* h, int64_t n_samples, int64_t* n_frames) {
  if (!h || !n_frames) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null argument");
  const int64_t Lw = h->plan.cfg.frame_length, H = h->plan.cfg.frame_step;
  *n_frames = n_samples < Lw ? 0 : 1 + (n_samples - Lw) / H;
  return B200MEL_OK;
}

int b200mel_feature_shape(const b200mel_handle* h, int32_t* frequency_dim, int32_t* feature_dim, int32_t* fft_length) {
  if (!h) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null handle");
  if (frequency_dim) *frequency_dim = h->plan.F;
  if (feature_dim) *feature_dim = h->plan.C;
  if (fft_length) *fft_length = h->plan.N;
  return B200MEL_OK;
}

int b200mel_mel_weights(const b200mel_handle* h, float* out_host) {
  if (!h || !out_host) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null argument");
  if (h->plan.mel_dense.empty()) return fail(B200MEL_ERR_INVALID_ARGUMENT, "feature type has no mel filterbank");
  std::memcpy(out_host, h->plan.mel_dense.data(), h->plan.mel_dense.size() * sizeof(float));
  return B200MEL_OK;
}

#if defined(TC_TRACE) && defined(B200MEL_WITH_TC)
// developer-only (variant builds): copy the role trace of CTA 0 to the host
int b200mel_debug_trace(long long* out_host) {
  return cudaMemcpyFromSymbol(out_host, g_tc_trace, sizeof(g_tc_trace)) == cudaSuccess ? 0 : 3;
}
#endif

int b200mel_tc_supported(const b200mel_handle* h) { return (h && h->tc_ok) ? 1 : 0; }

int64_t b200mel_launch_count(const b200mel_handle* h) { return h ? h->launches.load() : 0; }

static void fill_params(const b200mel_handle* h, KernelParams& kp) {
  const HostPlan& p = h->plan;
  std::memset(&kp, 0, sizeof(kp));
  kp.Lw = p.cfg.frame_length;
  kp.H = p.cfg.frame_step;
  kp.N = p.N;
  kp.K = p.K;
  kp.Lw_eff = p.Lw_eff;
  kp.Nc = p.Nc;
  kp.even = p.even;
  kp.n_radices = static_cast<int>(p.radices.size());
  for (size_t i = 0; i < p.radices.size(); ++i) kp.radices[i] = p.radices[i];
  kp.window = h->d_window;
  kp.twiddle = reinterpret_cast<const float2*>(h->d_twiddle);
  kp.post = reinterpret_cast<const float2*>(h->d_post);
  kp.feature_type = p.cfg.audio_feature_type;
  kp.M = p.cfg.num_mel_bins;
  kp.n_mfcc = p.n_mfcc_eff;
  kp.F = p.F;
  kp.C = p.C;
  kp.epsilon = p.cfg.epsilon;
  kp.mel_start = h->d_mel_start;
  kp.mel_count = h->d_mel_count;
  kp.mel_off = h->d_mel_off;
  kp.mel_w = h->d_mel_w;
  kp.dct = h->d_dct;
  kp.n_masks_f = p.n_masks_f;
  kp.n_masks_t = p.n_masks_t;
  kp.sa_F = p.cfg.sa_F;
  kp.sa_T = p.cfg.sa_T;
  kp.sa_p = p.cfg.sa_p;
  kp.frames_per_tile = h->g_ft;
  kp.halo = h->g_halo;
  kp.warps = h->g_warps;
  kp.smem_per_warp_floats = h->g_per_warp_floats;
  kp.gate = nullptr;
  kp.gate_value = 0;
  kp.f32_flag = nullptr;
}

int b200mel_forward(b200mel_handle* h, const void* pcm, int32_t pcm_dtype, const int64_t* offsets,
                    const int32_t* lengths, int32_t batch, int64_t row_stride, float* out, int64_t t_max,
                    int32_t* n_frames, const int32_t* mask_table, uint64_t seed, uint64_t first_utterance_id,
                    int32_t* status, int32_t kernel, void* stream) {
  if (!h) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null handle");
  if (batch < 0 || t_max < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative batch or t_max");
  if (pcm_dtype != B200MEL_PCM_I16 && pcm_dtype != B200MEL_PCM_F32)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "pcm_dtype must be B200MEL_PCM_I16 or B200MEL_PCM_F32");
  if (kernel < B200MEL_KERNEL_AUTO || kernel > B200MEL_KERNEL_TC)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "unknown kernel selector");
  if (kernel == B200MEL_KERNEL_TC && !h->tc_ok)
    return fail(B200MEL_ERR_UNSUPPORTED, "the tensor-core kernel does not support this configuration");
  if (kernel == B200MEL_KERNEL_TC && pcm_dtype != B200MEL_PCM_I16)
    return fail(B200MEL_ERR_UNSUPPORTED, "the tensor-core kernel takes int16 PCM (its operand split is exact for 16-bit samples)");
  if (batch == 0) return B200MEL_OK;
  if (!lengths || (!pcm) || (t_max > 0 && !out)) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  if (batch > 65535) return fail(B200MEL_ERR_UNSUPPORTED, "batch > 65535: split the call");
  if (t_max > (int64_t{1} << 30)) return fail(B200MEL_ERR_UNSUPPORTED, "t_max too large");
  DeviceGuard guard(h->device);
  if (!guard.ok) return fail(B200MEL_ERR_CUDA, "cudaSetDevice failed");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

  KernelParams kp;
  fill_params(h, kp);
  kp.pcm = pcm;
  kp.offsets = offsets;
  kp.lengths = lengths;
  kp.row_stride = row_stride;
  kp.pcm_is_f32 = pcm_dtype == B200MEL_PCM_F32;
  kp.batch = batch;
  kp.out = out;
  kp.t_max = t_max;
  kp.n_frames = n_frames;
  kp.status = status;
  kp.mask_table = mask_table;
  kp.seed = seed;
  kp.first_utt = first_utterance_id;

  if (t_max == 0) {
    // nothing to write but n_frames: one tile per utterance still reports the frame counts
    if (!n_frames) return B200MEL_OK;
  }
#ifdef B200MEL_WITH_TC
  // the tensor-core kernel stores rows with 16-byte bulk copies (C = 1) / 256-bit stores (C = 3)
  const bool tc_aligned = (reinterpret_cast<uintptr_t>(out) & (kp.C == 3 ? 31 : 15)) == 0;
  if (kernel == B200MEL_KERNEL_TC && !tc_aligned)
    return fail(B200MEL_ERR_INVALID_ARGUMENT,
                "the tensor-core kernel needs a 16-byte (with delta features: 32-byte) aligned output buffer");
  if (kernel == B200MEL_KERNEL_TC && kp.n_masks_f + kp.n_masks_t > 16)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "the tensor-core kernel handles at most 16 SpecAugment masks per utterance");
  // (one utterance must fit one launch of the tensor-core kernel: t_max <= ~9.5 M frames)
  const bool tc_fits = t_max > 0 && tc_max_tiles_per_launch(h->sm_count) / ((t_max + (kp.C == 3 ? 126 : 128) - 1) / (kp.C == 3 ? 126 : 128)) >= 1;
  if (kernel == B200MEL_KERNEL_TC && t_max > 0 && !tc_fits)
    return fail(B200MEL_ERR_UNSUPPORTED, "t_max is too large for the tensor-core kernel (use B200MEL_KERNEL_GENERIC)");
  // float32 PCM, dense rows, kernel = auto: the reference's audio is int16 / 32768 (data.py:98-116), which the
  // tensor-core kernel can take exactly.  Both kernels are launched and a device flag decides - no host round
  // trip.  16-byte aligned rows: the tensor-core kernel reads the floats itself (its producers convert and
  // test them, 369 MB moved instead of 614 MB) and sets the flag; otherwise a conversion kernel writes an
  // int16 copy first and the tensor-core launch is gated on the flag as well.
  AsyncScratch f32_guard;  // (freed behind both kernels of the pair, and on every error return)
  f32_guard.st = st;
  void*& f32_scratch = f32_guard.p;
  const KernelParams kp_f32 = kp;
  const bool f32_gated = h->tc_ok && kernel == B200MEL_KERNEL_AUTO && t_max > 0 && pcm_dtype == B200MEL_PCM_F32 && !offsets &&
                         tc_aligned && kp.n_masks_f + kp.n_masks_t <= 16 && tc_fits && row_stride > 0;
  static const bool f32_fused_ok = [] { const char* e = std::getenv("B200MEL_F32_TWO_PASS"); return !(e && std::atoi(e)); }();
  const bool f32_fused = f32_gated && f32_fused_ok && (reinterpret_cast<uintptr_t>(pcm) & 15) == 0 && (row_stride & 7) == 0 &&
                         static_cast<int64_t>(batch) * (row_stride >> 3) < 0x7fffffffLL;
  if (f32_gated) {
    const size_t n = static_cast<size_t>(batch) * static_cast<size_t>(row_stride);
    CU_CHECK(cudaMallocFromPoolAsync(&f32_scratch, 64 + (f32_fused ? 0 : n * sizeof(int16_t)), h->pool, st));
    int32_t* flag = static_cast<int32_t*>(f32_scratch);
    CU_CHECK(cudaMemsetAsync(flag, 0, sizeof(int32_t), st));
    if (f32_fused) {
      kp.f32_flag = flag;   // (pcm_is_f32 stays 1: tc_launch_one picks the float32 instantiation)
    } else {
      int16_t* pcm16 = reinterpret_cast<int16_t*>(static_cast<char*>(f32_scratch) + 64);
      const int64_t per_row_threads = (row_stride + 7) / 8;
      // (four groups of eight samples per thread and round: 1024 groups per block-round)
      dim3 cgrid(static_cast<unsigned>(std::min<int64_t>(16, (per_row_threads + 1023) / 1024)), static_cast<unsigned>(batch));
      pcm_f32_to_i16_kernel<<<cgrid, 256, 0, st>>>(static_cast<const float*>(pcm), lengths, row_stride, pcm16, flag);
      h->launches.fetch_add(1);
      kp.pcm = pcm16;
      kp.pcm_is_f32 = 0;
      kp.gate = flag;
      kp.gate_value = 0;
    }
  }
  // (the tensor-core kernel keeps one mask-table entry per lane: at most 16 masks per utterance)
  if (h->tc_ok && kernel != B200MEL_KERNEL_GENERIC && t_max > 0 && (pcm_dtype == B200MEL_PCM_I16 || f32_gated) && tc_aligned &&
      kp.n_masks_f + kp.n_masks_t <= 16 && tc_fits) {
    const int n_masks = kp.n_masks_f + kp.n_masks_t;
    int32_t* tmp_table = nullptr;
    if (n_masks > 0 && !mask_table && h->plan.cfg.audio_feature_type != B200MEL_FEATURE_MFCC &&
        !tc_draws_in_kernel(kp, h->sm_count)) {
      // pre-pass: the utterances' mask tables (same Philox stream as b200mel_specaug_draw), in a
      // stream-ordered scratch buffer so that concurrent callers of one handle do not share state
      CU_CHECK(cudaMallocAsync(reinterpret_cast<void**>(&tmp_table),
                               static_cast<size_t>(batch) * 2 * n_masks * sizeof(int32_t), st));
      specaug_draw_from_lengths_kernel<<<(batch + 127) / 128, 128, 0, st>>>(
          lengths, batch, kp.Lw, kp.H, kp.F, kp.sa_F, kp.n_masks_f, kp.sa_T, kp.sa_p, kp.n_masks_t, seed,
          first_utterance_id, tmp_table, status);
      kp.mask_table = tmp_table;
      h->launches.fetch_add(1);
    }
    // MFCC: the log-mel stage (C = 1, no masks) goes to a scratch buffer, the second kernel does the
    // DCT, the masks on the coefficients and the deltas
    const bool two_stage = h->plan.cfg.audio_feature_type == B200MEL_FEATURE_MFCC;
    float* lm = nullptr;
    KernelParams kp1 = kp;
    if (two_stage) {
      if (tmp_table) {  // (not reached: the pre-pass is only taken with masks, which stage 1 does not apply)
        cudaFreeAsync(tmp_table, st);
        tmp_table = nullptr;
      }
      CU_CHECK(cudaMallocFromPoolAsync(reinterpret_cast<void**>(&lm), static_cast<size_t>(batch) * t_max * kp.M * sizeof(float),
                                       h->pool, st));
      kp1.feature_type = B200MEL_FEATURE_LOG_MEL;
      kp1.F = kp.M;
      kp1.C = 1;
      kp1.n_masks_f = kp1.n_masks_t = 0;
      kp1.mask_table = nullptr;
      kp1.out = lm;
    }
    int n_launched = 0;
    int rc = tc_launch(h->tc, kp1, h->sm_count, st, &n_launched);
    if (tmp_table) cudaFreeAsync(tmp_table, st);
    h->launches.fetch_add(n_launched);
    if (two_stage) {
      // (float32 input: the second stage runs only when every sample was exact - the flag is final when the
      //  first stage, which sets it in the fused form, has completed)
      const int32_t* gate2 = f32_gated ? static_cast<const int32_t*>(f32_scratch) : nullptr;
      if (rc == 0) {
        if (h->d_dct_mma && (reinterpret_cast<uintptr_t>(out) & 31) == 0) {
          const MfccMmaArgs a{lm, out, lengths, t_max, kp.Lw, kp.H, h->d_dct_mma, kp.n_masks_f, kp.n_masks_t, kp.sa_F,
                              kp.sa_T, kp.sa_p, mask_table, seed, first_utterance_id, status, gate2, 0,
                              batch, h->sm_count};
          if (mfcc_mma(kp.n_mfcc, kp.C, 1, h->max_smem_optin, a, st) != cudaSuccess) rc = 1;
        } else {
          dim3 grid(static_cast<unsigned>((t_max + kMfccFrames - 1) / kMfccFrames), static_cast<unsigned>(batch));
          mfcc_post_kernel<<<grid, kMfccThreads, mfcc_post_smem_bytes(kp.M, kp.n_mfcc), st>>>(
              lm, out, lengths, t_max, kp.Lw, kp.H, kp.M, kp.n_mfcc, kp.C, kp.dct, kp.n_masks_f, kp.n_masks_t, kp.sa_F,
              kp.sa_T, kp.sa_p, mask_table, seed, first_utterance_id, status, gate2, 0);
          if (cudaGetLastError() != cudaSuccess) rc = 1;
        }
        h->launches.fetch_add(1);
      }
      cudaFreeAsync(lm, st);
    }
    if (rc == 2) return fail(B200MEL_ERR_CUDA, "cuTensorMapEncodeTiled failed for the output tensor");
    if (rc == 3) return fail(B200MEL_ERR_UNSUPPORTED, "t_max is too large for the tensor-core kernel (use B200MEL_KERNEL_GENERIC)");
    if (rc != 0) return cuda_fail(cudaGetLastError(), "tc_frontend_kernel launch");
    if (!f32_gated) return B200MEL_OK;
    // the other half of the pair: the generic kernel on the original float32 samples, if any was inexact
    kp = kp_f32;
    kp.gate = static_cast<const int32_t*>(f32_scratch);
    kp.gate_value = 1;
  }
#endif
  int64_t tiles = (t_max + h->g_ft - 1) / h->g_ft;
  if (tiles == 0) tiles = 1;
  dim3 grid(static_cast<unsigned>(tiles), static_cast<unsigned>(batch));
  generic_frontend_kernel<<<grid, h->g_warps * 32, h->g_smem, st>>>(kp);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "generic_frontend_kernel launch");
  h->launches.fetch_add(1);
  return B200MEL_OK;
}

// ---------------------------------------------------------------------------------------------
// b200mel_forward_mfcc_noalloc: the two-stage MFCC path of b200mel_forward with the log-mel stage in a
// file-scope device array instead of pool memory, for callers that may not allocate (CUDA graph
// capture, serving loops).  No counterpart in the reference.
// The array is per device (every GPU that loads the library holds it until exit) and
// shared by every call on that device.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_noalloc_logmel[B200MEL_NOALLOC_SCRATCH_BYTES / sizeof(float)];

// In host code `g_noalloc_logmel` is only a stand-in for the symbol: its device address is resolved
// once per device, with that device current (b200mel_create calls this under its DeviceGuard).
static std::mutex g_noalloc_mu;
static float* g_noalloc_ptr[64];
static bool noalloc_resolve(int device, float** out_ptr) {
  if (device < 0 || device >= 64) return false;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_logmel);
    if (e != cudaSuccess || !ptr) {
      cudaGetLastError();
      return false;
    }
    g_noalloc_ptr[device] = ptr;
  }
  *out_ptr = g_noalloc_ptr[device];
  return true;
}

int b200mel_forward_mfcc_noalloc(b200mel_handle* h, const int16_t* pcm, const int64_t* offsets, const int32_t* lengths,
                                 int32_t batch, int64_t row_stride, float* out, int64_t t_max, int32_t* n_frames,
                                 const int32_t* mask_table, uint64_t seed, uint64_t first_utterance_id, int32_t* status,
                                 void* stream) {
  if (!h) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null handle");
  if (batch < 0 || t_max < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative batch or t_max");
#ifdef B200MEL_WITH_TC
  if (h->plan.cfg.audio_feature_type != B200MEL_FEATURE_MFCC || !h->tc_ok)
    return fail(B200MEL_ERR_UNSUPPORTED,
                "b200mel_forward_mfcc_noalloc: the handle must be an MFCC configuration the tensor-core kernel supports");
  if (!h->d_noalloc_logmel)
    return fail(B200MEL_ERR_CUDA, "b200mel_forward_mfcc_noalloc: the static scratch array could not be resolved on this device");
  if (batch == 0 || t_max == 0) return B200MEL_OK;
  if (!lengths || !pcm || !out) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  if (batch > 65535) return fail(B200MEL_ERR_UNSUPPORTED, "batch > 65535: split the call");
  KernelParams kp;
  fill_params(h, kp);
  const int64_t max_rows = static_cast<int64_t>(B200MEL_NOALLOC_SCRATCH_BYTES / sizeof(float)) / kp.M;
  if (t_max > max_rows || static_cast<int64_t>(batch) * t_max > max_rows) {
    char msg[256];
    std::snprintf(msg, sizeof(msg),
                  "b200mel_forward_mfcc_noalloc: batch * t_max = %lld frames exceeds the static scratch (%lld frames of %d mel "
                  "bins, %u bytes): split the batch or use b200mel_forward",
                  static_cast<long long>(batch) * static_cast<long long>(t_max), static_cast<long long>(max_rows), kp.M,
                  static_cast<unsigned>(B200MEL_NOALLOC_SCRATCH_BYTES));
    return fail(B200MEL_ERR_UNSUPPORTED, msg);
  }
  if (kp.n_masks_f + kp.n_masks_t > 16)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "the tensor-core kernel handles at most 16 SpecAugment masks per utterance");
  if (reinterpret_cast<uintptr_t>(out) & 31)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "b200mel_forward_mfcc_noalloc needs a 32-byte aligned output buffer");
  const int64_t tiles_per_utt = (t_max + 127) / 128;
  if (tc_max_tiles_per_launch(h->sm_count) / tiles_per_utt < 1)
    return fail(B200MEL_ERR_UNSUPPORTED, "t_max is too large for the tensor-core kernel");
  DeviceGuard guard(h->device);
  if (!guard.ok) return fail(B200MEL_ERR_CUDA, "cudaSetDevice failed");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  kp.pcm = pcm;
  kp.offsets = offsets;
  kp.lengths = lengths;
  kp.row_stride = row_stride;
  kp.pcm_is_f32 = 0;
  kp.batch = batch;
  kp.out = out;
  kp.t_max = t_max;
  kp.n_frames = n_frames;
  kp.status = status;
  kp.mask_table = mask_table;
  kp.seed = seed;
  kp.first_utt = first_utterance_id;
  // stage 1: log-mel [batch, t_max, M] (C = 1, no masks) into the static array - every row below t_max
  // is written (padding rows as +0.0), so nothing of the previous call's contents is read
  float* lm = h->d_noalloc_logmel;
  KernelParams kp1 = kp;
  kp1.feature_type = B200MEL_FEATURE_LOG_MEL;
  kp1.F = kp.M;
  kp1.C = 1;
  kp1.n_masks_f = kp1.n_masks_t = 0;
  kp1.mask_table = nullptr;
  kp1.out = lm;
  int n_launched = 0;
  const int rc = tc_launch(h->tc, kp1, h->sm_count, st, &n_launched);
  h->launches.fetch_add(n_launched);
  if (rc == 2) return fail(B200MEL_ERR_CUDA, "cuTensorMapEncodeTiled failed for the scratch tensor");
  if (rc != 0) return cuda_fail(cudaGetLastError(), "tc_frontend_kernel launch");
  // stage 2: DCT, masks on the coefficients, deltas
  if (h->d_dct_mma) {
    const MfccMmaArgs a{lm, out, lengths, t_max, kp.Lw, kp.H, h->d_dct_mma, kp.n_masks_f, kp.n_masks_t, kp.sa_F,
                        kp.sa_T, kp.sa_p, mask_table, seed, first_utterance_id, status, nullptr, 0, batch, h->sm_count};
    const cudaError_t e = mfcc_mma(kp.n_mfcc, kp.C, 1, h->max_smem_optin, a, st);
    if (e != cudaSuccess) return cuda_fail(e, "mfcc_post_mma_kernel launch");
  } else {
    dim3 grid(static_cast<unsigned>((t_max + kMfccFrames - 1) / kMfccFrames), static_cast<unsigned>(batch));
    mfcc_post_kernel<<<grid, kMfccThreads, mfcc_post_smem_bytes(kp.M, kp.n_mfcc), st>>>(
        lm, out, lengths, t_max, kp.Lw, kp.H, kp.M, kp.n_mfcc, kp.C, kp.dct, kp.n_masks_f, kp.n_masks_t, kp.sa_F, kp.sa_T,
        kp.sa_p, mask_table, seed, first_utterance_id, status, nullptr, 0);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "mfcc_post_kernel launch");
  }
  h->launches.fetch_add(1);
  return B200MEL_OK;
#else
  (void)pcm; (void)offsets; (void)lengths; (void)row_stride; (void)out; (void)n_frames; (void)mask_table; (void)seed;
  (void)first_utterance_id; (void)status; (void)stream;
  return fail(B200MEL_ERR_UNSUPPORTED, "built without the tensor-core kernel");
#endif
}

static int host_wait_locked(b200mel_handle* h, int slot) {
  b200mel_handle::HostSlot& sl = h->slots[slot];
  if (!sl.pending) return B200MEL_OK;
  sl.pending = false;
  CU_CHECK(cudaEventSynchronize(sl.ev_done_k));
  CU_CHECK(cudaEventSynchronize(sl.ev_done_out));
  if (sl.user_n_frames) std::memcpy(sl.user_n_frames, sl.p_small + sl.user_batch, sl.user_batch * sizeof(int32_t));
  sl.user_n_frames = nullptr;
  // the step's device status word (b200mel_forward's `status`), delivered with the frame counts: the
  // results are in place, but the reference would have raised on this batch
  const int32_t st_word = sl.p_small[2 * sl.user_batch];
  if (st_word & 0x100)
    return fail(B200MEL_ERR_CUDA, "tensor-core kernel: a pipeline wait timed out (internal protocol error)");
  if (st_word & 1)
    return fail(B200MEL_ERR_INVALID_ARGUMENT,
                "SpecAugment: a draw had maxval <= minval (F > frequency_dim, or an utterance with too few frames); "
                "tf.random.uniform raises here (data.py:284-285, 297) - the mask was left empty");
  return B200MEL_OK;
}

int b200mel_forward_host_wait(b200mel_handle* h, int32_t slot) {
  if (!h) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null handle");
  if (slot < 0 || slot > 1) return fail(B200MEL_ERR_INVALID_ARGUMENT, "slot must be 0 or 1");
  std::lock_guard<std::mutex> lock(h->host_mu);
  return host_wait_locked(h, slot);
}

static int host_submit(b200mel_handle* h, const void* pcm_host, int32_t pcm_dtype, const int64_t* offsets_host,
                       const int32_t* lengths_host, int32_t batch, int64_t row_stride, int64_t total_samples,
                       float* out_host, int64_t t_max, int32_t* n_frames_host, uint64_t seed,
                       uint64_t first_utterance_id, int32_t kernel, int32_t slot, int default_chunks) {
  if (!h) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null handle");
  if (slot < 0 || slot > 1) return fail(B200MEL_ERR_INVALID_ARGUMENT, "slot must be 0 or 1");
  if (batch < 0 || t_max < 0 || total_samples < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative size");
  if (batch == 0) return B200MEL_OK;
  if (!pcm_host || !lengths_host || (t_max > 0 && !out_host)) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  if (pcm_dtype != B200MEL_PCM_I16 && pcm_dtype != B200MEL_PCM_F32)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "pcm_dtype must be B200MEL_PCM_I16 or B200MEL_PCM_F32");
  const size_t esz = pcm_dtype == B200MEL_PCM_F32 ? 4 : 2;
  if (!offsets_host && total_samples < static_cast<int64_t>(batch) * row_stride)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "total_samples < batch * row_stride");
  std::lock_guard<std::mutex> lock(h->host_mu);
  DeviceGuard guard(h->device);
  if (!guard.ok) return fail(B200MEL_ERR_CUDA, "cudaSetDevice failed");
  for (auto& s : h->streams)
    if (!s) CU_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  if (!h->ev_init) {
    for (auto& sl : h->slots) {
      for (auto& ev : sl.ev_in) CU_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
      CU_CHECK(cudaEventCreateWithFlags(&sl.ev_done_k, cudaEventDisableTiming));
      CU_CHECK(cudaEventCreateWithFlags(&sl.ev_done_out, cudaEventDisableTiming));
    }
    h->ev_init = true;
  }
  // the slot's previous step must have left its staging buffers (the caller has normally waited for it)
  {
    const int rc = host_wait_locked(h, slot);
    if (rc != B200MEL_OK) return rc;
  }
  b200mel_handle::HostSlot& sl = h->slots[slot];
  const HostPlan& p = h->plan;
  const size_t pcm_bytes = static_cast<size_t>(total_samples) * esz;
  const size_t per_utt_out = static_cast<size_t>(t_max) * p.F * p.C;
  const size_t out_bytes = per_utt_out * batch * sizeof(float);
  if (pcm_bytes > sl.s_pcm_bytes) {
    cudaFree(sl.s_pcm);
    sl.s_pcm = nullptr;
    sl.s_pcm_bytes = 0;
    CU_CHECK(cudaMalloc(&sl.s_pcm, pcm_bytes + 256));
    sl.s_pcm_bytes = pcm_bytes;
  }
  if (out_bytes > sl.s_out_bytes) {
    cudaFree(sl.s_out);
    sl.s_out = nullptr;
    sl.s_out_bytes = 0;
    CU_CHECK(cudaMalloc(reinterpret_cast<void**>(&sl.s_out), out_bytes + 256));
    sl.s_out_bytes = out_bytes;
  }
  if (static_cast<size_t>(batch) > sl.s_small_cap) {
    cudaFree(sl.s_small);
    cudaFree(sl.s_offsets);
    cudaFreeHost(sl.p_small);
    cudaFreeHost(sl.p_offsets);
    sl.s_small = nullptr;
    sl.s_offsets = nullptr;
    sl.p_small = nullptr;
    sl.p_offsets = nullptr;
    sl.s_small_cap = 0;
    CU_CHECK(cudaMallocHost(reinterpret_cast<void**>(&sl.p_small), (2 * static_cast<size_t>(batch) + 4) * sizeof(int32_t)));
    CU_CHECK(cudaMallocHost(reinterpret_cast<void**>(&sl.p_offsets), static_cast<size_t>(batch) * sizeof(int64_t)));
    CU_CHECK(cudaMalloc(reinterpret_cast<void**>(&sl.s_small), (2 * static_cast<size_t>(batch) + 4) * sizeof(int32_t)));
    CU_CHECK(cudaMalloc(reinterpret_cast<void**>(&sl.s_offsets), static_cast<size_t>(batch) * sizeof(int64_t)));
    sl.s_small_cap = batch;
  }
  int32_t* d_len = sl.s_small;
  int32_t* d_nf = sl.s_small + batch;
  int32_t* d_status = sl.s_small + 2 * static_cast<size_t>(batch);
  cudaStream_t s_in = h->streams[0], s_k = h->streams[1], s_out = h->streams[2];
  std::memcpy(sl.p_small, lengths_host, batch * sizeof(int32_t));
  sl.p_small[2 * static_cast<size_t>(batch)] = 0;
  CU_CHECK(cudaMemsetAsync(d_status, 0, sizeof(int32_t), s_in));  // (every kernel of the step waits for s_in)
  CU_CHECK(cudaMemcpyAsync(d_len, sl.p_small, batch * sizeof(int32_t), cudaMemcpyHostToDevice, s_in));
  if (offsets_host) {
    std::memcpy(sl.p_offsets, offsets_host, batch * sizeof(int64_t));
    CU_CHECK(cudaMemcpyAsync(sl.s_offsets, sl.p_offsets, batch * sizeof(int64_t), cudaMemcpyHostToDevice, s_in));
  }

  // Chunked 3-stage pipeline: H2D(chunk i+1) | kernel(chunk i) | D2H(chunk i-1) on three streams,
  // so PCIe traffic in both directions overlaps the kernel.
  // Chunks per step, measured on B200 (123 MB each way): a lone synchronous call is fastest with 8
  // (4: 1.20 M, 8: 1.28 M, 16: 1.23 M, 32: 1.11 M audio-s/s - the first upload and the last download cannot
  // overlap anything); back-to-back submits overlap across steps and are fastest with 2 (1: 1.43 M, 2: 1.47 M,
  // 4: 1.45 M, 8: 1.40 M).  The link carries ~49 GB/s per direction when both directions run.
  // B200MEL_HOST_CHUNKS overrides (developer knob for the e2e measurement).
  static const int forced_chunks = [] {
    const char* env = getenv("B200MEL_HOST_CHUNKS");
    return env ? atoi(env) : 0;
  }();
  int n_chunks = forced_chunks > 0 ? forced_chunks : default_chunks;
  if (n_chunks > batch) n_chunks = batch;
  if (n_chunks < 1) n_chunks = 1;
  if (n_chunks > 32) n_chunks = 32;
  const char* src = static_cast<const char*>(pcm_host);
  char* dst = static_cast<char*>(sl.s_pcm);
  for (int c = 0; c < n_chunks; ++c) {
    const int b0 = static_cast<int>(static_cast<int64_t>(batch) * c / n_chunks);
    const int b1 = static_cast<int>(static_cast<int64_t>(batch) * (c + 1) / n_chunks);
    if (b1 <= b0) continue;
    int64_t lo, hi;
    if (offsets_host) {
      lo = INT64_MAX;
      hi = 0;
      for (int b = b0; b < b1; ++b) {
        if (offsets_host[b] < 0 || offsets_host[b] + lengths_host[b] > total_samples)
          return fail(B200MEL_ERR_INVALID_ARGUMENT, "offsets/lengths outside total_samples");
        lo = offsets_host[b] < lo ? offsets_host[b] : lo;
        int64_t end = offsets_host[b] + lengths_host[b];
        hi = end > hi ? end : hi;
      }
      if (hi < lo) hi = lo;
    } else {
      lo = static_cast<int64_t>(b0) * row_stride;
      hi = static_cast<int64_t>(b1) * row_stride;
    }
    if (hi > lo)
      CU_CHECK(cudaMemcpyAsync(dst + lo * esz, src + lo * esz, static_cast<size_t>(hi - lo) * esz,
                               cudaMemcpyHostToDevice, s_in));
    CU_CHECK(cudaEventRecord(sl.ev_in[2 * c], s_in));
    CU_CHECK(cudaStreamWaitEvent(s_k, sl.ev_in[2 * c], 0));
    const int64_t* d_off = offsets_host ? sl.s_offsets + b0 : nullptr;
    const char* d_pcm = offsets_host ? dst : dst + static_cast<int64_t>(b0) * row_stride * esz;
    int rc = b200mel_forward(h, d_pcm, pcm_dtype, d_off, d_len + b0, b1 - b0, row_stride,
                             sl.s_out + per_utt_out * b0, t_max, d_nf + b0, nullptr, seed,
                             first_utterance_id + static_cast<uint64_t>(b0), d_status, kernel, s_k);
    if (rc != B200MEL_OK) return rc;
    CU_CHECK(cudaEventRecord(sl.ev_in[2 * c + 1], s_k));
    CU_CHECK(cudaStreamWaitEvent(s_out, sl.ev_in[2 * c + 1], 0));
    if (per_utt_out > 0)
      CU_CHECK(cudaMemcpyAsync(out_host + per_utt_out * b0, sl.s_out + per_utt_out * b0,
                               per_utt_out * (b1 - b0) * sizeof(float), cudaMemcpyDeviceToHost, s_out));
  }
  // frame counts and the status word behind them, in one copy
  CU_CHECK(cudaMemcpyAsync(sl.p_small + batch, d_nf, (static_cast<size_t>(batch) + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, s_k));
  sl.user_n_frames = n_frames_host;
  sl.user_batch = batch;
  CU_CHECK(cudaEventRecord(sl.ev_done_k, s_k));
  CU_CHECK(cudaEventRecord(sl.ev_done_out, s_out));
  sl.pending = true;
  return B200MEL_OK;
}

int b200mel_forward_host_submit(b200mel_handle* h, const void* pcm_host, int32_t pcm_dtype, const int64_t* offsets_host,
                                const int32_t* lengths_host, int32_t batch, int64_t row_stride, int64_t total_samples,
                                float* out_host, int64_t t_max, int32_t* n_frames_host, uint64_t seed,
                                uint64_t first_utterance_id, int32_t kernel, int32_t slot) {
  return host_submit(h, pcm_host, pcm_dtype, offsets_host, lengths_host, batch, row_stride, total_samples, out_host,
                     t_max, n_frames_host, seed, first_utterance_id, kernel, slot, 2);
}

int b200mel_forward_host(b200mel_handle* h, const void* pcm_host, int32_t pcm_dtype, const int64_t* offsets_host,
                         const int32_t* lengths_host, int32_t batch, int64_t row_stride, int64_t total_samples,
                         float* out_host, int64_t t_max, int32_t* n_frames_host, uint64_t seed,
                         uint64_t first_utterance_id, int32_t kernel) {
  const int rc = host_submit(h, pcm_host, pcm_dtype, offsets_host, lengths_host, batch, row_stride, total_samples,
                             out_host, t_max, n_frames_host, seed, first_utterance_id, kernel, 0, 8);
  if (rc != B200MEL_OK) return rc;
  return b200mel_forward_host_wait(h, 0);
}

int b200mel_specaug_draw(const int32_t* num_time, int32_t batch, int32_t v, int32_t F, int32_t m_F, int32_t T,
                         float p, int32_t m_T, uint64_t seed, uint64_t first_utterance_id, int32_t* table,
                         int32_t* status, void* stream) {
  if (batch < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative batch");
  // enable flags of spec_augment (data.py:267-269)
  int nf = (F != 0 && m_F != 0) ? m_F : 0;
  int nt = (T != 0 && p != 0.0f && m_T != 0) ? m_T : 0;
  if (nf < 0 || nt < 0 || F < 0 || T < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative SpecAugment parameter");
  if (nf + nt > kMaxMasks) return fail(B200MEL_ERR_UNSUPPORTED, "more than 16 SpecAugment masks");
  if (batch == 0 || nf + nt == 0) return B200MEL_OK;
  if (!num_time || !table) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  specaug_draw_kernel<<<(batch + 127) / 128, 128, 0, st>>>(num_time, batch, v, F, nf, T, p, nt, seed,
                                                          first_utterance_id, table, status);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "specaug_draw_kernel launch");
  return B200MEL_OK;
}

int b200mel_specaug_apply(const float* in, float* out, int32_t batch, int64_t t_max, int32_t v, int32_t channels,
                          const int32_t* table, int32_t m_F, int32_t m_T, void* stream) {
  if (batch < 0 || t_max < 0 || v < 0 || channels < 0 || m_F < 0 || m_T < 0)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative size");
  if (m_F + m_T > kMaxMasks) return fail(B200MEL_ERR_UNSUPPORTED, "more than 16 SpecAugment masks");
  if (batch > 65535) return fail(B200MEL_ERR_UNSUPPORTED, "batch > 65535: split the call");
  int64_t per = t_max * v * channels;
  if (batch == 0 || per == 0) return B200MEL_OK;
  if (!in || !out || (m_F + m_T > 0 && !table)) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int64_t blocks = (per + 255) / 256;
  if (blocks > 1184) blocks = 1184;
  specaug_apply_kernel<<<dim3(static_cast<unsigned>(blocks), batch), 256, 0, st>>>(in, out, t_max, v, channels, table,
                                                                                  m_F, m_T);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "specaug_apply_kernel launch");
  return B200MEL_OK;
}

int b200mel_delta(const float* in, float* out, const int32_t* n_frames, int32_t batch, int64_t t_max, int32_t v,
                  void* stream) {
  if (batch < 0 || t_max < 0 || v < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative size");
  if (batch > 65535) return fail(B200MEL_ERR_UNSUPPORTED, "batch > 65535: split the call");
  int64_t per = t_max * v;
  if (batch == 0 || per == 0) return B200MEL_OK;
  if (!in || !out) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  if (in == out) return fail(B200MEL_ERR_INVALID_ARGUMENT, "delta cannot run in place");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int64_t blocks = (per * 3 + 255) / 256;
  if (blocks > 1184) blocks = 1184;
  delta_kernel<<<dim3(static_cast<unsigned>(blocks), batch), 256, 0, st>>>(in, out, n_frames, t_max, v);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "delta_kernel launch");
  return B200MEL_OK;
}

int b200mel_time_warp(const float* in, float* out, const int32_t* n_frames, int32_t batch, int64_t t_max, int32_t v,
                      int32_t channels, int32_t W, const int32_t* warp_table, uint64_t seed, uint64_t first_utterance_id,
                      int32_t* status, void* stream) {
  if (batch < 0 || t_max < 0 || v < 0 || channels < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative size");
  if (batch > 65535) return fail(B200MEL_ERR_UNSUPPORTED, "batch > 65535: split the call");
  if (W <= 0 && !warp_table) return fail(B200MEL_ERR_INVALID_ARGUMENT, "time warp needs W > 0 or a warp_table");
  if (t_max > (int64_t{1} << 24)) return fail(B200MEL_ERR_UNSUPPORTED, "t_max too large for the float32 pixel grid");
  const int64_t pixels = t_max * v;
  if (batch == 0 || pixels == 0 || channels == 0) return B200MEL_OK;
  if (!in || !out || !n_frames) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  if (in == out) return fail(B200MEL_ERR_INVALID_ARGUMENT, "the time warp cannot run in place");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  AsyncScratch coef;   // [batch][kWarpCoefFloats], stream-ordered
  coef.st = st;
  CU_CHECK(cudaMallocAsync(&coef.p, static_cast<size_t>(batch) * kWarpCoefFloats * sizeof(float), st));
  time_warp_solve_kernel<<<(batch + 63) / 64, 64, 0, st>>>(n_frames, batch, t_max, v, W, warp_table, seed, first_utterance_id,
                                                          static_cast<float*>(coef.p), status);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "time_warp_solve_kernel launch");
  const int64_t blocks = (pixels + 255) / 256;
  if (blocks > 0x7fffffffLL) return fail(B200MEL_ERR_UNSUPPORTED, "t_max * v too large");
  time_warp_apply_kernel<<<dim3(static_cast<unsigned>(blocks), static_cast<unsigned>(batch)), 256, 0, st>>>(
      in, out, static_cast<const float*>(coef.p), t_max, v, channels);
  e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "time_warp_apply_kernel launch");
  return B200MEL_OK;
}

int b200mel_wav_parse(const void* file_bytes, int64_t n_bytes, int32_t* format, int32_t* channels,
                      int32_t* sample_rate, int32_t* bits_per_sample, int64_t* data_offset, int64_t* data_bytes) {
  if (!file_bytes || n_bytes < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null or negative-size file image");
  WavInfo w;
  int rc = wav_parse(static_cast<const uint8_t*>(file_bytes), n_bytes, &w);
  if (rc != 0) {
    static const char* why[] = {"", "not a RIFF/WAVE image", "data chunk before fmt chunk", "no data chunk",
                                "truncated fmt chunk"};
    return fail(B200MEL_ERR_INVALID_ARGUMENT, std::string("wav: ") + why[rc]);
  }
  if (format) *format = w.format;
  if (channels) *channels = w.channels;
  if (sample_rate) *sample_rate = w.sample_rate;
  if (bits_per_sample) *bits_per_sample = w.bits;
  if (data_offset) *data_offset = w.data_offset;
  if (data_bytes) *data_bytes = w.data_bytes;
  return B200MEL_OK;
}

int b200mel_ingest_pcm16(const void* bytes, const int64_t* byte_offsets, const int64_t* byte_lengths,
                         const int32_t* channels, int32_t batch, void* pcm_out, int32_t pcm_dtype,
                         const int64_t* out_offsets, int64_t row_stride, int32_t* lengths_out, void* stream) {
  if (batch < 0 || row_stride < 0) return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative size");
  if (pcm_dtype != B200MEL_PCM_I16 && pcm_dtype != B200MEL_PCM_F32)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "pcm_dtype must be B200MEL_PCM_I16 or B200MEL_PCM_F32");
  if (channels && pcm_dtype != B200MEL_PCM_F32)
    return fail(B200MEL_ERR_INVALID_ARGUMENT,
                "multi-channel payloads are channel means (data.py:116) and need float32 output");
  if (batch > 65535) return fail(B200MEL_ERR_UNSUPPORTED, "batch > 65535: split the call");
  if (batch == 0) return B200MEL_OK;
  if (!bytes || !byte_offsets || !byte_lengths || !pcm_out || !lengths_out)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // at most one resident wave (148 SMs x 8 blocks of 256 threads), split over the utterances
  unsigned bx = static_cast<unsigned>(std::max(1, 1184 / batch));
  dim3 grid(bx, static_cast<unsigned>(batch));
  const uint8_t* src = static_cast<const uint8_t*>(bytes);
  if (channels) {
    ingest_multi_kernel<<<grid, 256, 0, st>>>(src, byte_offsets, byte_lengths, channels,
                                              static_cast<float*>(pcm_out), out_offsets, row_stride, lengths_out);
  } else if (pcm_dtype == B200MEL_PCM_I16) {
    ingest_mono_kernel<int16_t><<<grid, 256, 0, st>>>(src, byte_offsets, byte_lengths,
                                                      static_cast<int16_t*>(pcm_out), out_offsets, row_stride,
                                                      lengths_out);
  } else {
    ingest_mono_kernel<float><<<grid, 256, 0, st>>>(src, byte_offsets, byte_lengths, static_cast<float*>(pcm_out),
                                                    out_offsets, row_stride, lengths_out);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "ingest kernel launch");
  return B200MEL_OK;
}

uint32_t b200mel_tfrecord_masked_crc(const void* data, int64_t n_bytes) {
  if (!data || n_bytes <= 0) return tfrecord_masked_crc(nullptr, 0);
  return tfrecord_masked_crc(static_cast<const uint8_t*>(data), n_bytes);
}

int b200mel_audio_mask_length(int64_t t_max, const int32_t* time_kernel_sizes, const int32_t* time_strides,
                              int32_t n_layers, int64_t* reduced_out, int32_t* group_out) {
  if (t_max < 0 || n_layers < 0 || (n_layers > 0 && (!time_kernel_sizes || !time_strides)))
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "bad audio-mask layer list");
  int64_t t = t_max, g = 1;
  for (int i = 0; i < n_layers; ++i) {
    int64_t k = time_kernel_sizes[i], s = time_strides[i];
    if (k < 1 || s < 1) return fail(B200MEL_ERR_INVALID_ARGUMENT, "kernel sizes and strides must be >= 1");
    t -= k - s;
    // Python floor division (models/las.py:210-213): negative lengths floor towards -inf, then clamp
    t = t >= 0 ? t / s : -((-t + s - 1) / s);
    g *= s;
    if (g > (1 << 20)) return fail(B200MEL_ERR_UNSUPPORTED, "stride product too large");
  }
  if (t < 0) t = 0;  // a slice [:negative * g] of the frame axis: clamp (an empty mask)
  if (reduced_out) *reduced_out = t;
  if (group_out) *group_out = static_cast<int32_t>(g);
  return B200MEL_OK;
}

int b200mel_audio_mask(const float* feat, int32_t batch, int64_t t_max, int64_t row_elems, int32_t group,
                       int64_t reduced, uint8_t* mask_out, void* stream) {
  if (batch < 0 || t_max < 0 || row_elems < 0 || group < 1 || reduced < 0)
    return fail(B200MEL_ERR_INVALID_ARGUMENT, "negative size");
  if (reduced * group > t_max) return fail(B200MEL_ERR_INVALID_ARGUMENT, "reduced * group exceeds t_max");
  if (batch == 0 || reduced == 0) return B200MEL_OK;
  if (!feat || !mask_out) return fail(B200MEL_ERR_INVALID_ARGUMENT, "null buffer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // one warp per pair of entries, at most one resident wave (148 SMs x 8 blocks of 8 warps)
  const int64_t want = (static_cast<int64_t>(batch) * ((reduced + 1) / 2) + 7) / 8;
  const unsigned bx = static_cast<unsigned>(std::min<int64_t>(want, 1184));
  audio_mask_kernel<<<bx, 256, 0, st>>>(feat, batch, t_max, row_elems, group, reduced, mask_out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "audio_mask_kernel launch");
  return B200MEL_OK;
}

}  // extern "C"
