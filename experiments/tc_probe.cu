// tc_probe.cu - tcgen05 feasibility probes for the tensor-core DFT kernel (not part of the product).
//
//  1. correctness of hand-built UMMA descriptors: control (A rows 128 B apart, SWIZZLE_128B) and
//     the "Hankel" view (A rows 64 B apart in ONE contiguous sample stream, SWIZZLE_64B, so that
//     consecutive frames overlap by half a row without being materialised twice);
//  2. throughput of M=128 x N x K=16 SS-mode MMAs for small N (operand smem bandwidth), alone and
//     with concurrent LSU shared-memory traffic.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tc_probe tc_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);  \
      exit(1);                                                                         \
    }                                                                                  \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// bounded wait: returns false on timeout instead of hanging the GPU
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity, uint32_t max_polls = (1u << 24)) {
  for (uint32_t i = 0; i < max_polls; ++i)
    if (mbar_try_wait(bar, parity)) return true;
  return false;
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout,
                                              uint32_t base_offset) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;  // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(base_offset & 7) << 49;
  d |= static_cast<uint64_t>(layout & 7) << 61;
  return d;
}
constexpr uint32_t LAYOUT_NONE = 0, LAYOUT_SW128 = 2, LAYOUT_SW64 = 4, LAYOUT_SW32 = 6;

__device__ __forceinline__ uint32_t make_idesc_f16(int M, int N) {
  return (1u << 4) | (0u << 7) | (0u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}

__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// --------------------------------------------------------------------------------------------
// correctness probe.  variant 0: control (explicit rows, SW128).  variant 1: Hankel SW64,
// base_offset 0.  variant 2: Hankel SW64, base_offset = (start >> 7) & 7.  variant 3: Hankel,
// no swizzle is impossible (row pitch 16 B) - skipped.  variant 4: Hankel SW64 with the tile
// starting one row (64 B) into the stream (tile base not 512-aligned), base_offset 0.
// D[m][n] = sum_k A[m][k] * Bt[n][k], A[m][k] = stream[32*(m + row0) + k], M=128, N=64, K=64.
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(160) probe_kernel(const __half* stream, const __half* Bt, float* D, int variant,
                                                    int* err) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                 // 32 KB region
  uint8_t* sB = smem + 32768;         // 8 KB: 64 rows x 128 B, SW128
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int row0 = variant == 4 ? 1 : 0;

  if (variant == 0) {
    // explicit rows: row m at 128*m bytes, SW128
    for (int i = tid; i < 128 * 64; i += blockDim.x) {
      int m = i / 64, k = i % 64;
      uint32_t o = 128 * m + 2 * k;
      uint32_t ph = o ^ (((o >> 7) & 7) << 4);
      *reinterpret_cast<__half*>(sA + ph) = stream[32 * m + k];
    }
  } else {
    // one contiguous stream, SW64 applied to the byte offset inside the 1024-aligned region
    for (int e = tid; e < 32 * 130 + 64; e += blockDim.x) {
      uint32_t o = 2 * e;
      uint32_t ph = o ^ (((o >> 7) & 3) << 4);
      *reinterpret_cast<__half*>(sA + ph) = stream[e];
    }
  }
  for (int i = tid; i < 64 * 64; i += blockDim.x) {
    int n = i / 64, k = i % 64;
    uint32_t o = 128 * n + 2 * k;
    uint32_t ph = o ^ (((o >> 7) & 7) << 4);
    *reinterpret_cast<__half*>(sB + ph) = Bt[n * 64 + k];
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  fence_async_smem();
  __syncthreads();
  if (warp == 4) {
    tmem_alloc(&tmem_base, 64);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;

  if (warp == 4 && (tid & 31) == 0) {
    const uint32_t idesc = make_idesc_f16(128, 64);
    const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
    for (int j = 0; j < 4; ++j) {
      uint64_t ad, bd;
      if (variant == 0) {
        ad = make_desc(a0 + 32 * j, 16, 1024, LAYOUT_SW128, 0);
      } else {
        uint32_t start = a0 + 64 * row0 + (j < 2 ? 32 * j : 64 + 32 * (j - 2));
        uint32_t bo = variant == 2 ? ((start >> 7) & 7) : 0;
        ad = make_desc(start, 16, 512, LAYOUT_SW64, bo);
      }
      bd = make_desc(b0 + 32 * j, 16, 1024, LAYOUT_SW128, 0);
      umma_f16(tbase, ad, bd, idesc, j > 0 ? 1u : 0u);
    }
    umma_commit(&bar);
  }
  if (warp < 4) {
    bool ok = mbar_wait(&bar, 0);
    if (!ok && (tid & 31) == 0) atomicExch(err, 1);
    tc_fence_after();
    if (ok) {
      const int m = tid;  // lane of TMEM == row
      uint32_t v[16];
      for (int c = 0; c < 4; ++c) {
        tmem_ld16(tbase + (static_cast<uint32_t>(warp * 32) << 16) + 16 * c, v);
        tmem_ld_wait();
        for (int i = 0; i < 16; ++i) D[m * 64 + 16 * c + i] = __uint_as_float(v[i]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tbase, 64);
}

// --------------------------------------------------------------------------------------------
// throughput probe: one CTA per SM; warp 4 issues `iters` batches of `per_batch` MMAs (M=128,
// N=n_dim, K=16, SS mode, A SW64 pitch 64 B, B SW128) and waits for each batch's commit; warps
// 0..3 optionally hammer shared memory through the LSU meanwhile.
// --------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(160) tput_kernel(int n_dim, int iters, int per_batch, int lsu_traffic, long long* cycles,
                                                   int* err, float* sink) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;           // 80 KB of "streams"
  uint8_t* sB = smem + 81920;   // 64 KB of B
  uint8_t* sL = smem + 81920 + 65536;  // 32 KB scratch for LSU traffic
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  __shared__ volatile int stop;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (81920 + 65536 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    stop = 0;
  }
  fence_async_smem();
  __syncthreads();
  if (warp == 4) tmem_alloc(&tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;
  if (warp == 4) {
    if ((tid & 31) == 0) {
      const uint32_t idesc = make_idesc_f16(128, n_dim);
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
      uint32_t parity = 0;
      long long t0 = clock64();
      bool ok = true;
      for (int it = 0; it < iters && ok; ++it) {
        for (int j = 0; j < per_batch; ++j) {
          uint32_t astart = a0 + ((j * 8192 + 32 * (j & 1)) % 65536);
          uint32_t bstart = b0 + ((j * 2048 + 32 * (j & 3)) % 32768);
          uint64_t ad = make_desc(astart, 16, 512, LAYOUT_SW64, 0);
          uint64_t bd = make_desc(bstart, 16, 1024, LAYOUT_SW128, 0);
          umma_f16(tbase + (j % 2) * 256, ad, bd, idesc, 1u);
        }
        umma_commit(&bar);
        ok = mbar_wait(&bar, parity);
        parity ^= 1;
      }
      long long t1 = clock64();
      if (!ok) atomicExch(err, 2);
      cycles[blockIdx.x] = t1 - t0;
      stop = 1;
    }
  } else if (lsu_traffic) {
    // 128-bit conflict-free loads + stores in a loop until the MMA warp is done
    float4* p = reinterpret_cast<float4*>(sL);
    float4 acc = make_float4(0, 0, 0, 0);
    int guard = 0;
    while (!stop && guard < (1 << 22)) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float4 v = p[(tid + 128 * u) & 2047];
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      if (lsu_traffic > 1) p[tid] = acc;
      ++guard;
    }
    if (acc.x == 12345.f) sink[tid] = acc.y;
    if (tid == 0) cycles[gridDim.x + blockIdx.x] = guard;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tbase, 512);
}


// --------------------------------------------------------------------------------------------
// throughput probe v2: descriptors precomputed, MMAs fully unrolled (the issuing thread must not be
// the bottleneck).  One "tile" = 5 residues x PASSES x 4 K-steps MMAs (M=128, N=64, K=16) into 5
// accumulators, then one commit + wait - the real kernel's pattern.
// --------------------------------------------------------------------------------------------
template <int PASSES, int NDIM>
__global__ void __launch_bounds__(160) tput2_kernel(int tiles, int lsu_traffic, long long* cycles, int* err, float* sink) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                   // 10 streams x 8448 B (hi/lo x 5 residues) = 84480 -> 86016
  uint8_t* sB = smem + 86016;           // 10 x 8 KB (NDIM<=64) or 10 x 16 KB
  uint8_t* sL = smem + 86016 + 10 * 128 * NDIM;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base;
  __shared__ volatile int stop;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int total_words = (86016 + 10 * 128 * NDIM + 16384) / 4;
  for (int i = tid; i < total_words; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    stop = 0;
  }
  fence_async_smem();
  __syncthreads();
  if (warp == 4) tmem_alloc(&tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_base;
  if (warp == 4) {
    if ((tid & 31) == 0) {
      const uint32_t idesc = make_idesc_f16(128, NDIM);
      const uint64_t a_base = make_desc(smem_u32(sA), 16, 512, LAYOUT_SW64, 0);
      const uint64_t b_base = make_desc(smem_u32(sB), 16, 1024, LAYOUT_SW128, 0);
      uint32_t parity = 0;
      bool ok = true;
      long long t0 = clock64();
      for (int t = 0; t < tiles && ok; ++t) {
#pragma unroll
        for (int r = 0; r < 5; ++r) {
#pragma unroll
          for (int ps = 0; ps < PASSES; ++ps) {
            // pass 0: hi x Bhi, pass 1: hi x Blo, pass 2: lo x Bhi, pass 3: lo x Blo
            const int a_term = ps >> 1, b_term = ps & 1;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t a_off = (a_term * 5 + r) * 8448 + (j < 2 ? 32 * j : 64 + 32 * (j - 2));
              const uint32_t b_off = (b_term * 5 + r) * 128 * NDIM + 32 * j;
              umma_f16(tbase + r * NDIM, a_base + (a_off >> 4), b_base + (b_off >> 4), idesc, (ps | j) ? 1u : 0u);
            }
          }
        }
        umma_commit(&bar);
        ok = mbar_wait(&bar, parity);
        parity ^= 1;
      }
      long long t1 = clock64();
      if (!ok) atomicExch(err, 2);
      cycles[blockIdx.x] = t1 - t0;
      stop = 1;
    }
  } else if (lsu_traffic) {
    float4* p = reinterpret_cast<float4*>(sL);
    float4 acc = make_float4(0, 0, 0, 0);
    int guard = 0;
    while (!stop && guard < (1 << 22)) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float4 v = p[(tid + 128 * u) & 1023];
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
      if (lsu_traffic > 1) {
#pragma unroll
        for (int u = 0; u < 8; ++u) p[(tid + 128 * u) & 1023] = acc;
      }
      ++guard;
    }
    if (acc.x == 12345.f) sink[tid] = acc.y;
    if (tid == 0) cycles[gridDim.x + blockIdx.x] = guard;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tbase, 512);
}

template <int PASSES, int NDIM>
void run_tput2(int G, long long* d_cyc, int* d_err, float* d_sink) {
  const int smem_bytes = 86016 + 10 * 128 * NDIM + 16384 + 1024;
  CK(cudaFuncSetAttribute(tput2_kernel<PASSES, NDIM>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  for (int lsu : {0, 1, 2}) {
    const int tiles = 100;
    CK(cudaMemset(d_err, 0, 4));
    CK(cudaMemset(d_cyc, 0, sizeof(long long) * 2 * G));
    tput2_kernel<PASSES, NDIM><<<G, 160, smem_bytes>>>(tiles, lsu, d_cyc, d_err, d_sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("tput2: CUDA error %s\n", cudaGetErrorString(e)); exit(3); }
    std::vector<long long> cyc(2 * G);
    int herr = 0;
    CK(cudaMemcpy(cyc.data(), d_cyc, sizeof(long long) * 2 * G, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost));
    double mean = 0;
    for (int i = 0; i < G; ++i) mean += cyc[i];
    mean /= G;
    const int n_mma = 5 * PASSES * 4;
    double per_tile = mean / tiles;
    double lsu_b = lsu ? (double)cyc[G] * 8 * 128 * 16 * (lsu > 1 ? 2.0 : 1.0) / (double)cyc[0] : 0.0;
    printf("tput2 passes=%d N=%3d lsu=%d: err=%d  %.0f cycles/tile (%d MMAs) = %.1f cycles/MMA (ideal %.1f), operand %.0f B/clk, lsu %.0f B/clk\n",
           PASSES, NDIM, lsu, herr, per_tile, n_mma, per_tile / n_mma, 128.0 * NDIM / 256.0,
           n_mma * (4096.0 + 32.0 * NDIM) / per_tile, lsu_b);
  }
}

int main() {
  int dev = 0;
  CK(cudaSetDevice(dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  printf("device %s sm_%d%d SMs %d smem optin %zu\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount,
         prop.sharedMemPerBlockOptin);
  // ---- correctness -----------------------------------------------------------------------
  const int NE = 32 * 130 + 64;
  std::vector<__half> h_stream(NE), h_B(64 * 64);
  srand(1);
  for (int i = 0; i < NE; ++i) h_stream[i] = __float2half(static_cast<float>((rand() % 2001) - 1000) / 8.0f);
  for (int i = 0; i < 64 * 64; ++i) h_B[i] = __float2half(static_cast<float>((rand() % 2001) - 1000) / 1024.0f);
  __half *d_stream, *d_B;
  float* d_D;
  int* d_err;
  CK(cudaMalloc(&d_stream, NE * 2));
  CK(cudaMalloc(&d_B, 64 * 64 * 2));
  CK(cudaMalloc(&d_D, 128 * 64 * 4));
  CK(cudaMalloc(&d_err, 4));
  CK(cudaMemcpy(d_stream, h_stream.data(), NE * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_B, h_B.data(), 64 * 64 * 2, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024));
  for (int variant : {0, 1, 2, 4}) {
    CK(cudaMemset(d_D, 0xFF, 128 * 64 * 4));
    CK(cudaMemset(d_err, 0, 4));
    probe_kernel<<<1, 160, 48 * 1024>>>(d_stream, d_B, d_D, variant, d_err);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("variant %d: CUDA error %s\n", variant, cudaGetErrorString(e));
      return 2;
    }
    std::vector<float> D(128 * 64);
    int herr = 0;
    CK(cudaMemcpy(D.data(), d_D, 128 * 64 * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost));
    const int row0 = variant == 4 ? 1 : 0;
    double maxerr = 0, maxref = 0;
    int bad = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < 64; ++n) {
        double ref = 0;
        for (int k = 0; k < 64; ++k)
          ref += static_cast<double>(__half2float(h_stream[32 * (m + row0) + k])) * static_cast<double>(__half2float(h_B[n * 64 + k]));
        double d = fabs(ref - D[m * 64 + n]);
        if (!(d <= 1e-2 * (1 + fabs(ref)))) ++bad;
        if (d > maxerr || d != d) maxerr = d;
        if (fabs(ref) > maxref) maxref = fabs(ref);
      }
    printf("variant %d: timeout=%d max_abs_err=%.4g (max |ref| %.4g) bad=%d/8192  D[0][0..3]=%g %g %g %g  D[1][0]=%g D[9][0]=%g\n",
           variant, herr, maxerr, maxref, bad, D[0], D[1], D[2], D[3], D[64], D[9 * 64]);
  }
  // ---- throughput ------------------------------------------------------------------------
  const int smem_bytes = 81920 + 65536 + 32768 + 1024;
  CK(cudaFuncSetAttribute(tput_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  long long* d_cyc;
  float* d_sink;
  const int G = prop.multiProcessorCount;
  CK(cudaMalloc(&d_cyc, sizeof(long long) * 2 * G));
  CK(cudaMalloc(&d_sink, 4 * 256));
  for (int lsu : {0, 1, 2})
    for (int n : {16, 32, 64, 128, 256}) {
      const int iters = 200, per_batch = 20;
      CK(cudaMemset(d_err, 0, 4));
      CK(cudaMemset(d_cyc, 0, sizeof(long long) * 2 * G));
      cudaEvent_t e0, e1;
      CK(cudaEventCreate(&e0));
      CK(cudaEventCreate(&e1));
      CK(cudaEventRecord(e0));
      tput_kernel<<<G, 160, smem_bytes>>>(n, iters, per_batch, lsu, d_cyc, d_err, d_sink);
      CK(cudaEventRecord(e1));
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) {
        printf("tput n=%d lsu=%d: CUDA error %s\n", n, lsu, cudaGetErrorString(e));
        return 3;
      }
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      std::vector<long long> cyc(2 * G);
      int herr = 0;
      CK(cudaMemcpy(cyc.data(), d_cyc, sizeof(long long) * 2 * G, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(&herr, d_err, 4, cudaMemcpyDeviceToHost));
      double mean = 0;
      long long mx = 0;
      for (int i = 0; i < G; ++i) { mean += cyc[i]; if (cyc[i] > mx) mx = cyc[i]; }
      mean /= G;
      double per_mma = mean / (iters * per_batch);
      double lsu_bytes_per_cycle = lsu ? (double)cyc[G] * 8 * 128 * 16 * (lsu > 1 ? 1.125 : 1.0) / (double)cyc[0] : 0.0;
      printf("tput N=%3d lsu=%d: err=%d  %.1f cycles/MMA (max CTA %.1f)  ideal %.1f  operand B/clk %.0f  lsu B/clk %.0f  kernel %.3f ms\n", n,
             lsu, herr, per_mma, (double)mx / (iters * per_batch), 128.0 * n / 256.0,
             (4096.0 + 32.0 * n) / per_mma, lsu_bytes_per_cycle, ms);
    }
  run_tput2<3, 64>(G, d_cyc, d_err, d_sink);
  run_tput2<4, 64>(G, d_cyc, d_err, d_sink);
  run_tput2<3, 32>(G, d_cyc, d_err, d_sink);
  run_tput2<3, 128>(G, d_cyc, d_err, d_sink);
  printf("done\n");
  return 0;
}
