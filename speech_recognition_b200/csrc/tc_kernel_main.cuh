// tc_kernel_main.cuh - the kernel body (roles, pipeline, epilogue).  See tc_kernel.cuh.
#pragma once

namespace b200mel {

struct TcCtl {
  // the A stage is handed over in two phases: [0] = streams of residues 0-2, [1] = residues 3-4
  uint64_t a_full[2];   // producers (256 arrivals) -> MMA: the phase's streams (hi, lo) are in smem
  uint64_t a_empty[2];  // tcgen05.commit -> producers: the MMAs that read the phase's streams are done
  uint64_t acc_full;    // tcgen05.commit -> epilogue: the 5 accumulators of this tile are complete
  uint64_t acc_empty;   // epilogue (384 arrivals) -> MMA: this tile's TMEM slots may be overwritten
  uint64_t b_full;      // TMA bulk copy -> MMA: the B operand image has landed in shared memory
  uint64_t list_full;   // epilogue warps (8 arrivals) -> everybody: the CTA's tile list is complete
  uint32_t tmem_base;
  int32_t n_tiles;      // length of the tile list
};

// log(mel + eps): one MUFU.LG2 + one FMUL (abs error ~ 2^-22 * |log|).  The epsilon is not added here: it
// is the initial value of the bin's accumulator (tc_tables_generated.inc), so `v` is mel + eps already,
// a normal float >= eps.  Silence must give exactly logf(eps) - the reference's golden fixture pins that
// value - and it does: the accumulator is then eps itself, and tc_eps_probe_kernel checks at handle
// creation that this expression reproduces logf(eps) bit-for-bit (true for the reference's 1e-12); a
// handle whose epsilon fails the probe does not get the tensor-core kernel.
__device__ __forceinline__ float tc_log_fast(float v) {
  float l2;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(v));
  return l2 * 0.69314718055994530942f;
}
__device__ __forceinline__ float tc_finish(float acc, float, float) { return tc_log_fast(acc); }
__global__ void tc_eps_probe_kernel(float eps, float* out) { out[0] = tc_log_fast(eps); }

#if defined(TC_ABLATE_ESTS)
#define TC_ROW_ST(PTR, V) if (eps == 12345.0f) *(PTR) = (V) /* (experiment: no staging stores) */
#else
#define TC_ROW_ST(PTR, V) *(PTR) = (V)
#endif
// SpecAugment (data.py:286-288, 299-301: audio *= mask) is applied to the finished value in registers:
// bit m of the row's 80-bit map (mb0..mb2, built at the top of the tile; all ones for a row inside a time
// mask) says "bin m is masked".  The bin index is a literal, so the test is one LOP3 into a predicate and
// one predicated FMUL by 0.0f - which keeps the sign, like the reference's multiply.  (The first version
// re-read the staged row: ~16 shared-memory round trips per tile on the critical epilogue group, +25 %.)
template <int M>
__device__ __forceinline__ float tc_masked(float v, uint32_t b0, uint32_t b1, uint32_t b2) {
  const uint32_t w = M < 32 ? b0 : (M < 64 ? b1 : b2);
  return (w & (1u << (M & 31))) ? v * 0.0f : v;
}
#define TC_FIN(M_, A) (kRegMask ? tc_masked<(M_)>(tc_finish(A, eps, log_eps), mb0, mb1, mb2) : tc_finish(A, eps, log_eps))
// generated epilogue (tc_tables_generated.inc): whole float4 groups of the staged row / of the hand-over zone
#define TC_ST4(G, A, B, C_, D) TC_ROW_ST(reinterpret_cast<float4*>(out_row + 4 * (G)), make_float4(A, B, C_, D))
#define TC_PARK4(S, A, B, C_, D) *reinterpret_cast<float4*>(park_row + 4 * (S)) = make_float4(A, B, C_, D)
#define TC_UNPARK4(S) *reinterpret_cast<const float4*>(park_row + 4 * (S))

#if defined(TC_TRACE)
// developer trace (experiments/trace_roles.py): clock64 stamps of CTA 0, [role][tile][event]
__device__ long long g_tc_trace[5][32][8];
#define TC_STAMP(ROLE, N, EV)                                                   \
  if (blockIdx.x == 0 && (N) < 32) g_tc_trace[ROLE][N][EV] = clock64();
__device__ __forceinline__ long long tc_clock_fenced() {
  long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory");
  return t;
}
#define TC_STAMP_F(ROLE, N, EV)                                                 \
  if (blockIdx.x == 0 && (N) < 32) g_tc_trace[ROLE][N][EV] = tc_clock_fenced();
#else
#define TC_STAMP(ROLE, N, EV)
#define TC_STAMP_F(ROLE, N, EV)
#endif

// SpecAugment on the thread's own staged row (C == 3 only; C == 1 masks in registers, see TC_FIN)
// (data.py:286-288, 299-301): audio *= mask, so a masked
// negative value becomes -0.0 exactly like the reference's multiply.  tb = (start, width) pairs,
// frequency masks first.  Out of line: it runs once per tile and must not bloat the hot loop.
__device__ __noinline__ void tc_mask_row(float* row, int32_t entry, int n_f, int n_t, int t, bool valid) {
  // `entry` = table entry number `lane` of the row's utterance (fetched at the top of the tile); all
  // 32 lanes call this, `valid` = the row is a real frame.  Multiplying by 0.0f keeps the sign like
  // the reference's `audio *= mask`.  Measured (trace stamps): what this costs is not instructions but
  // shared-memory round trips - the pipe is shared with the tensor core's operand fetch (~100 cycles
  // per access while MMAs run) - so whole float4 groups, as few accesses as possible.  Wider batches
  // of loads (8 groups in flight) were tried: the registers they need spill the epilogue's main loop.
  bool tm = false;
  for (int i = 0; i < n_t; ++i) {
    const int s = __shfl_sync(0xffffffffu, entry, 2 * (n_f + i)), w = __shfl_sync(0xffffffffu, entry, 2 * (n_f + i) + 1);
    tm = tm || (t >= s && t < s + w);
  }
  if (tm && valid) {
#pragma unroll 4
    for (int g = 0; g < 20; ++g) {
      float4 v = *reinterpret_cast<float4*>(row + 4 * g);
      v.x *= 0.0f; v.y *= 0.0f; v.z *= 0.0f; v.w *= 0.0f;
      *reinterpret_cast<float4*>(row + 4 * g) = v;
    }
  }
  for (int i = 0; i < n_f; ++i) {
    const int s = __shfl_sync(0xffffffffu, entry, 2 * i), w = __shfl_sync(0xffffffffu, entry, 2 * i + 1);
    const int e = s + w < 80 ? s + w : 80;
    if (valid && !tm && e > s) {
      const int g0 = s >> 2, g1 = (e + 3) >> 2;
#pragma unroll 2
      for (int g = g0; g < g1; ++g) {
        float4 v = *reinterpret_cast<float4*>(row + 4 * g);
        const int m = 4 * g;
        v.x *= (m + 0 >= s && m + 0 < e) ? 0.0f : 1.0f;
        v.y *= (m + 1 >= s && m + 1 < e) ? 0.0f : 1.0f;
        v.z *= (m + 2 >= s && m + 2 < e) ? 0.0f : 1.0f;
        v.w *= (m + 3 >= s && m + 3 < e) ? 0.0f : 1.0f;
        *reinterpret_cast<float4*>(row + 4 * g) = v;
      }
    }
  }
}

__device__ __noinline__ void tc_zero_row(float* row) {
  for (int g = 0; g < 20; ++g) *reinterpret_cast<float4*>(row + 4 * g) = make_float4(0.f, 0.f, 0.f, 0.f);
}

// Unaligned utterance start (ragged offsets that are not a multiple of 8 samples): element-wise
// producer, one group at a time.  Rare path, kept out of line.
__device__ __noinline__ void tc_produce_tile_slow(const short* __restrict__ pcm, int64_t s_begin, int32_t len,
                                                  uint8_t* sA, int ptid) {
  for (int g = ptid; g < 516; g += kTcProducerThreads) {
    uint4 raw[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) raw[q] = tc_load8_slow(pcm, s_begin + 40 * g + 8 * q, len);
    tc_produce_group(raw[0], raw[1], raw[2], raw[3], raw[4], sA, g);
  }
}


// ---------------------------------------------------------------------------------------------
// Work partition.  A tile (128 rows of one utterance) is ACTIVE when it holds at least one real
// frame below t_max.  The active tiles of the batch, in (utterance, tile) order, are split into
// contiguous equal ranges over the CTAs, so that ragged batches stay balanced (round-robin over all
// tiles left some CTAs with 40 % more real work than the average).  The rows behind an utterance's
// last active tile are pure padding: they are zero-filled at the end of the kernel by every warp, in
// 32-row blocks split the same way.  Each warp finds its range itself from the lengths (one pass,
// eight utterances per lane, shuffles): no pre-pass kernel, no global work counter.
// ---------------------------------------------------------------------------------------------
struct TcWalk {  // warp-uniform cursor over a contiguous range of units (active tiles, or padding blocks)
  int32_t b, j, na;  // current utterance, unit inside it, number of units of that utterance
  int32_t left;      // units left in the range, the current one included
  int32_t len, len_nb;  // length of utterance b; of utterance b + 1 (fetched one utterance ahead)
};
__device__ __forceinline__ int32_t tc_nact(int32_t len, int32_t t_max, int32_t tout) {
  int32_t f = frames_of(len, 320, 160);
  f = f < t_max ? f : t_max;
  return (f + tout - 1) / tout;
}
template <bool kPad>
__device__ __forceinline__ int32_t tc_units(int32_t len, int32_t t_max, int32_t tout) {
  const int32_t na = tc_nact(len, t_max, tout);
  if (!kPad) return na;
  const int32_t r = t_max - na * tout;  // rows behind the last active tile
  return r > 0 ? (r + 31) >> 5 : 0;
}
// Range number `part` of `n_parts`.  All 32 lanes call this.  n_frames_out (one warp of the grid only):
// also report every utterance's frame count, and flag the zero-frame utterances the reference's
// SpecAugment would have raised on (data.py:297: maxval <= minval).
template <bool kPad>
__device__ __forceinline__ void tc_walk_begin(TcWalk* dst, const int32_t* __restrict__ lengths, int32_t batch, int32_t t_max,
                                           int32_t tout, uint32_t part, uint32_t n_parts, int lane, int32_t* n_frames_out,
                                           int32_t* status_if_empty) {
  // (the result goes to the warp's slot in shared memory: a reference to a local object would put
  //  the caller's cursor on the stack, and the callers run under setmaxnreg)
  TcWalk w;
  long long total = 0;
  for (int base = 0; base < batch; base += 256) {
    int32_t s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int b = base + 8 * lane + k;
      if (b < batch) {
        const int32_t l = __ldg(lengths + b);
        s += tc_units<kPad>(l, t_max, tout);
        if (n_frames_out || status_if_empty) {
          const int32_t f = frames_of(l, 320, 160);
          if (n_frames_out) n_frames_out[b] = f;
          if (f == 0 && status_if_empty) atomicOr(status_if_empty, 1);
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    total += s;
  }
  const long long a0 = (total * part) / n_parts, a1 = (total * (part + 1)) / n_parts;
  w.left = static_cast<int32_t>(a1 - a0);
  w.b = 0;
  w.j = 0;
  w.na = 0;
  w.len = 0;
  w.len_nb = 0;
  if (w.left <= 0) {
    if (lane == 0) *dst = w;
    __syncwarp();
    return;
  }
  long long acc = 0;
  for (int base = 0; base < batch; base += 256) {
    int32_t c[8], s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int b = base + 8 * lane + k;
      c[k] = b < batch ? tc_units<kPad>(__ldg(lengths + b), t_max, tout) : 0;
      s += c[k];
    }
    int32_t incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const int32_t tot = __shfl_sync(0xffffffffu, incl, 31);
    if (acc + tot <= a0) {
      acc += tot;
      continue;
    }
    const uint32_t hit = __ballot_sync(0xffffffffu, acc + incl > a0);
    const int src = __ffs(hit) - 1;
    int32_t bb = 0, jj = 0;
    long long r = a0 - (acc + incl - s);  // (meaningful in lane `src`)
    bool found = false;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (!found && r < c[k]) {
        bb = base + 8 * lane + k;
        jj = static_cast<int32_t>(r);
        found = true;
      }
      r -= c[k];
    }
    w.b = __shfl_sync(0xffffffffu, bb, src);
    w.j = __shfl_sync(0xffffffffu, jj, src);
    break;
  }
  w.len = __shfl_sync(0xffffffffu, __ldg(lengths + w.b), 0);
  w.na = tc_units<kPad>(w.len, t_max, tout);
  w.len_nb = w.b + 1 < batch ? __ldg(lengths + w.b + 1) : 0;
  if (lane == 0) *dst = w;
  __syncwarp();
}
// next unit of the range (all lanes).  No load on the critical path: the next utterance's length was
// fetched when the current utterance was entered.
template <bool kPad>
__device__ __forceinline__ void tc_walk_step(TcWalk& w, const int32_t* __restrict__ lengths, int32_t batch, int32_t t_max,
                                             int32_t tout) {
  if (--w.left <= 0) return;
  if (++w.j < w.na) return;
  w.j = 0;
  do {  // (utterances without units are skipped; left > 0 guarantees that one with units follows)
    ++w.b;
    w.len = __shfl_sync(0xffffffffu, w.len_nb, 0);
    w.len_nb = w.b + 1 < batch ? __ldg(lengths + w.b + 1) : 0;
    w.na = tc_units<kPad>(w.len, t_max, tout);
  } while (w.na == 0 && w.b + 1 < batch);
}

// SpecAugment draws of this CTA's tiles, by the 384 epilogue threads while they wait for the first
// accumulators: raw Philox words first (one per thread), then one thread per tile turns them into
// the reference's (start, width) table in draw order (common.cuh: specaug_draw_table).
struct TcDrawArgs {
  int32_t* status;
  uint64_t seed, first_utt;
  int32_t n_masks_f, n_masks_t, sa_F, sa_T;
  float sa_p;
};
__device__ __noinline__ void tc_draw_tables(int32_t* tab, const TcDrawArgs p, const int4* list, int n_tiles, int etid) {
  const int n_e = 2 * (p.n_masks_f + p.n_masks_t);
  for (int k = etid; k < n_tiles * kTcMaskEntries; k += kTcEpilogueThreads) {
    const int i = k / kTcMaskEntries, d = k - i * kTcMaskEntries;
    if (d < n_e)
      tab[k] = static_cast<int32_t>(specaug_raw_draw(p.seed, p.first_utt + static_cast<uint64_t>(list[i].x), static_cast<uint32_t>(d)));
  }
  tc::named_bar_sync(13, kTcEpilogueThreads);
  for (int i = etid; i < n_tiles; i += kTcEpilogueThreads) {
    const int32_t num_time = frames_of(list[i].z, 320, 160);
    int32_t* e = tab + i * kTcMaskEntries;
    int bad = 0, o = 0;
    for (int q = 0; q < p.n_masks_f; ++q) {
      int32_t f = static_cast<int32_t>(static_cast<uint32_t>(e[o]) % static_cast<uint32_t>(p.sa_F));
      const uint32_t u = static_cast<uint32_t>(e[o + 1]);
      const int32_t range = 80 - f;
      int32_t f0 = 0;
      if (range <= 0) { bad = 1; f = 0; } else { f0 = static_cast<int32_t>(u % static_cast<uint32_t>(range)); }
      e[o++] = f0;
      e[o++] = f;
    }
    int32_t applied = 0;
    const int32_t cap = static_cast<int32_t>(static_cast<float>(num_time) * p.sa_p);
    for (int q = 0; q < p.n_masks_t; ++q) {
      int32_t t = static_cast<int32_t>(static_cast<uint32_t>(e[o]) % static_cast<uint32_t>(p.sa_T));
      const int32_t room = cap - applied;
      t = t < room ? t : room;
      applied += t;
      const uint32_t u = static_cast<uint32_t>(e[o + 1]);
      const int32_t range = num_time - t;
      int32_t t0 = 0;
      if (range <= 0) { bad = 1; t = 0; } else { t0 = static_cast<int32_t>(u % static_cast<uint32_t>(range)); }
      e[o++] = t0;
      e[o++] = t;
    }
    if (bad && p.status) atomicOr(p.status, 1);
  }
  tc::named_bar_sync(13, kTcEpilogueThreads);
}

// warpgroup-wide register re-allocation relative to the launch allocation (nothing to do when equal)
template <int N>
__device__ __forceinline__ void tc_set_regs() {
  if (N < kTcLaunchRegs) tc::setmaxnreg_dec<N>();
  if (N > kTcLaunchRegs) tc::setmaxnreg_inc<N>();
}

// kF32: the PCM is float32 (int16 / 32768 as load_audio_file gives it, data.py:98-116), dense rows, 16-byte
// aligned: the producers convert it back to the 16-bit samples on the fly (tc_f32_block) and OR 1 into
// p.f32_flag when a sample they used is not exact - the caller then runs the generic kernel behind this one.
template <int C, bool kMask, bool kF32>
__global__ void __launch_bounds__(kTcThreads, 1)
    tc_frontend_kernel(const KernelParams p, const TcParams tp, const __grid_constant__ CUtensorMap out_map) {
  extern __shared__ __align__(1024) uint8_t tc_smem_raw[];
  // 1024-byte alignment (swizzle atoms) without losing the shared address space of the pointer
  uint8_t* smem = tc_smem_raw + ((1024u - (tc::smem_u32(tc_smem_raw) & 1023u)) & 1023u);
  uint8_t* sB = smem + kTcSmemB;
  uint8_t* sA = smem + kTcSmemA;
  float* sOut = reinterpret_cast<float*>(smem + kTcSmemOut);
  float* sPark = reinterpret_cast<float*>(smem + kTcSmemPark);
  TcCtl* ctl = reinterpret_cast<TcCtl*>(smem + kTcSmemCtl);

  // warp index through a shuffle: the compiler then knows it is warp-uniform (uniform registers and
  // branches for everything derived from it: roles, TMEM addresses, chunk dispatch)
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;

  // ---- one-time setup ------------------------------------------------------------------------
  if (tid == 0) { TC_STAMP(1, 31, 0) }
  if (tid == 32) {
    tc::mbar_init(&ctl->b_full, 1);
    for (int r = 0; r < 2; ++r) {
      tc::mbar_init(&ctl->a_full[r], kTcProducerThreads);
      tc::mbar_init(&ctl->a_empty[r], 1);
    }
    tc::mbar_init(&ctl->acc_full, 1);
    tc::mbar_init(&ctl->acc_empty, kTcEpilogueThreads);
    tc::mbar_init(&ctl->list_full, 8);
    tc::fence_mbar_init();
    // B operand (80 KB, constant): two bulk copies by the TMA engine, only the MMA issuer waits for
    // them - producers and epilogue warps start on the first tile at once
    tc::mbar_expect_tx(&ctl->b_full, 10 * kTcBBytes);
    tc::bulk_copy_g2s(tc::smem_u32(sB), tp.b_image, 5 * kTcBBytes, &ctl->b_full);
    tc::bulk_copy_g2s(tc::smem_u32(sB) + 5 * kTcBBytes, tp.b_image + 5 * kTcBBytes, 5 * kTcBBytes, &ctl->b_full);
  }
  if (warp == kTcMmaWarp) tc::tmem_alloc(&ctl->tmem_base, 512);
  const int32_t t_max32 = p.t_max < 0x7fffffff ? static_cast<int32_t>(p.t_max) : 0x7fffffff;
  TcWalk* swk = reinterpret_cast<TcWalk*>(smem + kTcSmemWalk) + (warp & 15);  // (warps 0-7 use their cursor slot)
  // The CTA's tiles in processing order, (utterance, tile, length, 8-sample block offset or bit 31 =
  // "not 16-byte aligned"): built by the epilogue warps while they wait for the first accumulators,
  // read by every role (one 16-byte broadcast load per tile instead of cursor arithmetic).
  int4* sList = reinterpret_cast<int4*>(smem + kTcSmemList);
  const bool pcm_al = (reinterpret_cast<uintptr_t>(p.pcm) & 15) == 0;
  auto fetch_blk = [&](int bt) -> uint32_t {
    const int64_t o = p.offsets ? __ldg(p.offsets + bt) : static_cast<int64_t>(bt) * p.row_stride;
    return (pcm_al && (o & 7) == 0) ? static_cast<uint32_t>(o >> 3) : 0x80000000u;
  };
  if (tid == 64 && p.offsets == nullptr && (reinterpret_cast<uintptr_t>(p.pcm) & 15) == 0 && (p.row_stride & 7) == 0) {
    // Pull the PCM range of this CTA's first tile into L2 before anything else: its loads otherwise
    // start only after the (cold) pass over the lengths.  The tile is guessed as if every tile of the
    // batch were active (exact for a batch of full-length utterances) and the range is bounded by the
    // row, not by the utterance; an L2 prefetch has no architectural effect, so a wrong guess is
    // harmless and it may precede the dependency wait below.
    const int t0 = static_cast<int>((static_cast<long long>(tp.total_tiles) * blockIdx.x) / gridDim.x);
    const int b0 = t0 / tp.tiles_per_utt, j0 = t0 - b0 * tp.tiles_per_utt;
    int64_t s0 = 160 * static_cast<int64_t>(j0 * tp.tout - tp.halo), s1 = s0 + 129 * 160;
    if (s0 < 0) s0 = 0;
    if (s1 > p.row_stride) s1 = p.row_stride;
    if (s1 > s0)
      tc::prefetch_l2(static_cast<const char*>(p.pcm) + (static_cast<int64_t>(b0) * p.row_stride + s0) * (kF32 ? 4 : 2),
                      static_cast<uint32_t>((kF32 ? 4 : 2) * (s1 - s0)));
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = ctl->tmem_base;
  bool timed_out = false;
  // Programmatic dependent launch: everything above (barriers, TMEM, the constant B image) may overlap
  // the tail of the previous kernel in the stream; caller data (PCM, lengths, tables, the output) is
  // only touched after that kernel has completed.
  tc::pdl_launch_dependents();
  if (tid == 0) { TC_STAMP(1, 31, 1) }
  tc::pdl_wait();
  if (tid == 0) { TC_STAMP(1, 31, 2) }
  // float32 entry: when the conversion kernel found a sample that is not an exact int16, this launch
  // does nothing (an empty batch) and the generic kernel, launched behind it, takes the batch
  const int32_t batch_eff = (p.gate && (*p.gate != 0) != (p.gate_value != 0)) ? 0 : p.batch;

  if (warp >= 1 && warp < 8) {
    // ======================= PRODUCERS (warps 1-7: 224 threads) =================================
    tc_set_regs<kTcRegsProducer>();
    // One A stage, refilled in two phases: the streams of residues 0-2 as soon as the previous tile's
    // MMAs of those residues have completed, residues 3-4 after the rest (those MMAs wait for the
    // epilogue to release TMEM).  The raw PCM of the NEXT tile is loaded into the same registers at
    // the end of the tile, and its range was pulled into L2 one tile earlier.
    const int ptid = tid - 32;
    uint32_t n = 0;  // active tiles processed so far by this CTA
    // 516 groups of 40 samples = 129 hops (128 frames + the half-row of overlap): groups ptid and
    // ptid + 224 for every thread, ptid + 448 for the first 68 - all held in registers and preloaded for
    // the next tile.  (Loading the third group on the spot, out of line, to save 20 registers was measured:
    // the L2 round trips of those three warps sit on the a_full barrier - 83.1 vs 79.5 us per launch.)
    const bool third = ptid < 516 - 2 * kTcProducerThreads;
    uint32_t ph[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) ph[i] = tc_stream_offset(ptid + kTcProducerThreads * i);
    const uint4* pcm16 = reinterpret_cast<const uint4*>(p.pcm);
    uint32_t w[3][20];  // this thread's groups: raw words (two samples each), XOR-ed in place before use
    bool preloaded = false;
    TcF32Check f32_bad;  // (kF32) did a sample this thread converted fail to be an exact int16 / 32768?
    // raw words of one tile -> w: absolute 8-sample block index, clamped into the utterance
    // (a group whose five blocks all lie inside the utterance - every group of an interior tile - needs no
    //  clamping: one base address, immediate offsets; the clamped form is for the utterance's edges)
#define TC_LOAD_GROUP(I, BLK0, BLO, BHI)                                                               \
  {                                                                                                    \
    const int32_t g0 = (BLK0) + 5 * (ptid + kTcProducerThreads * (I));                                 \
    const bool inside = g0 >= (BLO) && g0 + 4 <= (BHI);                                                \
    if (kF32) { /* a block of 8 samples = two 128-bit loads of floats; the group's ten loads go out in a row */ \
      /* (a run of loads = one L2 round trip on the producers' critical path; 18 loads per run were tried:  */ \
      /*  the 96 registers the kernel is compiled for hold 12, the rest spills - three runs of ten it is)   */ \
      uint4 t32[10];                                                                                   \
      if (inside) {                                                                                    \
        const uint4* src = pcm16 + 2 * static_cast<int64_t>(g0);                                       \
        _Pragma("unroll") for (int q = 0; q < 10; ++q) t32[q] = tc::ldg_nc_v4(src + q);                \
      } else {                                                                                         \
        _Pragma("unroll") for (int q = 0; q < 5; ++q) {                                                \
          const int64_t blk = min(max(g0 + q, (BLO)), (BHI));                                          \
          t32[2 * q] = tc::ldg_nc_v4(pcm16 + 2 * blk);                                                 \
          t32[2 * q + 1] = tc::ldg_nc_v4(pcm16 + 2 * blk + 1);                                         \
        }                                                                                              \
      }                                                                                                \
      _Pragma("unroll") for (int q = 0; q < 5; ++q) tc_f32_block(t32[2 * q], t32[2 * q + 1], &w[I][4 * q], f32_bad); \
    } else {                                                                                           \
      uint4 t16[5];                                                                                    \
      if (inside) {                                                                                    \
        const uint4* src = pcm16 + g0;                                                                 \
        _Pragma("unroll") for (int q = 0; q < 5; ++q) t16[q] = __ldg(src + q);                         \
      } else {                                                                                         \
        _Pragma("unroll") for (int q = 0; q < 5; ++q) t16[q] = __ldg(pcm16 + min(max(g0 + q, (BLO)), (BHI))); \
      }                                                                                                \
      _Pragma("unroll") for (int q = 0; q < 5; ++q) {                                                  \
        w[I][4 * q + 0] = t16[q].x;                                                                    \
        w[I][4 * q + 1] = t16[q].y;                                                                    \
        w[I][4 * q + 2] = t16[q].z;                                                                    \
        w[I][4 * q + 3] = t16[q].w;                                                                    \
      }                                                                                                \
    }                                                                                                  \
  }
#if !defined(TC_NO_SPEC_FIRST)
    // Dense, 16-byte aligned input: load the CTA's first tile now, while the epilogue warps still build
    // the tile list - guessed as if every tile of the batch were active (exact for a batch of
    // full-length utterances; the same guess as the L2 prefetch above).  The block index is clamped
    // into the ROW here, not into the utterance: the samples between `len` and the end of the row
    // only feed frames >= n_frames, which the epilogue zero-fills or drops.  A wrong guess is found
    // by comparing with the list entry below and costs one wasted load.  Measured (r2k, 256 x 15 s):
    // 76.9 -> 76.1 us per launch in the 20-step bench, 72.4 -> 72.0 us in the 100-step one.
    int32_t spec_b = -1, spec_j = -1;
    // (not with float32 input: samples behind `len` would go through the exactness test)
    if (!kF32 && p.offsets == nullptr && pcm_al && (p.row_stride & 7) == 0 && batch_eff > 0) {
      const int t0 = static_cast<int>((static_cast<long long>(tp.total_tiles) * blockIdx.x) / gridDim.x);
      spec_b = t0 / tp.tiles_per_utt;
      spec_j = t0 - spec_b * tp.tiles_per_utt;
      if (spec_b < batch_eff && static_cast<int64_t>(spec_b + 1) * (p.row_stride >> 3) < 0x7fffffffLL) {
        const int32_t b_lo = spec_b * static_cast<int32_t>(p.row_stride >> 3), b_hi = b_lo + static_cast<int32_t>(p.row_stride >> 3) - 1;
        const int32_t b_0 = b_lo + 20 * (spec_j * tp.tout - tp.halo);
        TC_LOAD_GROUP(0, b_0, b_lo, b_hi)
        TC_LOAD_GROUP(1, b_0, b_lo, b_hi)
        if (third) { TC_LOAD_GROUP(2, b_0, b_lo, b_hi) }
      } else {
        spec_b = -1;
      }
    }
#endif
    if (!tc::mbar_wait(&ctl->list_full, 0)) timed_out = true;
    const int n_tiles = ctl->n_tiles;
#if !defined(TC_NO_SPEC_FIRST)
    if (spec_b >= 0 && n_tiles > 0) {
      const int4 first = sList[0];
      preloaded = first.x == spec_b && first.y == spec_j && first.w >= 0;
    }
#endif
    // Register budget: the 60 raw words of the NEXT tile are live across the top of the loop, so
    // nothing about the tiles is carried in registers - every use re-reads the list entry.
    // next tile: first block and clamp range, all in absolute 8-sample blocks (< 2^31 blocks)
#define TC_NEXT_RANGE()                                                                           \
  const int4 nxt = sList[ti + 1];                                                                 \
  const int32_t nb_lo = nxt.w, nb_hi = nb_lo + (nxt.z >> 3) - 1;                                  \
  const int32_t nb_0 = nb_lo + 20 * (nxt.y * tp.tout - tp.halo);
    for (int ti = 0; ti < n_tiles; ++ti) {
      const bool next_ok = ti + 1 < n_tiles && sList[ti + 1].w >= 0;  // (bit 31 = unaligned: slow path)
      if (next_ok && ptid == kTcProducerThreads - 1) {
        // ask the TMA engine to pull the next tile's PCM range into L2
        TC_NEXT_RANGE()
        const int32_t q0 = nb_0 < nb_lo ? nb_lo : nb_0;
        const int32_t q1 = nb_0 + 129 * 20 - 1 < nb_hi ? nb_0 + 129 * 20 - 1 : nb_hi;
        if (q1 >= q0) tc::prefetch_l2(pcm16 + (kF32 ? 2 : 1) * static_cast<int64_t>(q0), (kF32 ? 32u : 16u) * (q1 - q0 + 1));
      }
      const uint32_t bb = static_cast<uint32_t>(sList[ti].w);
      const uint32_t par = (n - 1) & 1;
      if (kF32 || !(bb & 0x80000000u)) {   // (float32: the host only takes this kernel for aligned dense rows)
        // Every frame t < n_fr ends at a multiple of 8 samples <= len, so clamping the 8-sample
        // block index into [0, len/8 - 1] never changes a sample a valid frame uses; blocks outside
        // only feed rows that the epilogue zero-fills (t >= n_fr) or drops (t < 0).
        if (ptid == 0) { TC_STAMP(0, n, 0) }
        if (!preloaded) {
          const int4 cur = sList[ti];
          const int32_t b_lo = cur.w, b_hi = b_lo + (cur.z >> 3) - 1;  // n_fr >= 1: len >= 320
          const int32_t b_0 = b_lo + 20 * (cur.y * tp.tout - tp.halo);
          TC_LOAD_GROUP(0, b_0, b_lo, b_hi)
          TC_LOAD_GROUP(1, b_0, b_lo, b_hi)
          if (third) { TC_LOAD_GROUP(2, b_0, b_lo, b_hi) }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
          for (int k = 0; k < 20; ++k) w[i][k] ^= 0x80008000u;
        if (ptid == 0) { TC_STAMP(0, n, 1) }
        // ---- phase 1: residues 0-2
        if (n > 0 && !timed_out && !tc::mbar_wait(&ctl->a_empty[0], par)) timed_out = true;
        if (ptid == 0) { TC_STAMP(0, n, 2) }
        tc_produce_res<0>(w[0], sA, ph[0]);
        tc_produce_res<1>(w[0], sA, ph[0]);
        tc_produce_res<2>(w[0], sA, ph[0]);
        tc_produce_res<0>(w[1], sA, ph[1]);
        tc_produce_res<1>(w[1], sA, ph[1]);
        tc_produce_res<2>(w[1], sA, ph[1]);
        if (third) {
          tc_produce_res<0>(w[2], sA, ph[2]);
          tc_produce_res<1>(w[2], sA, ph[2]);
          tc_produce_res<2>(w[2], sA, ph[2]);
        }
        tc::fence_async_smem();  // generic-proxy writes -> visible to the tensor core (async proxy)
        tc::mbar_arrive(&ctl->a_full[0]);
        if (ptid == 0) { TC_STAMP(0, n, 3) }
        // ---- phase 2: residues 3-4, then reload the registers with the next tile's PCM
        if (n > 0 && !timed_out && !tc::mbar_wait(&ctl->a_empty[1], par)) timed_out = true;
        if (ptid == 0) { TC_STAMP(0, n, 4) }
        tc_produce_res<3>(w[0], sA, ph[0]);
        tc_produce_res<4>(w[0], sA, ph[0]);
        tc_produce_res<3>(w[1], sA, ph[1]);
        tc_produce_res<4>(w[1], sA, ph[1]);
        if (third) {
          tc_produce_res<3>(w[2], sA, ph[2]);
          tc_produce_res<4>(w[2], sA, ph[2]);
        }
        tc::fence_async_smem();
        tc::mbar_arrive(&ctl->a_full[1]);
        if (ptid == 96) { TC_STAMP(0, n, 5) }
        // the registers are free: start loading the next tile's PCM (its range is in L2 by now).
        // After the fence on purpose - the fence waits for this thread's outstanding loads.
        if (next_ok) {
          TC_NEXT_RANGE()
          TC_LOAD_GROUP(0, nb_0, nb_lo, nb_hi)
          TC_LOAD_GROUP(1, nb_0, nb_lo, nb_hi)
          if (third) { TC_LOAD_GROUP(2, nb_0, nb_lo, nb_hi) }
        }
        preloaded = next_ok;
        if (ptid == 0) { TC_STAMP(0, n, 6) }
      } else {
        for (int r = 0; r < 2; ++r)
          if (n > 0 && !timed_out && !tc::mbar_wait(&ctl->a_empty[r], par)) timed_out = true;
        const int4 cur = sList[ti];
        const int64_t base = p.offsets ? p.offsets[cur.x] : static_cast<int64_t>(cur.x) * p.row_stride;
        tc_produce_tile_slow(reinterpret_cast<const short*>(p.pcm) + base, 160 * (cur.y * tp.tout - tp.halo), cur.z, sA, ptid);
        tc::fence_async_smem();
        for (int r = 0; r < 2; ++r) tc::mbar_arrive(&ctl->a_full[r]);
        preloaded = false;
      }
      ++n;
    }
#undef TC_NEXT_RANGE
#undef TC_LOAD_GROUP
    if (kF32 && f32_bad.bad() && p.f32_flag) atomicOr(p.f32_flag, 1);
  } else if (warp == 0) {
    // ======================= MMA ISSUER (one thread of warp 0) ==================================
    // Warp 0 on purpose: the issuing thread needs an instruction slot every few cycles to keep the
    // tensor pipe fed, and the warp schedulers favour the oldest (lowest) warp of a sub-partition -
    // as the youngest warp it was starved by the epilogue warps (96 instead of 60 cycles per MMA).
    tc_set_regs<kTcRegsProducer>();
    if (tc::elect_one()) {
      // the issuer only needs the NUMBER of tiles of this CTA (every tile of the list is active)
      if (!tc::mbar_wait(&ctl->list_full, 0)) timed_out = true;
      const uint32_t n_tiles = static_cast<uint32_t>(ctl->n_tiles);
      const uint32_t idesc = tc::make_idesc_f16(128, 64);
      const uint64_t a_base = tc::make_smem_desc(tc::smem_u32(sA), 16, 512, tc::kLayoutSw64);
      const uint64_t b_base = tc::make_smem_desc(tc::smem_u32(sB), 16, 1024, tc::kLayoutSw128);
      const uint32_t a_lo0 = static_cast<uint32_t>(a_base), a_hi = static_cast<uint32_t>(a_base >> 32);
      const uint32_t b_lo0 = static_cast<uint32_t>(b_base), b_hi = static_cast<uint32_t>(b_base >> 32);
      if (!tc::mbar_wait(&ctl->b_full, 0)) timed_out = true;
      for (uint32_t n = 0; n < n_tiles; ++n) {
#pragma unroll 1
        for (int r = 0; r < 5; ++r) {
          if (r == 0) {
            TC_STAMP(1, n, 0)
            if (!timed_out && !tc::mbar_wait(&ctl->a_full[0], n & 1)) timed_out = true;
            tc::tc_fence_after();
            TC_STAMP(1, n, 1)
          }
          if (r == 3) {
            TC_STAMP(1, n, 2)
            // slots of residues 3, 4 were residues 0, 1 of the previous tile: wait for its epilogue
            if (!timed_out && !tc::mbar_wait(&ctl->a_full[1], n & 1)) timed_out = true;
            TC_STAMP(1, n, 5)
            if (n > 0 && !timed_out && !tc::mbar_wait(&ctl->acc_empty, (n - 1) & 1)) timed_out = true;
            tc::tc_fence_after();
          }
          if (r == 3) { TC_STAMP(1, n, 3) }
          const uint32_t d_tmem = tmem_base + (((5u * n + r) & 7u) << 6);
          // passes, smallest magnitude first: lo x Bhi, hi x Blo, hi x Bhi; descriptor offsets in
          // 16-byte units: stream (term, r) at (5 term + r) * 544, matrix (term, r) at (5 term + r) * 512
          const uint32_t a_r = a_lo0 + 544u * r, b_r = b_lo0 + 512u * r;
#pragma unroll
          for (int ps = 0; ps < 3; ++ps) {
            const uint32_t a_t = ps == 0 ? 5u * 544u : 0u, b_t = ps == 1 ? 5u * 512u : 0u;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              if (ps == 0 && ks == 0)
                tc::umma_f16<false>(d_tmem, a_r + a_t + 2u * ks, a_hi, b_r + b_t + 2u * ks, b_hi, idesc);
              else
                tc::umma_f16<true>(d_tmem, a_r + a_t + 2u * ks, a_hi, b_r + b_t + 2u * ks, b_hi, idesc);
            }
          }
          // commits only at the two hand-over points: a commit between residues drains the tensor
          // pipe (measured: ~300 cycles per commit)
          if (r == 2) tc::umma_commit(&ctl->a_empty[0]);  // the streams of residues 0-2 may be refilled
        }
        tc::umma_commit(&ctl->a_empty[1]);
        tc::umma_commit(&ctl->acc_full);
        TC_STAMP(1, n, 4)
      }
    }
    __syncwarp();
  } else {
    // ======================= EPILOGUE (warps 8-19) ==============================================
    // thread = frame row = TMEM lane.  Three groups share every tile: group A (warps 8-11) reduces the
    // accumulator quads [0, TC_SPLIT_A) of each residue, B (12-15) [TC_SPLIT_A, TC_SPLIT_B), C (16-19) the
    // rest; warps w, w + 4, w + 8 work on the same 32 rows.  Each group has its own straight-line code and
    // its own mel accumulators (only the bins its FFT bins feed).  Values travel down the chain A -> B -> C
    // through the hand-over zone: partial sums of bins two groups feed, and finished values on their way to
    // the group that stores their float4 group.  C then stores the 32 rows with one TMA tensor store.
    tc_set_regs<kTcRegsEpilogue>();
    // C == 1: masks are applied to the finished values in registers (measured 102.7 -> 96.7 us on config 3);
    // C == 3: on the staged row before the copy-out
    constexpr bool kRegMask = kMask && C == 1;
    // (two groups: the second one is "group 2" - the last group - in the generated code as well)
    // (measured r2k: the last group of the chain on the OLDEST epilogue warps instead - no change, 72.36 vs 72.44 us)
    const int group = TC_NGROUPS == 3 ? (warp - 8) >> 2 : ((warp - 8) >> 2) * 2, wq = warp & 3;
    const int etid = tid - 256;
    const int row = wq * 32 + lane;
    uint32_t n = 0;
    const float eps = p.epsilon, log_eps = tp.log_eps;
    const int n_masks = p.n_masks_f + p.n_masks_t;
    float* out_row = sOut + row * kTcOutPitch;
    float* park_row = sPark + row * TC_PARK_PITCH;
    const uint32_t sOut_s = tc::smem_u32(sOut);
    const uint32_t lane_base = static_cast<uint32_t>(wq * 32) << 16;
    // barriers of the warps that share 32 rows: "rows free" C -> A (, B); hand-overs A -> B, B -> C (64 threads;
    // two groups: A -> C on bar_bc)
    const uint32_t bar_free = 1 + wq, bar_ab = 5 + wq, bar_bc = 9 + wq;
    constexpr uint32_t kFreeCount = 32 * TC_NGROUPS;
    bool issuer = false;  // (group C) this thread issued the trio's TMA stores (bulk async-groups are per thread)
    int32_t* sMask = reinterpret_cast<int32_t*>(smem + kTcSmemMask);
    // the first epilogue warp of the grid also reports the frame counts, and flags the zero-frame
    // utterances on which the reference's time-mask draw raises
    const bool reporter = blockIdx.x == 0 && warp == 8;
    constexpr int kTout = kTcTileM - (C == 3 ? 2 : 0);  // = tp.tout, as a constant: divisions by it are multiplications
    if (warp < 16) {
      // ---- the CTA's tile list, in one pass over the lengths (warps 8-15): thread (warp e, lane l) owns
      //      utterance 256 c + 8 l + e of every chunk c of 256; a warp scan of the per-lane counts gives
      //      the index of its first active tile, and it writes those of its tiles that fall into this
      //      CTA's range.  Ready ~one L2 round trip after the kernel starts. ----
      const int ew = warp - 8;
      if (etid == 0) { TC_STAMP_F(2, 30, 0) }
      int32_t* nfo = reporter ? p.n_frames : nullptr;
      int32_t* sfe = (reporter && p.n_masks_t > 0) ? p.status : nullptr;
      long long total = 0;
      int32_t c[8], mylen = 0;
      for (int base = 0; base < batch_eff; base += 256) {
        int32_t sum = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int bq = base + 8 * lane + k;
          const int32_t l = bq < batch_eff ? __ldg(p.lengths + bq) : 0;
          c[k] = bq < batch_eff ? tc_nact(l, t_max32, kTout) : 0;
          sum += c[k];
          if (k == ew) mylen = l;
          if (bq < batch_eff && (nfo || sfe)) {
            const int32_t f = frames_of(l, 320, 160);
            if (nfo) nfo[bq] = f;
            if (f == 0 && sfe) atomicOr(sfe, 1);
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        total += sum;
      }
      // (64-bit divisions are subroutine calls: the usual batch takes the 32-bit ones)
      long long a0, a1r;
      if (total < (1 << 20) && gridDim.x < (1 << 11)) {
        const uint32_t tt = static_cast<uint32_t>(total);
        a0 = (tt * blockIdx.x) / gridDim.x;
        a1r = (tt * (blockIdx.x + 1)) / gridDim.x;
      } else {
        a0 = (total * blockIdx.x) / gridDim.x;
        a1r = (total * (blockIdx.x + 1)) / gridDim.x;
      }
      const long long a1 = a1r - a0 > kTcListTiles ? a0 + kTcListTiles : a1r;  // (the host keeps ranges within the list)
      if (etid == 0) { TC_STAMP_F(2, 30, 1) }
      if (ew == 0 && lane == 0) ctl->n_tiles = static_cast<int32_t>(a1 - a0);
      long long acc = 0;
      for (int base = 0; base < batch_eff; base += 256) {
        if (batch_eff > 256) {  // more than one chunk: the counts of this one again (L1 / L2 hits)
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int bq = base + 8 * lane + k;
            const int32_t l = bq < batch_eff ? __ldg(p.lengths + bq) : 0;
            c[k] = bq < batch_eff ? tc_nact(l, t_max32, kTout) : 0;
            if (k == ew) mylen = l;
          }
        }
        int32_t sum = 0, before = 0, mine = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (k < ew) before += c[k];
          if (k == ew) mine = c[k];
          sum += c[k];
        }
        int32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int32_t v = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += v;
        }
        const long long first = acc + (incl - sum) + before;  // active-tile index of my utterance's tile 0
        acc += __shfl_sync(0xffffffffu, incl, 31);
        const int bq = base + 8 * lane + ew;
        long long jl = a0 - first, jh = a1 - first;
        if (jl < 0) jl = 0;
        if (jh > mine) jh = mine;
        if (jl < jh) {
          const int bbv = static_cast<int>(fetch_blk(bq));
          for (int j = static_cast<int>(jl); j < static_cast<int>(jh); ++j)
            sList[first + j - a0] = make_int4(bq, j, mylen, bbv);
        }
      }
      __syncwarp();
      if (etid == 0) { TC_STAMP_F(2, 30, 2) }
      if (lane == 0) tc::mbar_arrive(&ctl->list_full);
    }
    if (!tc::mbar_wait(&ctl->list_full, 0)) timed_out = true;
    if (etid == 0) { TC_STAMP_F(2, 30, 3) }
    const int n_tiles = ctl->n_tiles;
    if (n_masks > 0 && !p.mask_table) {
      TcDrawArgs da;
      da.status = p.status;
      da.seed = p.seed;
      da.first_utt = p.first_utt;
      da.n_masks_f = p.n_masks_f;
      da.n_masks_t = p.n_masks_t;
      da.sa_F = p.sa_F;
      da.sa_T = p.sa_T;
      da.sa_p = p.sa_p;
      tc_draw_tables(sMask, da, sList, n_tiles, etid);
    }
    for (int ti = 0; ti < n_tiles; ++ti) {
      const int4 cur = sList[ti];
      const int b = cur.x, j = cur.y;
      const int32_t len = cur.z;
      // SpecAugment: lane i fetches entry i of the utterance's mask table now (at most 16 masks on this
      // path): C == 1 every group (masks in registers), C == 3 group C (masks on the staged row)
      int32_t mask_entry = 0;
      if (kMask && (kRegMask || group == 2) && lane < 2 * n_masks)
        mask_entry = p.mask_table ? __ldg(p.mask_table + static_cast<size_t>(b) * 2 * n_masks + lane)
                                  : sMask[ti * kTcMaskEntries + lane];
      const int n_fr = frames_of(len, 320, 160);
      const int t_first = j * tp.tout;  // first OUTPUT frame of this tile; row i holds frame t_first - halo + i
      const int t_row = t_first - tp.halo + row;
      // this row's mask map: frequency masks [f0, f0 + f) of the utterance, everything for a row inside a
      // time mask [t0, t0 + t).  (Rows that are not real frames are zero-filled or dropped anyway.)
      uint32_t mb0 = 0, mb1 = 0, mb2 = 0;
      if (kRegMask) {
        for (int i = 0; i < p.n_masks_f; ++i) {
          const int s = __shfl_sync(0xffffffffu, mask_entry, 2 * i), w = __shfl_sync(0xffffffffu, mask_entry, 2 * i + 1);
          const int e = s + w;  // (bits >= 80 are never tested)
          if (w > 0) {
            // bits [s, e) of a 96-bit map, word by word: [lo, hi) clipped to the word
            const int l0 = s < 32 ? s : 32, h0 = e < 32 ? e : 32;
            const int l1 = s < 32 ? 0 : (s < 64 ? s - 32 : 32), h1 = e < 32 ? 0 : (e < 64 ? e - 32 : 32);
            const int l2 = s < 64 ? 0 : (s < 96 ? s - 64 : 32), h2 = e < 64 ? 0 : (e < 96 ? e - 64 : 32);
            if (h0 > l0) mb0 |= (0xffffffffu >> (32 - (h0 - l0))) << l0;
            if (h1 > l1) mb1 |= (0xffffffffu >> (32 - (h1 - l1))) << l1;
            if (h2 > l2) mb2 |= (0xffffffffu >> (32 - (h2 - l2))) << l2;
          }
        }
        bool tm = false;
        for (int i = 0; i < p.n_masks_t; ++i) {
          const int s = __shfl_sync(0xffffffffu, mask_entry, 2 * (p.n_masks_f + i)),
                    w = __shfl_sync(0xffffffffu, mask_entry, 2 * (p.n_masks_f + i) + 1);
          tm = tm || (t_row >= s && t_row < s + w);
        }
        if (tm) mb0 = mb1 = mb2 = 0xffffffffu;
      }
      if (lane == 0 && wq == 0) { TC_STAMP(2 + group, n, 0) }
      if (!timed_out && !tc::mbar_wait(&ctl->acc_full, n & 1)) timed_out = true;
      tc::tc_fence_after();
      if (lane == 0 && wq == 0) { TC_STAMP(2 + group, n, 1) }
      const uint32_t t_tile = tmem_base + lane_base;
      // One group's pass over its quads: five TMEM loads (one per residue) of four columns, the packed
      // radix-5 + power step, the quad's mel FFMAs (immediates, named accumulators).  The TMEM slots go
      // back to the MMA issuer right after the group's last load; all finishing work comes after that.
      // C == 1: the staged rows and the hand-over zone of the previous tile are free once group C's TMA
      // store has read them and C's phase 2 is over - C says so at its first quad (`CFREE`), A and B check
      // before their first write (`ABFREE`), long after.
#ifndef TC_LOOP_MODE
#define TC_LOOP_MODE 0   // measured: rolled 78.7 us, straight-line + software-pipelined 87.9 us per launch (see below)
#endif
#ifndef TC_CFREE_Q
#define TC_CFREE_Q 0     // quad (counted from the group's first) after which group C frees the previous tile's rows (r2k: 2 = no change)
#endif
#define TC_QUAD_ADDR(Q, R) (t_tile + (((5u * n + (R)) & 7u) << 6) + 4 * (Q))
#if TC_LOOP_MODE == 0
      // rolled loop, one quad per iteration: load, wait, compute - except the last TC_TAIL_K quads, which are
      // loaded together into registers BEFORE the arithmetic on them: the chain "TMEM read phase -> release
      // -> the next tile's MMAs of residues 3-4 -> acc_full" is what the tile period is made of, and every
      // quad of arithmetic taken out of the read phase shortens it.
#define TC_QUAD_MATH(G, QQ, D)                                                                    \
    if ((G) == 0 && (QQ) == 0) tc_radix5_power_special(D, P);                                     \
    else tc_radix5_power_pair(D, P);
#define TC_GROUP_LOOP(G, Q0, Q1, CFREE)                                                           \
  float P[11], d[TC_TAIL_K][5][4];                                                                \
  constexpr int kTail = (Q1) - (Q0) < TC_TAIL_K ? (Q1) - (Q0) : TC_TAIL_K;                        \
  _Pragma("unroll 1") for (int q = (Q0); q < (Q1) - kTail; ++q) {                                 \
    _Pragma("unroll") for (int r = 0; r < 5; ++r) tc::tmem_ld4(TC_QUAD_ADDR(q, r), d[0][r]);      \
    tc::tmem_ld_wait();                                                                           \
    TC_QUAD_MATH(G, q, d[0])                                                                      \
    if (q == (Q0) + TC_CFREE_Q) { CFREE }                                                         \
    TC_G##G##_SWITCH(q, P);                                                                       \
  }                                                                                               \
  _Pragma("unroll") for (int k = 0; k < kTail; ++k)                                               \
    _Pragma("unroll") for (int r = 0; r < 5; ++r) tc::tmem_ld4(TC_QUAD_ADDR((Q1) - kTail + k, r), d[k][r]); \
  tc::tmem_ld_wait();                                                                             \
  tc::tc_fence_before();                                                                          \
  tc::mbar_arrive(&ctl->acc_empty);                                                               \
  if (lane == 0 && wq == 0) { TC_STAMP(2 + (G), n, 2) }                                           \
  _Pragma("unroll") for (int k = 0; k < kTail; ++k) {                                             \
    TC_QUAD_MATH(G, (Q1) - kTail + k, d[k])                                                       \
    if ((Q1) - kTail + k == (Q0)) { CFREE }                                                       \
    TC_G##G##_SWITCH((Q1) - kTail + k, P);                                                        \
  }
#else
      // straight-line code, software-pipelined: the next quad's five TMEM loads are in flight while the
      // current quad's butterflies and mel FFMAs run.  Measured (r2d): the groups' read phase shrinks from
      // 3.3-3.9 k to 2.8-3.2 k cycles, but the producers, on the same sub-partitions, take 2.9 k instead of
      // 1.5 k cycles for their first phase and the launch gets slower (87.9 vs 78.7 us): the sub-partitions'
      // issue slots / FMA pipe are what the roles compete for, not latency.  Kept for experiments.
#define TC_GROUP_LOOP(G, Q0, Q1, CFREE)                                                           \
  float P[11], d[2][5][4];                                                                        \
  _Pragma("unroll") for (int r = 0; r < 5; ++r) tc::tmem_ld4(TC_QUAD_ADDR((Q0), r), d[0][r]);     \
  _Pragma("unroll") for (int q = (Q0); q < (Q1); ++q) {                                           \
    tc::tmem_ld_wait();                                                                           \
    if (q + 1 < (Q1)) {                                                                           \
      _Pragma("unroll") for (int r = 0; r < 5; ++r) tc::tmem_ld4(TC_QUAD_ADDR(q + 1, r), d[(q + 1 - (Q0)) & 1][r]); \
    } else {                                                                                      \
      tc::tc_fence_before();                                                                      \
      tc::mbar_arrive(&ctl->acc_empty);                                                           \
      if (lane == 0 && wq == 0) { TC_STAMP(2 + (G), n, 2) }                                       \
    }                                                                                             \
    if ((G) == 0 && q == 0) tc_radix5_power_special(d[(q - (Q0)) & 1], P);                        \
    else tc_radix5_power_pair(d[(q - (Q0)) & 1], P);                                              \
    if (q == (Q0)) { CFREE }                                                                      \
    TC_G##G##_SWITCH(q, P);                                                                       \
  }
#endif
      if (group == 0) {
        TC_G0_DECL
        TC_GROUP_LOOP(0, 0, TC_SPLIT_A, )
        if (lane == 0 && wq == 0) { TC_STAMP(2, n, 3) }
        if (C == 1) tc::named_bar_sync(bar_free, kFreeCount);
        TC_G0_PHASE1;
        tc::named_bar_arrive(TC_NGROUPS == 3 ? bar_ab : bar_bc, 64);
#if TC_NGROUPS == 3
      } else if (group == 1) {
        TC_G1_DECL
        TC_GROUP_LOOP(1, TC_SPLIT_A, TC_SPLIT_B, )
        if (lane == 0 && wq == 0) { TC_STAMP(3, n, 3) }
        if (C == 1) tc::named_bar_sync(bar_free, kFreeCount);
        TC_G1_PHASE1;
        if (lane == 0 && wq == 0) { TC_STAMP(3, n, 4) }
        tc::named_bar_sync(bar_ab, 64);
        if (lane == 0 && wq == 0) { TC_STAMP(3, n, 6) }
        { TC_G1_PHASE2; }
        tc::named_bar_arrive(bar_bc, 64);
#endif
      } else {
        TC_G2_DECL
        TC_GROUP_LOOP(2, TC_SPLIT_B, 16,
                      if (C == 1) {
                        if (issuer) tc::bulk_wait_read(); /* async-groups belong to the thread that committed them */
                        __syncwarp();
                        tc::named_bar_arrive(bar_free, kFreeCount);
                      })
        if (lane == 0 && wq == 0) { TC_STAMP(4, n, 3) }
        TC_G2_PHASE1;
        if (lane == 0 && wq == 0) { TC_STAMP(4, n, 4) }
        tc::named_bar_sync(bar_bc, 64);
        if (lane == 0 && wq == 0) { TC_STAMP(4, n, 6) }
        { TC_G2_PHASE2; }
        if (kMask && !kRegMask) tc_mask_row(out_row, mask_entry, p.n_masks_f, p.n_masks_t, t_row, t_row >= 0 && t_row < n_fr);
      }
#undef TC_GROUP_LOOP
#undef TC_QUAD_ADDR
#undef TC_QUAD_MATH
      if (lane == 0 && wq == 0) { TC_STAMP(2 + group, n, 5) }
      ++n;

      if (C == 1) {
        // ---- group C stores the trio's 32 staged rows with ONE TMA tensor store: the box is
        //      [32 rows x 84 floats] (the padded staging pitch) over the [B, t_max, 80] output, so
        //      the 4 pad floats of every row and the rows >= t_max fall outside the tensor and are
        //      not written.  No CTA-wide copy-out phase. ----
        if (group == 2) {
          const int wq_u = __shfl_sync(0xffffffffu, wq, 0);  // provably warp-uniform
          const int t_warp = t_first + 32 * wq_u;
          if (t_warp < p.t_max) {
            if (t_row >= n_fr) tc_zero_row(out_row);  // padding is +0.0 (padded_batch, train.py:192-197)
            tc::fence_async_smem();
            __syncwarp();
            if (tc::elect_one()) {
              tc::tma_store_3d(&out_map, sOut_s + 4 * kTcOutPitch * 32 * wq_u, 0, t_warp, b);
              tc::bulk_commit();
              issuer = true;
            }
          }
          if (lane == 0 && wq == 0) { TC_STAMP(4, n - 1, 7) }
        }
      } else {
        tc::named_bar_sync(14, kTcEpilogueThreads);  // staged (and masked) rows visible to the 12 epilogue warps
        // ---- copy-out with delta / delta-delta: staged tile -> HBM.  A work unit is a PAIR of float4 groups
        //      (8 mel bins) over a RUN of four consecutive rows: per row its 24 outputs are 96 contiguous bytes =
        //      three whole 32-byte sectors, written with three 256-bit stores, and the two rows before it stay
        //      in registers - every staged value is read once per run (1.5 times on average) instead of three
        //      times.  126 rows = 32 runs x 10 pairs = 320 units over the 384 epilogue threads; the next row's
        //      two shared-memory loads are issued before the current row's arithmetic and stores (a
        //      shared-memory round trip costs >100 cycles while the MMAs stream their operands).
        //      (First version: one (row, pair) item at a time, six loads per item: 135.1 us per config-2 + delta
        //      batch.) ----
        float* out_b = p.out + static_cast<int64_t>(b) * p.t_max * 240;
        for (int u = etid; u < 320; u += kTcEpilogueThreads) {
          const int q = u / 10, pr = u - 10 * q;
          const int i0 = tp.halo + 4 * q;                    // first staged row of the run (halo = 2: rows 2..127)
          const int t0 = t_first + 4 * q;                    // its frame
          const float* col = sOut + 8 * pr;
          float* dst = out_b + (static_cast<int64_t>(t0) * 80 + 8 * pr) * 3;
          const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
          float4 w2[2], w1[2], w0[2], nx[2];
          // staged row i holds frame t_first - halo + i: real frames only, everything else reads as zero
#define TC_CO_ROW(V, I)                                                                                \
  {                                                                                                    \
    const int tt = t_first - tp.halo + (I);                                                            \
    if (tt >= 0 && tt < n_fr) {                                                                        \
      V[0] = *reinterpret_cast<const float4*>(col + (I) * kTcOutPitch);                                \
      V[1] = *reinterpret_cast<const float4*>(col + (I) * kTcOutPitch + 4);                            \
    } else {                                                                                           \
      V[0] = z4;                                                                                       \
      V[1] = z4;                                                                                       \
    }                                                                                                  \
  }
          TC_CO_ROW(w2, i0 - 2)
          TC_CO_ROW(w1, i0 - 1)
          TC_CO_ROW(w0, i0)
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int i = i0 + c, t = t0 + c;
            if (i >= kTcTileM || t >= p.t_max) break;
            if (c < 3 && i + 1 < kTcTileM) TC_CO_ROW(nx, i + 1)
            // (x, x - x1, (x - x1) - (x1 - x2)) per bin, interleaved [.., m, 3]; padding frames are +0.0
            const bool live = t < n_fr;
            const float x0[8] = {w0[0].x, w0[0].y, w0[0].z, w0[0].w, w0[1].x, w0[1].y, w0[1].z, w0[1].w};
            const float x1[8] = {w1[0].x, w1[0].y, w1[0].z, w1[0].w, w1[1].x, w1[1].y, w1[1].z, w1[1].w};
            const float x2[8] = {w2[0].x, w2[0].y, w2[0].z, w2[0].w, w2[1].x, w2[1].y, w2[1].z, w2[1].w};
            float o[24];
#pragma unroll
            for (int m = 0; m < 8; ++m) {
              const float dd = x0[m] - x1[m], e = x1[m] - x2[m];
              o[3 * m] = live ? x0[m] : 0.0f;
              o[3 * m + 1] = live ? dd : 0.0f;
              o[3 * m + 2] = live ? dd - e : 0.0f;
            }
            tc::st_global_v8(dst + 240 * c, o);
            tc::st_global_v8(dst + 240 * c + 8, o + 8);
            tc::st_global_v8(dst + 240 * c + 16, o + 16);
            w2[0] = w1[0]; w2[1] = w1[1];
            w1[0] = w0[0]; w1[1] = w0[1];
            w0[0] = nx[0]; w0[1] = nx[1];
          }
#undef TC_CO_ROW
        }
        tc::named_bar_sync(14, kTcEpilogueThreads);  // the staging tile and the hand-over zone may be overwritten
      }
    }
    if (C == 1 && issuer) tc::bulk_wait_all();
  }

  // ---- the rows behind every utterance's last active tile: +0.0 like padded_batch
  //      (run/train.py:192-197).  32-row blocks, split evenly over the producer and issuer warps of
  //      the grid: they are done about one tile before the epilogue, which is left alone. ----
  if (warp < 8) {
    tc_walk_begin<true>(swk, p.lengths, batch_eff, t_max32, tp.tout, blockIdx.x * 8u + static_cast<uint32_t>(warp),
                        gridDim.x * 8u, lane, nullptr, nullptr);
    TcWalk z = *swk;
    for (; z.left > 0; tc_walk_step<true>(z, p.lengths, batch_eff, t_max32, tp.tout)) {
      const int32_t r0 = tc_nact(z.len, t_max32, tp.tout) * tp.tout + 32 * z.j;
      const int32_t rows = t_max32 - r0 < 32 ? t_max32 - r0 : 32;
      float4* dst = reinterpret_cast<float4*>(p.out + (static_cast<int64_t>(z.b) * p.t_max + r0) * (80 * C));
      for (int q = lane; q < rows * 20 * C; q += 32) dst[q] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
  }

  if (timed_out && p.status) atomicOr(p.status, 0x100);
  tc::tc_fence_before();
  __syncthreads();
  if (tid == 0) { TC_STAMP(1, 31, 3) }
  if (warp == kTcMmaWarp) tc::tmem_dealloc(tmem_base, 512);
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency).
typedef CUresult (*TcEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline TcEncodeTiledFn tc_encode_fn() {
  static TcEncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<TcEncodeTiledFn>(f);
  }();
  return fn;
}

// SpecAugment without a caller-supplied table: the kernel draws the tables of a CTA's tiles itself when
// they fit its shared-memory table (<= 8 masks, <= 128 tiles per CTA); otherwise the API layer runs
// specaug_draw_from_lengths_kernel as a pre-pass and passes the table.
inline bool tc_draws_in_kernel(const KernelParams& kp, int sm_count) {
  const int halo = kp.C == 3 ? 2 : 0, tout = kTcTileM - halo;
  const int64_t total = ((kp.t_max + tout - 1) / tout) * kp.batch;  // (an upper bound per launch)
  const int64_t grid = total < sm_count ? total : sm_count;
  return 2 * (kp.n_masks_f + kp.n_masks_t) <= kTcMaskEntries && grid > 0 && (total + grid - 1) / grid <= kTcMaskTiles;
}

// One launch over utterances [0, kp.batch) of kp.
inline int tc_launch_one(const TcPlan& tc, const KernelParams& kp, int sm_count, cudaStream_t st) {
  TcParams tp;
  tp.b_image = tc.d_b;
  tp.log_eps = tc.log_eps;
  tp.halo = kp.C == 3 ? 2 : 0;
  tp.tout = kTcTileM - tp.halo;
  tp.tiles_per_utt = static_cast<int32_t>((kp.t_max + tp.tout - 1) / tp.tout);
  const int64_t total = static_cast<int64_t>(tp.tiles_per_utt) * kp.batch;
  if (total <= 0 || total > 0x7FFFFFFF) return total == 0 ? 0 : 1;
  tp.total_tiles = static_cast<int32_t>(total);
  const int grid = total < sm_count ? static_cast<int>(total) : sm_count;
  alignas(64) CUtensorMap map;
  std::memset(&map, 0, sizeof(map));
  cudaLaunchConfig_t lc;
  std::memset(&lc, 0, sizeof(lc));
  lc.gridDim = dim3(grid);
  lc.blockDim = dim3(kTcThreads);
  lc.dynamicSmemBytes = kTcSmemTotal;
  lc.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  lc.attrs = attr;
  lc.numAttrs = 1;
  const bool masks = kp.n_masks_f + kp.n_masks_t > 0, f32 = kp.pcm_is_f32 != 0;
  if (kp.C == 3) {
    const cudaError_t e =
        f32 ? (masks ? cudaLaunchKernelEx(&lc, tc_frontend_kernel<3, true, true>, kp, tp, map)
                     : cudaLaunchKernelEx(&lc, tc_frontend_kernel<3, false, true>, kp, tp, map))
            : (masks ? cudaLaunchKernelEx(&lc, tc_frontend_kernel<3, true, false>, kp, tp, map)
                     : cudaLaunchKernelEx(&lc, tc_frontend_kernel<3, false, false>, kp, tp, map));
    if (e != cudaSuccess) return 1;
  } else {
    // output viewed as a 3-D tensor [B][t_max][80] f32; box = [1][32][84]
    TcEncodeTiledFn enc = tc_encode_fn();
    if (!enc) return 2;
    const cuuint64_t dims[3] = {80, static_cast<cuuint64_t>(kp.t_max), static_cast<cuuint64_t>(kp.batch)};
    const cuuint64_t strides[2] = {320, static_cast<cuuint64_t>(kp.t_max) * 320};
    const cuuint32_t box[3] = {static_cast<cuuint32_t>(kTcOutPitch), 32, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, kp.out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return 2;
    const cudaError_t e =
        f32 ? (masks ? cudaLaunchKernelEx(&lc, tc_frontend_kernel<1, true, true>, kp, tp, map)
                     : cudaLaunchKernelEx(&lc, tc_frontend_kernel<1, false, true>, kp, tp, map))
            : (masks ? cudaLaunchKernelEx(&lc, tc_frontend_kernel<1, true, false>, kp, tp, map)
                     : cudaLaunchKernelEx(&lc, tc_frontend_kernel<1, false, false>, kp, tp, map));
    if (e != cudaSuccess) return 1;
  }
  return cudaGetLastError() == cudaSuccess ? 0 : 1;
}

// Tiles one launch may cover: every CTA keeps its tile list in shared memory (kTcListTiles entries).
// B200MEL_TC_MAX_TILES lowers it (tests of the chunked path; read once).
inline int64_t tc_max_tiles_per_launch(int sm_count) {
  static const int64_t forced = [] {
    const char* e = std::getenv("B200MEL_TC_MAX_TILES");
    return e ? std::atoll(e) : 0LL;
  }();
  const int64_t cap = static_cast<int64_t>(kTcListTiles) * sm_count;
  return forced > 0 && forced < cap ? forced : cap;
}

// Host launcher.  Returns 0 on success, 2 when the tensor map cannot be encoded, 3 when one utterance
// alone has more tiles than a launch can cover.  A batch with more tiles than that is processed in
// consecutive utterance ranges, one launch each (same stream, so they overlap like any two steps).
// With SpecAugment enabled and no kp.mask_table the kernel draws the tables itself: the caller checks
// tc_draws_in_kernel() first.
inline int tc_launch(const TcPlan& tc, const KernelParams& kp, int sm_count, cudaStream_t st, int* n_launches) {
  const int halo = kp.C == 3 ? 2 : 0, tout = kTcTileM - halo;
  const int64_t tiles_per_utt = (kp.t_max + tout - 1) / tout;
  if (n_launches) *n_launches = 0;
  if (tiles_per_utt <= 0 || kp.batch <= 0) return 0;
  const int64_t per_launch = tc_max_tiles_per_launch(sm_count) / tiles_per_utt;
  if (per_launch < 1) return 3;
  const int n_masks = kp.n_masks_f + kp.n_masks_t;
  for (int64_t b0 = 0; b0 < kp.batch; b0 += per_launch) {
    KernelParams sub = kp;
    sub.batch = static_cast<int32_t>(kp.batch - b0 < per_launch ? kp.batch - b0 : per_launch);
    sub.lengths = kp.lengths + b0;
    if (kp.offsets) {
      sub.offsets = kp.offsets + b0;  // (element offsets into the same buffer)
    } else {
      sub.pcm = static_cast<const char*>(kp.pcm) + b0 * kp.row_stride * (kp.pcm_is_f32 ? 4 : 2);
    }
    sub.out = kp.out + b0 * kp.t_max * kp.F * kp.C;
    if (kp.n_frames) sub.n_frames = kp.n_frames + b0;
    if (kp.mask_table) sub.mask_table = kp.mask_table + b0 * 2 * n_masks;
    sub.first_utt = kp.first_utt + static_cast<uint64_t>(b0);
    const int rc = tc_launch_one(tc, sub, sm_count, st);
    if (rc != 0) return rc;
    if (n_launches) ++*n_launches;
  }
  return 0;
}

}  // namespace b200mel
