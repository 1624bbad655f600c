// RowParallelLinear (o_proj / down_proj) GEMM + tensor-parallel all-reduce in ONE launch
// per rank, over NVLink peer memory.  Declarations and reference citations: include/nvllm_attn.h (6).
//
// reference: nanovllm/layers/linear.py:185-190 — y = F.linear(x, W_r); dist.all_reduce(y)   (two launches
// plus an NCCL kernel per call; 2 calls per decoder layer: qwen3.py:118 and qwen3.py:147).
//
// One CTA per (feature tile, k split, token block); the ksplit CTAs of a tile form one thread-block CLUSTER:
//   1. partial[tok, f] = sum_k x[tok, k] * W[f, k] over its k range: mma.sync m16n8k16 bf16 -> fp32, operands
//      streamed with 16-byte cp.async.cg through a 4/6/8-stage swizzled shared-memory ring, as deep as two CTAs per
//      SM allow (the op is bound by streaming W once from HBM and by latency: N tokens <= 512 is N flop/byte,
//      under the tensor ridge; ~100 KB in flight per SM hides the HBM latency);
//   2. the k splits are summed INSIDE the cluster through distributed shared memory (CTA ks owns a slice of the
//      tile's token rows and reads that slice from every CTA of the cluster, fixed order), so one fp32 partial per
//      rank and tile crosses NVLink, not ksplit of them; the slice is PUSHED with 16-byte stores into the inbox of
//      every rank ("one-shot", mode 0) or of the tile's owner rank only ("two-shot", mode 1), then published with
//      one st.release.sys flag per target;
//   3. the CTA waits (ld.acquire.sys) for the `world` flags of its row slice, sums the ranks' partials in rank
//      order in fp32 (bit-identical on every rank), adds the bias, rounds to bf16 once and writes y (mode 0), or
//      writes the reduced bf16 slice into every rank's result area, publishes a second flag, and every rank copies
//      its slice into y (mode 1).
//   Decode-sized pushes (<= 1 MiB per rank) use mode 3 instead of 0: the slice travels as 8-byte {payload, call
//   epoch} words (two per 16-byte volatile store, the NCCL "LL" layout) and the receiving thread polls the words of
//   its own elements until every epoch matches -- no fence, no separate flag, no CTA barrier on either side.
//   Hardware assumption, the one NCCL's LL protocol makes with the very same instructions (its storeLL / readLL are
//   st/ld.volatile.global.v4.u32 over {data1, flag1, data2, flag2}): each naturally aligned 8-byte half of a 16-byte
//   vector access is single-copy atomic over NVLink and in L2, so a half whose epoch matches carries that epoch's
//   payload.  tests/test_gpu_rowparallel_worlds.py::test_ll_exchange_stress hammers it (worlds 2, 4, 8).
//   These words live in a region of their own -- anchored at the END of each parity, while the fp32 inboxes and the
//   two-shot result area of every other plan grow from its start, and nvllm_rowparallel_comm_layout sizes the buffer
//   so the two never meet -- zeroed at allocation, so nothing but an older epoch can ever sit there.
// Token batches from 64-128 up on big shards run step 1 on the 5th-generation tensor cores instead ("swap-AB" tcgen05
// tile, template flag TC): A = 128 output features of W (UMMA M = 128), B = 128 token rows of x (UMMA N), both K-major
// with 128-byte swizzle.  Warp 0 is the TMA producer (one elected lane: two cp.async.bulk.tensor.2d per 64-wide k stage
// into a 6-stage, 192 KB ring, mbarrier complete_tx; token rows past n are out of the tensor map's bounds and arrive as
// zeros), warp 1 the MMA issuer (one elected lane: four tcgen05.mma.kind::f16 per stage, tcgen05.commit hands the
// stage back), the fp32 accumulator [128 features x 128 tokens] lives in TMEM, and all four warps read it with
// tcgen05.ld (lane = feature) and write the transposed tile into the same staged [token][feature] layout steps 2-3
// consume -- so the cluster reduction and all three exchanges are shared with the mma.sync tile.  One CTA per SM.
// The first cut fed the UMMA from the cp.async ring (same swizzled layout): LSU request tracking capped it near
// 5 TB/s of L2+HBM traffic however deep the ring (profiles/r02_rowparallel_local_tile128_tcgen05_cpasync_v1.jsonl,
// r02_rowparallel_tcgen05_n256_ncu_full_metrics.csv); bulk copies took the 50 MB shard at 256 tokens from 37 to 24 us.
// Flags carry a call epoch kept in device memory (graph replays advance it), data areas are double-buffered by
// epoch parity; nothing is reset, nothing is allocated per call.  A pushed store never waits on anything, so a
// waiting CTA only depends on CTAs that cannot block before publishing; the launcher additionally keeps the grid
// within one co-resident wave of clusters.
// Why two parities suffice: a rank finishes call e only after it has received, for every tile it reduces, what each
// peer pushed in call e -- so every peer has STARTED call e and therefore (stream order) finished call e-1.  A rank
// that is already pushing call e+1 (same parity as e-1) can thus only meet peers that are done reading e-1.
#include <cooperative_groups.h>

#include <cstdlib>
#include <cstring>
#include <unordered_map>

#include "common.cuh"
#include "tcgen05.cuh"

namespace nvllm {
namespace {

constexpr int RP_THREADS = 128;
constexpr int RP_BK_MIN = 64;      // k elements per pipeline stage: 64 (128-byte rows) or 128 (256-byte rows)
constexpr int RP_TOKB = 128;       // tokens per CTA
constexpr int RP_MAX_WORLD = 8;
constexpr int RP_MAX_KSPLIT = 8;
constexpr int RP_MAX_SLOTS = RP_MAX_WORLD * RP_MAX_KSPLIT;
constexpr int RP_MAX_TILES = 1024;
constexpr int RP_MAX_TOKENS = 512;
constexpr int RP_PLAN_CAP = 296;    // 2 CTAs x 148 SMs
constexpr size_t RP_HDR_BYTES = 4096;                                       // control words
constexpr size_t RP_FLAG1_BYTES = size_t(RP_MAX_SLOTS) * RP_MAX_TILES * 4;  // [slot][tile] partial arrived
constexpr size_t RP_FLAG2_BYTES = size_t(RP_MAX_KSPLIT) * RP_MAX_TILES * 4; // [ks][tile] reduced slice arrived
constexpr size_t RP_DATA_OFF = RP_HDR_BYTES + RP_FLAG1_BYTES + RP_FLAG2_BYTES;

struct RpControl {
  uint32_t epoch;   // calls completed on this rank
  uint32_t done;    // CTAs of the running call that have finished (self-resetting)
  uint32_t error;   // 0, or 1 + the tile whose wait timed out
};

struct RpParams {
  const __nv_bfloat16* x;
  const __nv_bfloat16* w;
  const __nv_bfloat16* bias;
  __nv_bfloat16* y;
  int64_t x_stride, w_stride, y_stride;
  uint8_t* comm[RP_MAX_WORLD];
  int rank, world, n_tokens, ksplit, kc, h_tiles, n_tiles, mode;
  unsigned long long par_bytes;   // bytes of one parity's data area
  unsigned long long res_off;     // offset of the result area inside a parity (mode 1)
  unsigned long long ll_off;      // offset of the flag-in-data inbox inside a parity (mode 3)
  unsigned long long timeout_ns;
  uint32_t* host_err;             // mapped pinned host word: a timeout is visible to the host without a sync
  // tcgen05 tile only: TMA descriptors of W [out, k/tp] and x [n, k/tp] (box 128 rows x 64 k, 128-byte swizzle)
  alignas(64) CUtensorMap tm_w;
  alignas(64) CUtensorMap tm_x;
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;\n" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_v4(void* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_v2u(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.global.v2.u32 [%0], {%1,%2};\n" ::"l"(p), "r"(a), "r"(b) : "memory");
}

__device__ __forceinline__ void st_volatile_v4u(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_v4u(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// Spin until *flag has reached `epoch` (wrap-safe); on timeout record the tile and give up.
__device__ __forceinline__ void latch_timeout(RpControl* ctrl, uint32_t* host_err, int tile) {
  atomicCAS(&ctrl->error, 0u, 1u + static_cast<uint32_t>(tile));
  if (host_err) {
    *reinterpret_cast<volatile uint32_t*>(host_err) = 1u + static_cast<uint32_t>(tile);
    __threadfence_system();
  }
}

__device__ __forceinline__ void wait_flag(const uint32_t* flag, uint32_t epoch, RpControl* ctrl, int tile,
                                          unsigned long long timeout_ns, uint32_t* host_err) {
  if (static_cast<int32_t>(ld_acquire_sys(flag) - epoch) >= 0) return;
  const unsigned long long t0 = globaltimer_ns();
  while (static_cast<int32_t>(ld_acquire_sys(flag) - epoch) < 0) {
    __nanosleep(32);
    if (globaltimer_ns() - t0 > timeout_ns) {
      latch_timeout(ctrl, host_err, tile);
      return;
    }
  }
}

__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release;\n" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire;\n" ::: "memory"); }

// Ring depth: as deep as lets two CTAs share an SM's shared memory (RP_SMEM_BUDGET each).
constexpr size_t RP_SMEM_BUDGET = size_t(100) * 1024;
__host__ __device__ constexpr int rp_stages(int bh, int mt, int bk) {
  const size_t stage = size_t(bh + mt * 16) * bk * 2;
  return 8 * stage <= RP_SMEM_BUDGET ? 8 : 6 * stage <= RP_SMEM_BUDGET ? 6 : 4;
}

// BH output features x (MT*16) token rows per CTA, BK k elements per pipeline stage.
// tcgen05 tile: ONE CTA per SM with a 192 KB ring (6 stages of 16 KB W + 16 KB x), filled by TMA.  ncu of the first cut (3 stages, two
// CTAs per SM by shared memory, profiles/r02_rowparallel_tcgen05_n256_ncu_full_metrics.csv): the cluster capacity
// halved the k split, 128 CTAs of 4 warps ran one per SM with 64 KB in flight each and DRAM sat at 1.35 TB/s -- the
// tile is bound by bytes in flight per SM, not by the tensor pipe (19 % active), so depth goes to the ring.
__host__ __device__ constexpr int rp_tc_stages(int /*mt*/) { return 6; }
constexpr int RP_TC_PLAN_CAP = 148;   // co-resident CTAs of the tcgen05 tile (one per SM)

template <int BH, int BK, int MT, bool TC = false>
__global__ void __launch_bounds__(RP_THREADS) rowparallel_allreduce_kernel(const __grid_constant__ RpParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  // UMMA descriptors with 128-byte swizzle want 1024-byte aligned tiles: the TC launcher over-allocates 1 KiB
  uint8_t* const smem = TC ? reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023))
                           : smem_raw;
  static_assert(!TC || (BH == 128 && BK == 64 && MT == 8), "tcgen05 tile: 128 features x 128 tokens x k64 per stage");
  constexpr int STAGES = TC ? rp_tc_stages(MT) : rp_stages(BH, MT, BK);
  constexpr int NT = MT * 16;             // staged token rows (rows past the last token are zero-filled)
  // warp grid: 4 warps split the BH features (decode tiles), or 2 x 2 over (token tiles, features) for 64-feature
  // tiles with >= 32 tokens -- each A fragment then feeds 4 MMAs and each B fragment MT/2, which halves the
  // shared-memory reads per MMA
  constexpr int WM = (BH == 64 && MT >= 2) ? 2 : 1, WN = 4 / WM;
  constexpr int MW = MT / WM;             // 16-token m-tiles per warp
  constexpr int FW = BH / WN;             // features per warp
  constexpr int NJ = FW / 8;              // 8-feature n-tiles per warp
  constexpr int F4 = BH / 4;              // float4 chunks per tile row
  constexpr int EP_STRIDE = BH + 8;       // floats per staged epilogue row (padding breaks bank conflicts)
  constexpr size_t TILE_BYTES = size_t(RP_TOKB) * BH * 4;   // inbox footprint of one (slot, tile)
  constexpr size_t RES_TILE_BYTES = size_t(RP_TOKB) * BH * 2;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ks = blockIdx.x % p.ksplit;
  const int tile = blockIdx.x / p.ksplit;
  const int ht = tile % p.h_tiles;
  const int tok0 = (tile / p.h_tiles) * RP_TOKB;
  const int nt = min(min(RP_TOKB, NT), p.n_tokens - tok0);

  uint8_t* const local = p.comm[p.rank];
  RpControl* const ctrl = reinterpret_cast<RpControl*>(local);
  __shared__ uint32_t s_epoch;
  if (tid == 0) s_epoch = *reinterpret_cast<volatile uint32_t*>(&ctrl->epoch) + 1u;

  // ---------------- 1. partial GEMM over k in [ks*kc, (ks+1)*kc) ----------------
  constexpr int ROWB = BK * 2;            // bytes per staged row
  constexpr int CPR = BK / 8;             // 16-byte chunks per staged row; chunks are XOR-swizzled by (row & 7)
  constexpr int RSTEP = RP_THREADS / CPR; // a thread keeps its chunk column and walks rows r0, r0+RSTEP, ... (RSTEP % 8 == 0)
  constexpr uint32_t W_STAGE = BH * ROWB, STAGE_BYTES = (BH + NT) * ROWB;
  const uint32_t sbase = smem_u32(smem);
  const int nk = p.kc / BK;
  // per-thread copy assignments, fixed for the whole k loop: only the k offset moves
  const int r0 = tid / CPR, ch0 = tid % CPR;
  const uint32_t dst0 = r0 * ROWB + ((ch0 ^ (r0 & 7)) << 4);
  const __nv_bfloat16* const wsrc0 = p.w + (static_cast<int64_t>(ht) * BH + r0) * p.w_stride +
                                     static_cast<int64_t>(ks) * p.kc + ch0 * 8;
  const __nv_bfloat16* const xbase = p.x + static_cast<int64_t>(tok0) * p.x_stride + static_cast<int64_t>(ks) * p.kc + ch0 * 8;
  const __nv_bfloat16* const xsrc0 = xbase + static_cast<int64_t>(r0) * p.x_stride;
  const int64_t w_step = static_cast<int64_t>(RSTEP) * p.w_stride, x_step = static_cast<int64_t>(RSTEP) * p.x_stride;

  auto load_stage = [&](int i) {
    const uint32_t wb = sbase + (i % STAGES) * STAGE_BYTES + dst0, xb = wb + W_STAGE;
    const __nv_bfloat16* ws = wsrc0 + i * BK;
    const __nv_bfloat16* xs = xsrc0 + i * BK;
#pragma unroll
    for (int j = 0; j < BH / RSTEP; ++j) cp_async16(wb + j * (RSTEP * ROWB), ws + j * w_step, 16);
#pragma unroll
    for (int j = 0; j < NT / RSTEP; ++j) {
      const bool ok = r0 + j * RSTEP < nt;
      cp_async16(xb + j * (RSTEP * ROWB), ok ? xs + j * x_step : xbase, ok ? 16 : 0);
    }
  };

  float* const ep = reinterpret_cast<float*>(smem);
  if constexpr (TC) {
    // ---- tcgen05 path: TMA ring -> UMMA (one elected thread each) -> TMEM accumulator -> staged fp32 tile
    __shared__ __align__(8) uint64_t tc_full[STAGES], tc_empty[STAGES], tc_done;
    __shared__ uint32_t tc_tmem_slot;
    constexpr uint32_t TMEM_COLS = 128;
    constexpr uint32_t IDESC = tc::make_idesc_bf16(BH, NT, /*b_mn_major=*/false);
    if (tid == 0) {
#pragma unroll
      for (int s = 0; s < STAGES; ++s) {
        tc::mbar_init(smem_u32(&tc_full[s]), 1);
        tc::mbar_init(smem_u32(&tc_empty[s]), 1);
      }
      tc::mbar_init(smem_u32(&tc_done), 1);
      tc::fence_barrier_init();
      tc::tma_prefetch_desc(&p.tm_w);
      tc::tma_prefetch_desc(&p.tm_x);
    }
    if (warp == 0) {
      tc::tmem_alloc(smem_u32(&tc_tmem_slot), TMEM_COLS);
      tc::tmem_relinquish();
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tc_tmem_slot);
    const int k_begin = ks * p.kc;
    if (warp == 0) {
      // producer: one lane feeds the ring with two bulk tensor copies per stage (W tile + x tile; rows of x past the
      // last token are out of bounds of the tensor map and arrive as zeros).  Bulk copies need no LSU request
      // tracking, so the whole 192 KB ring is in flight (the cp.async loader saturated near 5 TB/s of L2+HBM traffic).
      if (tc::elect_one()) {
        for (int i = 0; i < nk; ++i) {
          const int s = i % STAGES;
          if (i >= STAGES) tc::mbar_wait(smem_u32(&tc_empty[s]), static_cast<uint32_t>((i / STAGES - 1) & 1));
          const uint32_t full = smem_u32(&tc_full[s]);
          const uint32_t dst = sbase + s * STAGE_BYTES;
          tc::mbar_arrive_expect_tx(full, STAGE_BYTES);
          tc::tma_load_2d(dst, &p.tm_w, k_begin + i * BK, ht * BH, full);
          tc::tma_load_2d(dst + W_STAGE, &p.tm_x, k_begin + i * BK, tok0, full);
        }
      }
    } else if (warp == 1) {
      if (tc::elect_one()) {
        for (int i = 0; i < nk; ++i) {
          const int s = i % STAGES;
          tc::mbar_wait(smem_u32(&tc_full[s]), static_cast<uint32_t>((i / STAGES) & 1));
          tc::fence_after_sync();
          const uint32_t sb = sbase + s * STAGE_BYTES;
          const uint64_t da = tc::make_smem_desc(sb, 16, 1024), db = tc::make_smem_desc(sb + W_STAGE, 16, 1024);
#pragma unroll
          for (int ks16 = 0; ks16 < BK / 16; ++ks16)   // +32 bytes per k16 step inside the swizzle atom
            tc::mma_ss(tmem, da + 2 * ks16, db + 2 * ks16, IDESC, (i > 0 || ks16 > 0) ? 1u : 0u);
          tc::mma_commit(smem_u32(&tc_empty[s]));   // the stage may be refilled once these MMAs have retired
        }
        tc::mma_commit(smem_u32(&tc_done));
      }
    }
    __syncwarp();
    tc::mbar_wait(smem_u32(&tc_done), 0);   // every MMA retired: the accumulator is final and the ring is free
    tc::fence_after_sync();
    __syncthreads();
    // epilogue: warp w owns TMEM lanes (= features) 32w..32w+31; 32 token columns per tcgen05.ld
#pragma unroll
    for (int c0 = 0; c0 < NT; c0 += 32) {
      uint32_t r[32];
      tc::tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + static_cast<uint32_t>(c0), r);
      tc::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) ep[(c0 + j) * EP_STRIDE + warp * 32 + lane] = __uint_as_float(r[j]);
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, TMEM_COLS);
  } else {
  const int wm = warp / WN, wn = warp % WN;
  float acc[MW][NJ][4];
#pragma unroll
  for (int m = 0; m < MW; ++m)
#pragma unroll
    for (int j = 0; j < NJ; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[m][j][e] = 0.f;

  // ldmatrix lane addresses inside a stage (the k16 step only flips chunk bits: XOR with a constant)
  const int brow = wn * FW + (lane & 7);
  const uint32_t boff = brow * ROWB, bsw = ((lane >> 3) ^ (brow & 7)) << 4;          // chunk kk2*4 + (lane>>3)
  const int arow = lane & 15;
  const uint32_t aoff = W_STAGE + (wm * MW * 16 + arow) * ROWB, asw = ((lane >> 4) ^ (arow & 7)) << 4;  // chunk kk*2 + (lane>>4)

#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < nk) load_stage(s);
    cp_async_commit();
  }
  for (int i = 0; i < nk; ++i) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();   // stage i landed for every thread; stage i-1 is no longer being read
    if (i + STAGES - 1 < nk) load_stage(i + STAGES - 1);
    cp_async_commit();
    const uint32_t sb = sbase + (i % STAGES) * STAGE_BYTES;
#pragma unroll
    for (int kk2 = 0; kk2 < BK / 32; ++kk2) {
      uint32_t b[NJ][4];   // {b0,b1} of k16 step 2*kk2, {b0,b1} of step 2*kk2+1
#pragma unroll
      for (int j = 0; j < NJ; ++j)
        ldmatrix_x4(b[j][0], b[j][1], b[j][2], b[j][3], sb + boff + j * (8 * ROWB) + (bsw ^ ((kk2 * 4) << 4)));
#pragma unroll
      for (int kh = 0; kh < 2; ++kh) {
        const int kk = kk2 * 2 + kh;
#pragma unroll
        for (int m = 0; m < MW; ++m) {
          uint32_t a0, a1, a2, a3;
          ldmatrix_x4(a0, a1, a2, a3, sb + aoff + m * (16 * ROWB) + (asw ^ ((kk * 2) << 4)));
#pragma unroll
          for (int j = 0; j < NJ; ++j) mma_bf16_16816(acc[m][j], a0, a1, a2, a3, b[j][2 * kh], b[j][2 * kh + 1]);
        }
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();   // the ring is free: reuse it to stage the fp32 tile [NT][EP_STRIDE]

  {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int m = 0; m < MW; ++m) {
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int col = wn * FW + j * 8 + 2 * t, row = (wm * MW + m) * 16 + g;
        *reinterpret_cast<float2*>(ep + row * EP_STRIDE + col) = make_float2(acc[m][j][0], acc[m][j][1]);
        *reinterpret_cast<float2*>(ep + (row + 8) * EP_STRIDE + col) = make_float2(acc[m][j][2], acc[m][j][3]);
      }
    }
  }
  }   // !TC
  // every CTA of the cluster has staged its k-split partial (release/acquire at cluster scope covers the
  // distributed-shared-memory reads below)
  cluster_arrive();
  cluster_wait();

  // ---------------- 2. sum the k splits of this CTA's row slice, push it, publish it ----------------
  // the ksplit CTAs of a tile share its reduction: CTA ks owns token rows [r_lo, r_hi)
  const uint32_t epoch = s_epoch;
  const size_t par_off = RP_DATA_OFF + static_cast<size_t>(epoch & 1u) * p.par_bytes;
  const int owner = tile % p.world;
  const int rows_per = (nt + p.ksplit - 1) / p.ksplit;
  const int r_lo = min(nt, ks * rows_per), r_hi = min(nt, r_lo + rows_per);
  const bool has_rows = r_hi > r_lo;   // CTA-uniform, and the same on every rank
  const size_t inbox_off = par_off + (static_cast<size_t>(p.rank) * p.n_tiles + tile) * TILE_BYTES;
  const size_t ll_off = par_off + p.ll_off + (static_cast<size_t>(p.rank) * p.n_tiles + tile) * (2 * TILE_BYTES);
  {
    const float* epc[RP_MAX_KSPLIT];
    cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
#pragma unroll
    for (int c = 0; c < RP_MAX_KSPLIT; ++c) epc[c] = c < p.ksplit ? cluster.map_shared_rank(ep, c) : ep;
    uint8_t* const dst1 = p.comm[owner] + inbox_off;
    for (int i = r_lo * F4 + tid; i < r_hi * F4; i += RP_THREADS) {
      const int off = (i / F4) * EP_STRIDE + (i % F4) * 4;
      float4 s = *reinterpret_cast<const float4*>(epc[0] + off);
#pragma unroll
      for (int c = 1; c < RP_MAX_KSPLIT; ++c) {
        if (c < p.ksplit) {
          const float4 v = *reinterpret_cast<const float4*>(epc[c] + off);
          s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
      }
      if (p.mode == 0) {
        for (int d = 0; d < p.world; ++d) st_v4(p.comm[d] + inbox_off + static_cast<size_t>(i) * 16, s);
      } else if (p.mode == 1) {
        st_v4(dst1 + static_cast<size_t>(i) * 16, s);
      } else if (p.mode == 3) {
        // flag-in-data ("LL"): every 8 bytes carry 4 bytes of payload and the call epoch, so the receiver needs
        // neither a separate flag nor a fence on this side -- an 8-byte half is valid as soon as its epoch matches
        const size_t off = ll_off + static_cast<size_t>(i) * 32;
        for (int d = 0; d < p.world; ++d) {
          st_volatile_v4u(p.comm[d] + off, __float_as_uint(s.x), epoch, __float_as_uint(s.y), epoch);
          st_volatile_v4u(p.comm[d] + off + 16, __float_as_uint(s.z), epoch, __float_as_uint(s.w), epoch);
        }
      } else {   // mode 2 (profiling only): this rank's partial straight to y, no exchange
        const int c = (i % F4) * 4;
        if (p.bias) {
          const __nv_bfloat16* bp = p.bias + ht * BH + c;
          s.x += __bfloat162float(bp[0]); s.y += __bfloat162float(bp[1]);
          s.z += __bfloat162float(bp[2]); s.w += __bfloat162float(bp[3]);
        }
        st_v2u(p.y + static_cast<int64_t>(tok0 + i / F4) * p.y_stride + ht * BH + c, pack_bf16x2(s.x, s.y),
               pack_bf16x2(s.z, s.w));
      }
    }
  }
  cluster_arrive();   // this CTA no longer reads its peers' shared memory (waited for before exit)
  const size_t flag_off = RP_HDR_BYTES + (static_cast<size_t>(p.rank * p.ksplit + ks) * RP_MAX_TILES + tile) * 4;
  if (has_rows && p.mode < 2) {
    // the release below is cumulative over the CTA barrier: it orders every thread's pushed stores before the flag
    __syncthreads();
    if (p.mode == 0) {
      if (tid < p.world) st_release_sys(reinterpret_cast<uint32_t*>(p.comm[tid] + flag_off), epoch);
    } else if (tid == 0) {
      st_release_sys(reinterpret_cast<uint32_t*>(p.comm[owner] + flag_off), epoch);
    }
  }

  // ---------------- 3. reduce over ranks ----------------
  const int col0 = ht * BH;
  if (has_rows && (p.mode == 0 || (p.mode == 1 && owner == p.rank))) {
    if (tid < p.world)
      wait_flag(reinterpret_cast<const uint32_t*>(local + RP_HDR_BYTES +
                                                  (static_cast<size_t>(tid * p.ksplit + ks) * RP_MAX_TILES + tile) * 4),
                epoch, ctrl, tile, p.timeout_ns, p.host_err);
    __syncthreads();
    const uint8_t* const in0 = local + par_off + static_cast<size_t>(tile) * TILE_BYTES;
    const size_t slot_stride = static_cast<size_t>(p.n_tiles) * TILE_BYTES;
    for (int i = r_lo * F4 + tid; i < r_hi * F4; i += RP_THREADS) {
      float4 s = __ldcg(reinterpret_cast<const float4*>(in0 + static_cast<size_t>(i) * 16));
      for (int sl = 1; sl < p.world; ++sl) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(in0 + sl * slot_stride + static_cast<size_t>(i) * 16));
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      const int row = i / F4, c = (i % F4) * 4;
      if (p.bias) {
        const __nv_bfloat16* bp = p.bias + col0 + c;
        s.x += __bfloat162float(bp[0]); s.y += __bfloat162float(bp[1]);
        s.z += __bfloat162float(bp[2]); s.w += __bfloat162float(bp[3]);
      }
      const uint32_t lo = pack_bf16x2(s.x, s.y), hi = pack_bf16x2(s.z, s.w);
      if (p.mode == 0) {
        st_v2u(p.y + static_cast<int64_t>(tok0 + row) * p.y_stride + col0 + c, lo, hi);
      } else {
        const size_t off = par_off + p.res_off + static_cast<size_t>(tile) * RES_TILE_BYTES + static_cast<size_t>(i) * 8;
        for (int d = 0; d < p.world; ++d) st_v2u(p.comm[d] + off, lo, hi);
      }
    }
    if (p.mode != 0) {
      __syncthreads();
      if (tid < p.world)
        st_release_sys(reinterpret_cast<uint32_t*>(p.comm[tid] + RP_HDR_BYTES + RP_FLAG1_BYTES +
                                                   (static_cast<size_t>(ks) * RP_MAX_TILES + tile) * 4), epoch);
    }
  }
  if (has_rows && p.mode == 3) {
    const uint8_t* const in0 = local + par_off + p.ll_off + static_cast<size_t>(tile) * (2 * TILE_BYTES);
    const size_t slot_stride = static_cast<size_t>(p.n_tiles) * (2 * TILE_BYTES);
    for (int i = r_lo * F4 + tid; i < r_hi * F4; i += RP_THREADS) {
      uint4 a[RP_MAX_WORLD], b[RP_MAX_WORLD];
      unsigned long long t0 = 0;
      for (int spins = 0;; ++spins) {
        bool ready = true;
#pragma unroll
        for (int sl = 0; sl < RP_MAX_WORLD; ++sl) {
          if (sl < p.world) {
            const uint8_t* src = in0 + sl * slot_stride + static_cast<size_t>(i) * 32;
            a[sl] = ld_volatile_v4u(src);
            b[sl] = ld_volatile_v4u(src + 16);
            ready = ready && a[sl].y == epoch && a[sl].w == epoch && b[sl].y == epoch && b[sl].w == epoch;
          }
        }
        if (ready) break;
        if ((spins & 63) == 63) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (now - t0 > p.timeout_ns) {
            latch_timeout(ctrl, p.host_err, tile);
            break;
          }
        }
      }
      float4 s = make_float4(__uint_as_float(a[0].x), __uint_as_float(a[0].z), __uint_as_float(b[0].x), __uint_as_float(b[0].z));
#pragma unroll
      for (int sl = 1; sl < RP_MAX_WORLD; ++sl) {
        if (sl < p.world) {
          s.x += __uint_as_float(a[sl].x); s.y += __uint_as_float(a[sl].z);
          s.z += __uint_as_float(b[sl].x); s.w += __uint_as_float(b[sl].z);
        }
      }
      const int row = i / F4, c = (i % F4) * 4;
      if (p.bias) {
        const __nv_bfloat16* bp = p.bias + col0 + c;
        s.x += __bfloat162float(bp[0]); s.y += __bfloat162float(bp[1]);
        s.z += __bfloat162float(bp[2]); s.w += __bfloat162float(bp[3]);
      }
      st_v2u(p.y + static_cast<int64_t>(tok0 + row) * p.y_stride + col0 + c, pack_bf16x2(s.x, s.y), pack_bf16x2(s.z, s.w));
    }
  }
  if (has_rows && p.mode == 1) {
    if (tid == 0)
      wait_flag(reinterpret_cast<const uint32_t*>(local + RP_HDR_BYTES + RP_FLAG1_BYTES +
                                                  (static_cast<size_t>(ks) * RP_MAX_TILES + tile) * 4),
                epoch, ctrl, tile, p.timeout_ns, p.host_err);
    __syncthreads();
    const uint8_t* const res = local + par_off + p.res_off + static_cast<size_t>(tile) * RES_TILE_BYTES;
    for (int i = r_lo * F4 + tid; i < r_hi * F4; i += RP_THREADS) {
      const uint2 v = __ldcg(reinterpret_cast<const uint2*>(res + static_cast<size_t>(i) * 8));
      st_v2u(p.y + static_cast<int64_t>(tok0 + i / F4) * p.y_stride + col0 + (i % F4) * 4, v.x, v.y);
    }
  }
  cluster_wait();   // every CTA of the cluster is done with this CTA's shared memory

  // ---------------- call epoch: the last CTA to finish advances it ----------------
  __syncthreads();
  if (tid == 0) {
    const uint32_t prev = atomicAdd(&ctrl->done, 1u);
    if (prev == gridDim.x - 1) {
      ctrl->done = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(&ctrl->epoch) = epoch;
    }
  }
}

// ---- host side ---------------------------------------------------------------------------------------------
struct RpPlan {
  int bh, ksplit, mode, h_tiles, tok_blocks, n_tiles, kc, mt, bk;
  size_t smem, res_off;
  size_t reg_need;   // bytes this plan uses at the START of a parity (fp32 inboxes + two-shot result area); 0 in mode 3
  size_t ll_need;    // bytes this plan uses at the END of a parity (flag-in-data inbox); 0 unless mode 3
};

int g_force_bh = 0, g_force_ksplit = 0, g_force_mode = -1;
constexpr int RP_TC_DEFAULT_MIN_TOKENS = 128;   // see rp_plan; 0 = tcgen05 tile only when forced
int g_tc_min_tokens = -1;   // tcgen05 tile from this many tokens up; -1 = not read yet (env NVLLM_RP_TC_MIN_TOKENS)
std::unordered_map<void*, size_t> g_comm_bytes;   // buffers from nvllm_tp_comm_alloc (reset clears them whole)
int g_force_bk = -1;   // tuning only: NVLLM_RP_BK=64|128 (read once)
unsigned long long g_timeout_ns = 20ull * 1000 * 1000 * 1000;

// One mapped, pinned host word per process: a kernel whose peer never arrived writes 1 + tile into it besides the
// device-side latch, so the host notices at its next poll (nvllm_rowparallel_poll_error: a plain memory read, no
// CUDA call, no synchronisation -- usable after every launch or graph replay).  nullptr if pinned memory is refused.
uint32_t* rp_host_error_word() {
  static uint32_t* word = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    if (cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
      memset(p, 0, 64);
      word = static_cast<uint32_t*>(p);
    } else {
      cudaGetLastError();
    }
  }
  return word;
}

int rp_nt16(int n_tokens) { return n_tokens >= RP_TOKB ? RP_TOKB : ((n_tokens + 15) & ~15); }

int rp_mt(int n_tokens) {   // 16-row MMA tiles per CTA: 1, 2, 4 or 8 (bucketed so the kernel loops are static)
  const int nt16 = rp_nt16(n_tokens);
  return nt16 <= 16 ? 1 : nt16 <= 32 ? 2 : nt16 <= 64 ? 4 : 8;
}

// 256-byte rows per stage (longer contiguous runs per W row, fewer barriers) whenever a 4-deep ring of them fits
int rp_bk(int bh, int mt, int k_per_rank) {
  if (g_force_bk < 0) {
    const char* e = getenv("NVLLM_RP_BK");
    g_force_bk = e ? atoi(e) : 0;
  }
  if (g_force_bk == 64) return 64;
  const bool fits = 4 * size_t(bh + mt * 16) * 256 <= RP_SMEM_BUDGET && k_per_rank % 128 == 0;
  return fits ? 128 : 64;
}

size_t rp_smem_bytes(int bh, int mt, int bk) {
  if (bh == 128)   // tcgen05 tile: ring of (128 W rows + mt*16 token rows) x 128 bytes, + 1 KiB to align it to 1024
    return size_t(rp_tc_stages(mt)) * (size_t(128) + 128) * 128 + 1024;
  const size_t ring = size_t(rp_stages(bh, mt, bk)) * (size_t(bh) + mt * 16) * bk * 2;
  const size_t epi = size_t(mt) * 16 * (bh + 8) * 4;
  return ring > epi ? ring : epi;
}

template <int BH, int BK>
const void* rp_kernel_m(int mt) {
  return mt == 1   ? reinterpret_cast<const void*>(rowparallel_allreduce_kernel<BH, BK, 1>)
         : mt == 2 ? reinterpret_cast<const void*>(rowparallel_allreduce_kernel<BH, BK, 2>)
         : mt == 4 ? reinterpret_cast<const void*>(rowparallel_allreduce_kernel<BH, BK, 4>)
                   : reinterpret_cast<const void*>(rowparallel_allreduce_kernel<BH, BK, 8>);
}

const void* rp_kernel(int bh, int bk, int mt) {
  if (bh == 128) return reinterpret_cast<const void*>(rowparallel_allreduce_kernel<128, 64, 8, true>);
  if (bh == 32) return bk == 128 ? rp_kernel_m<32, 128>(mt) : rp_kernel_m<32, 64>(mt);
  return bk == 128 ? rp_kernel_m<64, 128>(mt) : rp_kernel_m<64, 64>(mt);
}

// Load every instantiation on the current device and opt it into large dynamic shared memory, so that the first
// launch on a device never triggers lazy module loading (which may synchronise the context) while a peer's kernel is
// already spinning on this rank's flags.  Called from nvllm_tp_comm_alloc / nvllm_tp_comm_reset and, once per device,
// from the launcher.
cudaError_t rp_warm_device() {
  static bool done[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && done[dev]) return cudaSuccess;
  for (int bh = 32; bh <= 64; bh += 32)
    for (int mt = 1; mt <= 8; mt *= 2)
      for (int bk = 64; bk <= 128; bk += 64) {
        if (rp_smem_bytes(bh, mt, bk) > RP_SMEM_BUDGET) continue;   // never planned
        cudaFuncAttributes fa;
        e = cudaFuncSetAttribute(rp_kernel(bh, bk, mt), cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
        if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, rp_kernel(bh, bk, mt));
        if (e != cudaSuccess) return e;
      }
  {   // the tcgen05 instantiation
    cudaFuncAttributes fa;
    e = cudaFuncSetAttribute(rp_kernel(128, 64, 8), cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, rp_kernel(128, 64, 8));
    if (e != cudaSuccess) return e;
  }
  // mbarrier-timeout latch of the tcgen05 tile -> mapped host word (tcgen05.cuh); a refusal keeps the device-only latch
  if (tc::tc_error_arm() != cudaSuccess) cudaGetLastError();
  if (dev < 64) done[dev] = true;
  return cudaSuccess;
}

void rp_launch_config(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, int n_ctas, int ksplit, size_t smem,
                      cudaStream_t s) {
  *cfg = cudaLaunchConfig_t{};
  cfg->gridDim = dim3(n_ctas);
  cfg->blockDim = dim3(RP_THREADS);
  cfg->dynamicSmemBytes = smem;
  cfg->stream = s;
  attr->id = cudaLaunchAttributeClusterDimension;
  attr->val.clusterDim.x = ksplit;
  attr->val.clusterDim.y = 1;
  attr->val.clusterDim.z = 1;
  cfg->attrs = attr;
  cfg->numAttrs = 1;
}

// Co-resident cluster capacity of the current device for (feature tile, token bucket, cluster size).  Cached, so a
// call whose shape was seen before (the eager warm-up before graph capture) issues no CUDA runtime call besides the
// launch itself.
int rp_max_clusters(int bh, int mt, int bk, size_t smem, int ksplit, int* out) {
  static int cache[64][3][2][4][4] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return record_cuda_error(e);
  const int lg = ksplit == 1 ? 0 : ksplit == 2 ? 1 : ksplit == 4 ? 2 : 3;
  int* slot = dev < 64 ? &cache[dev][bh == 128 ? 2 : bh == 64][bk == 128][mt == 1 ? 0 : mt == 2 ? 1 : mt == 4 ? 2 : 3][lg] : nullptr;
  if (slot && *slot > 0) {
    *out = *slot;
    return NVLLM_OK;
  }
  e = rp_warm_device();
  if (e != cudaSuccess) return record_cuda_error(e);
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr;
  rp_launch_config(&cfg, &attr, RP_PLAN_CAP / ksplit * ksplit, ksplit, smem, nullptr);
  int n = 0;
  e = cudaOccupancyMaxActiveClusters(&n, rp_kernel(bh, bk, mt), &cfg);
  if (e != cudaSuccess) return record_cuda_error(e);
  *out = n;
  if (slot) *slot = n;
  return NVLLM_OK;
}

// Shape-only plan (identical on every rank, computable without a device so nvllm_rowparallel_comm_bytes can size
// the buffer): feature tile, k split (= cluster size), exchange mode; at most RP_PLAN_CAP CTAs.  The launcher halves
// the k split while the grid exceeds the device's real co-resident cluster capacity (the same on every rank of a
// homogeneous node; the inbox layout does not depend on the k split).
int rp_plan(int n_tokens, int k_per_rank, int out_features, int world, RpPlan* pl) {
  if (n_tokens <= 0 || n_tokens > RP_MAX_TOKENS || world < 2 || world > RP_MAX_WORLD) return NVLLM_ERR_UNSUPPORTED;
  if (k_per_rank <= 0 || k_per_rank % RP_BK_MIN || out_features <= 0 || out_features % 64) return NVLLM_ERR_UNSUPPORTED;
  const int tok_blocks = (n_tokens + RP_TOKB - 1) / RP_TOKB;
  int cap = RP_PLAN_CAP;
  // 64-feature tiles (2 x 2 warp grid, half the x re-reads) above 32 tokens when they still give >= 128 CTAs, and
  // always above 64 tokens or when 32-feature tiles would not fit one wave
  auto ksplit_for = [&](int tiles) {
    int ks = 1;
    for (int c = 2; c <= RP_MAX_KSPLIT; c *= 2) {
      if (k_per_rank % (c * RP_BK_MIN) || k_per_rank / c < 256 || tiles * c > cap) break;
      ks = c;
    }
    return ks;
  };
  const int tiles64 = (out_features / 64) * tok_blocks;
  const bool wide = n_tokens > 64 || (n_tokens > 32 && tiles64 * ksplit_for(tiles64) >= 128);
  int bh = (!wide && (out_features / 32) * tok_blocks <= cap) ? 32 : 64;
  // tcgen05 tile (128 features x 64|128 tokens) from g_tc_min_tokens tokens up: above that the mma.sync tile is bound by
  // its shared-memory reads and instruction issue, not by streaming W
  if (g_tc_min_tokens < 0) {   // first use: NVLLM_RP_TC_MIN_TOKENS overrides the built-in threshold (0 = never)
    const char* e = getenv("NVLLM_RP_TC_MIN_TOKENS");
    const int v = e ? atoi(e) : RP_TC_DEFAULT_MIN_TOKENS;
    g_tc_min_tokens = v <= 0 ? (1 << 30) : (v < 33 ? 33 : v);
  }
  // measured on B200, GEMM half alone (profiles/r02_rowparallel_local_tile64_mma_sync.jsonl vs
  // r02_rowparallel_local_tile128_tcgen05_tma.jsonl), 8B-shape down_proj shard (50 MB) at 64 / 128 / 256 / 512 tokens:
  // 16.1 / 17.5 / 23.7 / 35.7 us against 19.1 / 26.4 / 44.7 / 80.7 us for the mma.sync tile (cuBLAS alone: 12.2 / 12.7 /
  // 14.2 / 20.7); o_proj shard (16 MB): 10.6 / 12.4 / 15.6 / 20.2 vs 10.3 / 13.4 / 19.9 / 32.7.  So: from 128 tokens on
  // every big shard, from 64 where k per rank is long.
  // Small shards stay on the mma.sync tile at any token count: 128-feature tiles leave them too few CTAs (0.6B o_proj,
  // 256 tokens, 2 GPUs: 19.2 us vs 13.3 us end to end, profiles/r02_bench_n2_tc_small_shards_REJECTED.json).
  const bool big_shard = static_cast<int64_t>(k_per_rank) * out_features >= (int64_t(8) << 20);
  if (out_features % 128 == 0 && big_shard && (out_features / 128) * tok_blocks <= RP_TC_PLAN_CAP &&
      (n_tokens >= g_tc_min_tokens || (n_tokens >= g_tc_min_tokens / 2 && k_per_rank >= 4096)))
    bh = 128;
  if (g_force_bh == 32 || g_force_bh == 64) bh = g_force_bh;
  if (g_force_bh == 128) {
    if (out_features % 128 || n_tokens <= 32) return NVLLM_ERR_UNSUPPORTED;   // pointless below a few MMA rows
    bh = 128;
  }
  if (bh == 128) cap = RP_TC_PLAN_CAP;
  const int h_tiles = out_features / bh, n_tiles = h_tiles * tok_blocks;
  if (n_tiles > RP_MAX_TILES || n_tiles > cap) return NVLLM_ERR_UNSUPPORTED;
  int ksplit = ksplit_for(n_tiles);
  if (g_force_ksplit > 0) {
    ksplit = g_force_ksplit;
    if (ksplit > RP_MAX_KSPLIT || k_per_rank % (ksplit * RP_BK_MIN) || n_tiles * ksplit > cap) return NVLLM_ERR_UNSUPPORTED;
  }
  int mt = rp_mt(n_tokens);
  if (bh == 128) mt = 8;   // the tcgen05 tile always stages 128 token rows (UMMA N = 128; TMA zero-fills the rest)
  int bk = bh == 128 ? 64 : rp_bk(bh, mt, k_per_rank);
  if ((k_per_rank / ksplit) % bk) bk = 64;
  // one-shot while the fp32 partials a rank pushes over NVLink stay small; else reduce-scatter + all-gather
  const size_t pushed = size_t(world - 1) * n_tokens * out_features * 4;
  // flag-in-data while the payload is small (it doubles the bytes but drops a fence + flag round trip), plain
  // one-shot up to 4 MiB, then reduce-scatter + all-gather
  int mode = pushed <= (size_t(1) << 20) ? 3 : pushed <= (size_t(4) << 20) ? 0 : 1;
  if (g_force_mode >= 0 && g_force_mode <= 3) mode = g_force_mode;
  pl->bh = bh; pl->ksplit = ksplit; pl->mode = mode; pl->h_tiles = h_tiles; pl->tok_blocks = tok_blocks;
  pl->n_tiles = n_tiles; pl->kc = k_per_rank / ksplit;
  pl->bk = bk;
  pl->mt = mt;
  pl->smem = rp_smem_bytes(bh, mt, bk);
  // one parity: [fp32 inbox, one slot per rank | bf16 result area (mode 1) | ... free ... | flag-in-data inbox
  // (mode 3)].  The mode-3 inbox is anchored at the END of the parity whatever the plan, and the buffer is sized
  // (nvllm_rowparallel_comm_layout) as max(regular) + max(flag-in-data) over every plan, so no fp32 partial or bf16
  // result of ANY call can land in a word a mode-3 call reads as an epoch.
  pl->res_off = size_t(world) * n_tiles * (size_t(RP_TOKB) * bh * 4);
  const size_t regular = pl->res_off + size_t(n_tiles) * (size_t(RP_TOKB) * bh * 2);
  pl->reg_need = mode == 3 ? 0 : regular;
  pl->ll_need = mode == 3 ? 2 * pl->res_off : 0;
  return NVLLM_OK;
}

}  // namespace

// test hook (nvllm_tc_latch_selftest, unit 1): a wait in THIS unit that times out after 1 ms
int rp_tc_latch_selftest(cudaStream_t stream) { return record_cuda_error(tc::tc_latch_selftest(stream)); }
}  // namespace nvllm

using namespace nvllm;

extern "C" {

int nvllm_rowparallel_comm_layout(int32_t world, int32_t max_tokens, int32_t in_features_per_rank,
                                  int32_t out_features, size_t* header_bytes_out, size_t* regular_bytes_out,
                                  size_t* flag_in_data_bytes_out) {
  // worst case of the shape-only plan over every token count and both feature tiles (a k split or mode forced through
  // nvllm_rowparallel_set_plan is honoured: size the buffer after forcing it)
  if (max_tokens > RP_MAX_TOKENS) max_tokens = RP_MAX_TOKENS;
  const int saved_bh = g_force_bh;
  size_t reg = 0, ll = 0;
  bool any = false;
  for (int n = 1; n <= max_tokens; ++n)
    for (int bh = 32; bh <= 128; bh *= 2) {
      RpPlan pl;
      g_force_bh = bh;
      if (rp_plan(n, in_features_per_rank, out_features, world, &pl) != NVLLM_OK) continue;
      any = true;
      if (pl.reg_need > reg) reg = pl.reg_need;
      if (pl.ll_need > ll) ll = pl.ll_need;
    }
  g_force_bh = saved_bh;
  if (!any) return NVLLM_ERR_UNSUPPORTED;
  if (header_bytes_out) *header_bytes_out = RP_DATA_OFF;
  if (regular_bytes_out) *regular_bytes_out = reg + 16;   // the launcher keeps 16 bytes of slack per parity
  if (flag_in_data_bytes_out) *flag_in_data_bytes_out = ll;
  return NVLLM_OK;
}

size_t nvllm_rowparallel_comm_bytes(int32_t world, int32_t max_tokens, int32_t in_features_per_rank,
                                    int32_t out_features) {
  size_t hdr = 0, reg = 0, ll = 0;
  if (nvllm_rowparallel_comm_layout(world, max_tokens, in_features_per_rank, out_features, &hdr, &reg, &ll) != NVLLM_OK)
    return 0;
  return hdr + 2 * (reg + ll);
}

int nvllm_tp_comm_alloc(size_t bytes, void** dev_ptr, void* ipc_handle_out) {
  if (!dev_ptr || bytes < RP_DATA_OFF) return NVLLM_ERR_BAD_ARG;
  void* p = nullptr;
  cudaError_t e = rp_warm_device();
  if (e != cudaSuccess) return record_cuda_error(e);
  e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) return record_cuda_error(e);
  e = cudaMemset(p, 0, bytes);   // epoch 0, flags 0, and no word of the flag-in-data inboxes may look like an epoch
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e == cudaSuccess && ipc_handle_out) {
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, p);
    if (e == cudaSuccess) memcpy(ipc_handle_out, &h, sizeof(h));
  }
  if (e != cudaSuccess) {
    cudaFree(p);
    return record_cuda_error(e);
  }
  *dev_ptr = p;
  g_comm_bytes[p] = bytes;
  return NVLLM_OK;
}

int nvllm_tp_comm_open(const void* ipc_handle, void** peer_ptr) {
  if (!ipc_handle || !peer_ptr) return NVLLM_ERR_BAD_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return record_cuda_error(e);
  *peer_ptr = p;
  return NVLLM_OK;
}

int nvllm_tp_comm_close(void* peer_ptr) {
  if (!peer_ptr) return NVLLM_OK;
  return record_cuda_error(cudaIpcCloseMemHandle(peer_ptr));
}

int nvllm_tp_comm_free(void* dev_ptr) {
  if (!dev_ptr) return NVLLM_OK;
  g_comm_bytes.erase(dev_ptr);
  return record_cuda_error(cudaFree(dev_ptr));
}

int nvllm_tp_comm_reset(void* dev_ptr, void* stream) {
  if (!dev_ptr) return NVLLM_ERR_BAD_ARG;
  const cudaError_t e = rp_warm_device();
  if (e != cudaSuccess) return record_cuda_error(e);
  // buffers of our own are cleared whole (the flag-in-data inboxes must not keep words of an earlier epoch count);
  // for memory allocated elsewhere the caller zeroes the data area itself before this call
  const auto it = g_comm_bytes.find(dev_ptr);
  const size_t n = it != g_comm_bytes.end() ? it->second : RP_DATA_OFF;
  return record_cuda_error(cudaMemsetAsync(dev_ptr, 0, n, static_cast<cudaStream_t>(stream)));
}

int nvllm_tp_comm_enable_peer(int32_t peer_device) {
  int dev = 0, can = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return record_cuda_error(e);
  if (peer_device == dev) return NVLLM_OK;
  e = cudaDeviceCanAccessPeer(&can, dev, peer_device);
  if (e != cudaSuccess) return record_cuda_error(e);
  if (!can) return NVLLM_ERR_UNSUPPORTED;
  e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    e = cudaSuccess;
  }
  return record_cuda_error(e);
}

int nvllm_tp_comm_status(const void* dev_ptr, uint32_t* epoch_out, uint32_t* error_out) {
  if (!dev_ptr) return NVLLM_ERR_BAD_ARG;
  RpControl c;
  cudaError_t e = cudaMemcpy(&c, dev_ptr, sizeof(c), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return record_cuda_error(e);
  if (epoch_out) *epoch_out = c.epoch;
  if (error_out) *error_out = c.error;
  return NVLLM_OK;
}

int nvllm_rowparallel_poll_error(uint32_t* error_out, int32_t clear) {
  if (!error_out) return NVLLM_ERR_BAD_ARG;
  uint32_t* w = rp_host_error_word();
  const uint32_t peer = w ? *reinterpret_cast<volatile uint32_t*>(w) : 0u;
  // the tcgen05 tile's mbarrier-timeout latch (bit 31 set) reports through the same call: both are host memory reads
  const uint32_t tile = tc::tc_error_poll();
  *error_out = peer ? peer : tile;
  if (w && clear) *reinterpret_cast<volatile uint32_t*>(w) = 0u;
  if (tile && clear) {   // error path only: re-arms the device-side latch too (synchronises the current device)
    unsigned int v = 0;
    return record_cuda_error(tc::tc_error_fetch(&v));
  }
  return NVLLM_OK;
}

int nvllm_rowparallel_set_plan(int32_t feature_tile, int32_t ksplit, int32_t mode, int64_t timeout_ms) {
  if (!(feature_tile == 0 || feature_tile == 32 || feature_tile == 64 || feature_tile == 128)) return NVLLM_ERR_BAD_ARG;
  if (ksplit < 0 || ksplit > RP_MAX_KSPLIT || (ksplit & (ksplit - 1))) return NVLLM_ERR_BAD_ARG;
  if (mode < -1 || mode > 3) return NVLLM_ERR_BAD_ARG;   // 2: local partial only (profiling), 3: flag-in-data
  g_force_bh = feature_tile;
  g_force_ksplit = ksplit;
  g_force_mode = mode;
  if (timeout_ms > 0) g_timeout_ns = static_cast<unsigned long long>(timeout_ms) * 1000000ull;
  return NVLLM_OK;
}

int nvllm_rowparallel_set_tc_min_tokens(int32_t min_tokens) {
  g_tc_min_tokens = min_tokens <= 0 ? (1 << 30) : (min_tokens < 33 ? 33 : min_tokens);
  return NVLLM_OK;
}

int nvllm_rowparallel_plan(int32_t n_tokens, int32_t in_features_per_rank, int32_t out_features, int32_t world,
                           int32_t* feature_tile_out, int32_t* ksplit_out, int32_t* mode_out) {
  RpPlan pl;
  const int st = rp_plan(n_tokens, in_features_per_rank, out_features, world, &pl);
  if (st != NVLLM_OK) return st;
  if (feature_tile_out) *feature_tile_out = pl.bh;
  if (ksplit_out) *ksplit_out = pl.ksplit;
  if (mode_out) *mode_out = pl.mode;
  return NVLLM_OK;
}

int nvllm_rowparallel_linear_allreduce(void* y, int64_t y_row_stride, const void* x, int64_t x_row_stride,
                                       const void* weight, int64_t w_row_stride, const void* bias,
                                       int32_t n_tokens, int32_t in_features_per_rank, int32_t out_features,
                                       void* const* comm_ptrs, size_t comm_bytes, int32_t rank, int32_t world,
                                       void* stream) {
  if (n_tokens == 0) return NVLLM_OK;
  if (!y || !x || !weight || !comm_ptrs || rank < 0 || rank >= world) return NVLLM_ERR_BAD_ARG;
  if (!aligned16(y) || !aligned16(x) || !aligned16(weight) || (y_row_stride % 8) || (x_row_stride % 8) ||
      (w_row_stride % 8) || (bias && (reinterpret_cast<uintptr_t>(bias) & 7u)))
    return NVLLM_ERR_ALIGNMENT;
  if (y_row_stride < out_features || x_row_stride < in_features_per_rank || w_row_stride < in_features_per_rank)
    return NVLLM_ERR_BAD_ARG;
  RpPlan pl;
  int st = rp_plan(n_tokens, in_features_per_rank, out_features, world, &pl);
  if (st != NVLLM_OK) return st;
  if (comm_bytes < RP_DATA_OFF + 2 * (pl.reg_need + pl.ll_need + 16)) return NVLLM_ERR_WORKSPACE;

  // every CTA of the grid must be co-resident (cross-rank waits inside the kernel): shrink the cluster while the
  // device cannot hold all of them at once
  for (;;) {
    int max_clusters = 0;
    st = rp_max_clusters(pl.bh, pl.mt, pl.bk, pl.smem, pl.ksplit, &max_clusters);
    if (st != NVLLM_OK) return st;
    if (pl.n_tiles <= max_clusters) break;
    if (pl.ksplit == 1 || g_force_ksplit > 0) return NVLLM_ERR_UNSUPPORTED;
    pl.ksplit /= 2;
    pl.kc *= 2;
  }

  RpParams p{};
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.w = static_cast<const __nv_bfloat16*>(weight);
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.y = static_cast<__nv_bfloat16*>(y);
  p.x_stride = x_row_stride; p.w_stride = w_row_stride; p.y_stride = y_row_stride;
  for (int r = 0; r < world; ++r) {
    if (!comm_ptrs[r] || !aligned16(comm_ptrs[r])) return NVLLM_ERR_BAD_ARG;
    p.comm[r] = static_cast<uint8_t*>(comm_ptrs[r]);
  }
  p.rank = rank; p.world = world; p.n_tokens = n_tokens; p.ksplit = pl.ksplit; p.kc = pl.kc;
  p.h_tiles = pl.h_tiles; p.n_tiles = pl.n_tiles; p.mode = pl.mode;
  p.par_bytes = (comm_bytes - RP_DATA_OFF) / 2 / 16 * 16;
  p.res_off = pl.res_off;
  p.ll_off = (p.par_bytes - pl.ll_need) / 16 * 16;   // end-anchored: see rp_plan
  p.timeout_ns = g_timeout_ns;
  p.host_err = rp_host_error_word();
  if (pl.bh == 128) {
    // TMA descriptors of this call's operands (host-side encode, ~1 us each; baked into a captured graph's node)
    if (!tc::make_tile_map_bf16(&p.tm_w, weight, static_cast<uint64_t>(in_features_per_rank),
                                static_cast<uint64_t>(out_features), static_cast<uint64_t>(w_row_stride)) ||
        !tc::make_tile_map_bf16(&p.tm_x, x, static_cast<uint64_t>(in_features_per_rank), static_cast<uint64_t>(n_tokens),
                                static_cast<uint64_t>(x_row_stride)))
      return NVLLM_ERR_UNSUPPORTED;
  }
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr;
  rp_launch_config(&cfg, &attr, pl.n_tiles * pl.ksplit, pl.ksplit, pl.smem, static_cast<cudaStream_t>(stream));
  void* args[] = {&p};
  const cudaError_t le = cudaLaunchKernelExC(&cfg, rp_kernel(pl.bh, pl.bk, pl.mt), args);
  if (le != cudaSuccess) return record_cuda_error(le);
  note_launch();
  return record_cuda_error(cudaGetLastError());
}

}  // extern "C"
