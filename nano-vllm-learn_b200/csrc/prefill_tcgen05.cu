// Varlen causal prefill on the 5th-gen tensor cores (sm_100a): tcgen05.mma with TMEM accumulators,
// operands staged by TMA.  Replaces flash_attn_varlen_func (reference nanovllm/layers/attention.py:89-96)
// including the paged-prefix (prefix cache) form: K/V tiles come either from this step's dense k/v or
// from the paged cache through block_tables, both via 2-D tensor maps with 128-byte swizzle.
//
// Persistent kernel: one CTA per SM loops over work items (sequence, q head, PAIR of 128-query tiles A and
// B); barrier phases run on across items, so the TMA producer prefetches the next item's Q/K/V while the
// softmax warpgroups drain the current item's output.  640 threads:
//   warpgroups 0-3 (warps 0-15) softmax: warpgroup g owns columns 32g..32g+31 of EVERY S tile (A and B) and of
//                               every O tile, one query row per thread (warps w, w+4, w+8, w+12 share TMEM lanes)
//   warp 16                     TMA producer (one elected lane)
//   warp 17                     tcgen05.mma issuer (one elected lane) + TMEM allocator
//   warps 18-19                 idle (complete the fifth warpgroup for setmaxnreg)
// Both tiles share every K/V tile (loaded once).  Per 128-key tile j and query tile t in {A, B}:
//   S_t = Q_t . K_j^T      SS MMA, 8 x (M128 N128 K16); Q_t, K_j K-major SW128 in shared memory
//   softmax (both groups): half an S row -> registers, running max (halves agree through shared memory) with lazy
//                          rescale (only when the max grows by more than 2^8), P = exp2(S*c - m*c) in bf16 written
//                          back over S (P aliases S); part of the exponentials run on the FMA pipe (exp2_poly2)
//   O_t += P . V_j         TS MMA: A = P from TMEM, B = V_j MN-major SW128 in shared memory; O_t stays in TMEM
// The issuer alternates A and B, so the tensor pipe runs one tile's MMAs while all eight softmax warps work on the
// other tile (ping-pong).  TMEM: S_A S_B O_A O_B = 4 x 128 columns = all 512.
#include "prefill_params.cuh"
#include "tcgen05.cuh"
#ifdef NVLLM_TOOLS_BUILD
#include "../../include/nvllm_tools.h"
#endif

namespace nvllm {

namespace {

constexpr int TM = 128;                      // query rows per tile (two tiles per CTA)
constexpr int TN = 128;                      // keys per tile
constexpr int kHalfBytes = 128 * 128;        // [128 rows][64 bf16] = 16 KB, one SW128 half of a tile
constexpr int kTileBytes = 2 * kHalfBytes;   // 32 KB
constexpr int kStages = 2;                   // K/V ring depth
constexpr int kSoftmaxGroups = 2;            // softmax warpgroups
constexpr int kTeams = 1;                    // 2: warpgroups 0-1 work on query tile A, 2-3 on tile B; 1: all on every tile
constexpr int kSoftmaxWarps = 4 * kSoftmaxGroups;
constexpr int kThreads = (kSoftmaxWarps + 4) * 32;   // + one warpgroup: TMA producer, MMA issuer, two idle warps
constexpr int kRegsSoftmax = kSoftmaxGroups == 2 ? 200 : 104;   // 512 * 104 + 128 * 64 < 64 K: asking for exactly the whole file never completes   // setmaxnreg budgets: 128 * (kColGroups * soft + other) <= 64 K
constexpr int kRegsOther = kSoftmaxGroups == 2 ? 96 : 64;
constexpr int kSmemQ = 0;                                    // Q_A, Q_B
constexpr int kSmemK = kSmemQ + 2 * kTileBytes;
constexpr int kSmemV = kSmemK + kStages * kTileBytes;
constexpr int kSmemBar = kSmemV + kStages * kTileBytes;      // 192 KB
constexpr int kSmemXchg = kSmemBar + 256;                    // row max / row sum exchange between the column halves
constexpr int kSmemStage = kSmemXchg + 2 * kSoftmaxGroups * 128 * 4;   // per-warp output staging: 32 rows x 80 B
constexpr int kStagePitch = 80;
constexpr int kSmemBytes = kSmemStage + kSoftmaxWarps * 32 * kStagePitch + 1024;   // + alignment slack
constexpr uint32_t kTmemCols = 512;
constexpr float kRescaleThreshold = 8.f;     // log2 units: P stays <= 2^8 relative to the stale max

constexpr uint32_t kIdescQK = tc::make_idesc_bf16(TM, TN, /*b_mn_major=*/false);
constexpr uint32_t kIdescPV = tc::make_idesc_bf16(TM, kHeadDim, /*b_mn_major=*/true);

enum Bar : int {
  kQFull = 0,                    // [2]  per query tile
  kKFull = 2,                    // [kStages]
  kVFull = 2 + kStages,          // [kStages]
  kKVEmpty = 2 + 2 * kStages,    // [kStages]
  kSFull = 2 + 3 * kStages,      // [2]
  kPFull = kSFull + 2,           // [2]
  kOFull = kPFull + 2,           // [2]
  kQEmpty = kOFull + 2,          // [2]  Q_t no longer read by any pending MMA (persistent loop)
  kOEmpty = kQEmpty + 2,         // [2]  epilogue of tile t has drained O_t from TMEM
  kNumBars = kOEmpty + 2,
};

// Work item = (sequence, q head, pair of query tiles).  Items are enumerated heavy-first (late pairs see
// the most keys under the causal mask) and adjacent ids share the sequence and kv head, so CTAs running
// side by side hit the same K/V tiles in L2.  Every role of a CTA decodes the same item the same way.
// The two query tiles of an item are either two consecutive 128-row blocks of ONE q head (row pairing) or the SAME
// 128 rows of two q heads that share the kv head (head pairing, p.pair_heads; needs an even GQA group).  Head pairing
// gives both tiles the same K/V tiles AND the same causal mask, so n_t[0] == n_t[1]: with row pairing tile B always
// runs one KV tile more than A, and that last tile has no partner to ping-pong with -- 1 of only 2-8 tiles per item
// at 1024-token prompts.
struct Item {
  int seq, h, kvh, q0, k0, Lq, Lk, m0, shift, n_t[2], n_tiles;
  int mr[2];   // first query row (inside the sequence) of tile t
  int hq[2];   // q head of tile t
  bool live;
};

// `slot` = CTA + round * grid.  Rounds alternate direction (CTA i takes item i of an even round and item grid-1-i of an
// odd one): with heavy-first items a plain round-robin hands CTA 0 the heaviest item of every round (1 x 8192: the
// busiest CTA carried 29 % more than the mean, 4 % this way).
__device__ __forceinline__ Item decode_item(const PrefillParams& p, int slot, int num_pairs, int batch) {
  Item it;
  const int heads = p.pair_heads ? p.num_q_heads / 2 : p.num_q_heads;   // head slots per (row tile | tile pair, sequence)
  const int per_pair = heads * batch;
  const int grid = static_cast<int>(gridDim.x);
  const int round = slot / grid, cta = slot % grid;
  const int w = round * grid + ((round & 1) ? grid - 1 - cta : cta);
  if (w >= num_pairs * per_pair) {   // the last round may be partial
    it.seq = it.h = it.kvh = it.q0 = it.k0 = it.Lq = it.Lk = it.m0 = it.shift = it.n_tiles = 0;
    it.n_t[0] = it.n_t[1] = 0;
    it.mr[0] = it.mr[1] = it.hq[0] = it.hq[1] = 0;
    it.live = false;
    return it;
  }
  const int mbp = num_pairs - 1 - w / per_pair;
  const int rem = w % per_pair;
  it.seq = rem / heads;
  it.h = p.pair_heads ? 2 * (rem % heads) : rem % heads;
  it.kvh = it.h / p.group;
  it.q0 = __ldg(p.cu_q + it.seq);
  it.Lq = __ldg(p.cu_q + it.seq + 1) - it.q0;
  it.k0 = __ldg(p.cu_k + it.seq);
  it.Lk = __ldg(p.cu_k + it.seq + 1) - it.k0;
  it.m0 = p.pair_heads ? mbp * TM : mbp * 2 * TM;
  it.shift = it.Lk - it.Lq;
  it.live = it.m0 < it.Lq;
  it.mr[0] = it.m0;
  it.hq[0] = it.h;
  const int hiA = min(it.Lk, it.shift + it.m0 + TM);
  it.n_t[0] = (it.live && hiA > 0) ? (hiA + TN - 1) / TN : 0;
  if (p.pair_heads) {
    it.mr[1] = it.m0;
    it.hq[1] = it.h + 1;
    it.n_t[1] = it.n_t[0];
  } else {
    it.mr[1] = it.m0 + TM;
    it.hq[1] = it.h;
    const int hiB = (it.m0 + TM < it.Lq) ? min(it.Lk, it.shift + it.m0 + 2 * TM) : 0;
    it.n_t[1] = (it.live && hiB > 0) ? (hiB + TN - 1) / TN : 0;
  }
  it.n_tiles = max(it.n_t[0], it.n_t[1]);
  it.live = it.live && it.n_tiles > 0;
  return it;
}


// exp2 of a pair on the FMA pipe (no MUFU): round to the nearest integer n by the 1.5 * 2^23 trick, a cubic for
// 2^f on f in [-0.5, 0.5] (max relative error 7.5e-5 = 2^-13.7, far below the bf16 rounding of P), n added to the
// exponent field.  x is clamped to >= -125 (2^-125 stands in for anything smaller, -inf included); x <= ~9 here.
// The MUFU unit makes 4 ex2 per clock per SM sub-partition: for a 128 x 128 S tile that is as long as the tile's two
// MMAs take, so part of every row goes through this path (FFMA2 / FADD2 are two-wide) to pull the softmax off the
// critical path.
__device__ __forceinline__ float2 exp2_poly2(float2 x) {
  constexpr float kMagic = 12582912.f;   // 1.5 * 2^23
  x.x = fmaxf(x.x, -125.f);
  x.y = fmaxf(x.y, -125.f);
  const float2 t = __fadd2_rn(x, make_float2(kMagic, kMagic));
  const float2 r = __fadd2_rn(t, make_float2(-kMagic, -kMagic));
  const float2 f = __ffma2_rn(r, make_float2(-1.f, -1.f), x);
  float2 q = __ffma2_rn(make_float2(0.05517147481441498f, 0.05517147481441498f), f,
                        make_float2(0.242610901594162f, 0.242610901594162f));
  q = __ffma2_rn(q, f, make_float2(0.6932609677314758f, 0.6932609677314758f));
  q = __ffma2_rn(q, f, make_float2(0.9999281167984009f, 0.9999281167984009f));
  float2 out;
  out.x = __int_as_float(__float_as_int(q.x) + (__float_as_int(t.x) << 23));
  out.y = __int_as_float(__float_as_int(q.y) + (__float_as_int(t.y) << 23));
  return out;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;\n" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// Debug-build timeline (tools/trace_prefill.py): CTA 0 stamps clock64() at the protocol points of its first
// work item into debug[32768 + (role * 64 + step) * 8 + k]; role 0 = MMA issuer, 1 = softmax A, 2 = softmax B.
__device__ __forceinline__ void trace_stamp(float* debug, int role, int step, int k) {
  if (step < 64) reinterpret_cast<uint32_t*>(debug)[32768 + (role * 64 + step) * 8 + k] = static_cast<uint32_t>(clock64());
}

// kEmu: of every 8 column pairs of an unmasked S tile, how many take exp2_poly2 instead of the MUFU unit.
// Item-level timeline of CTA 0 (tools/trace_prefill.py --items): debug[32768 + (4 * 64 + item) * 8 + k], k: 0 softmax
// threads start the item, 1 its first S tile is ready, 2 its last P has been handed over, 3 the O tiles are final,
// 4 the output stores are issued; 5 the MMA issuer starts the item, 6 its first QK is issued (Q and K landed), 7 its
// last PV is issued.
__device__ __forceinline__ void item_stamp(float* debug, int item, int k) {
  if (item < 64) reinterpret_cast<uint32_t*>(debug)[32768 + (4 * 64 + item) * 8 + k] = static_cast<uint32_t>(clock64());
}

template <bool kDebug, int kEmu>
__global__ void __launch_bounds__(kThreads, 1)
prefill_tcgen05_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const PrefillParams p, const int num_pairs,
                       const int batch, float* __restrict__ debug) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SW128 tiles need 1024-byte alignment
  unsigned char* smem_gen = smem_raw + (smem - smem_u32(smem_raw));

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  // slots of this launch: whole rounds of the grid (decode_item marks the tail of the last round dead)
  const int total_items = (num_pairs * p.num_q_heads * batch + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x) *
                          static_cast<int>(gridDim.x);

  const uint32_t bar0 = smem + kSmemBar;
  auto bar = [&](int i) -> uint32_t { return bar0 + 8u * static_cast<uint32_t>(i); };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem_gen + kSmemBar + 8 * kNumBars);

  if (warp == kSoftmaxWarps && lane == 0) {
    for (int t = 0; t < 2; ++t) {
      tc::mbar_init(bar(kQFull + t), 1);
      tc::mbar_init(bar(kSFull + t), 1);
      tc::mbar_init(bar(kPFull + t), kSoftmaxWarps * 32 / kTeams);
      tc::mbar_init(bar(kOFull + t), 1);
      tc::mbar_init(bar(kQEmpty + t), 1);
      tc::mbar_init(bar(kOEmpty + t), kSoftmaxWarps * 32 / kTeams);
    }
    for (int s = 0; s < kStages; ++s) {
      tc::mbar_init(bar(kKFull + s), 1);
      tc::mbar_init(bar(kVFull + s), 1);
      tc::mbar_init(bar(kKVEmpty + s), 1);
    }
    tc::fence_barrier_init();
  }
  if (warp == kSoftmaxWarps + 1) {
    tc::tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), kTmemCols);
    tc::tmem_relinquish();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;
  // debug build only: bit0 = softmax skips exp2 + P stores, bit1 = also skips the S loads, bit2 = the producer
  // arrives on K/V barriers without loading (tools/trace_prefill.py uses these to time the MMA pipe alone)
  const uint32_t dbg_mode =
      (kDebug && debug != nullptr) ? reinterpret_cast<const uint32_t*>(debug)[32768 + 3 * 64 * 8] : 0u;

  if (warp >= kSoftmaxWarps) {
    tc::reg_dealloc<kRegsOther>();
    if (warp == kSoftmaxWarps) {
      // ================================ TMA producer (warp-wide, issue by one elected lane) ==========
      const bool lead = tc::elect_one();
      if (lead) {
        tc::tma_prefetch_desc(&tm_q);
        tc::tma_prefetch_desc(&tm_k);
        tc::tma_prefetch_desc(&tm_v);
      }
      uint32_t kv_iter = 0;        // K/V tiles loaded so far (ring position + phase)
      uint32_t q_uses[2] = {0, 0};  // how often Q_t has been filled
      for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
        const Item it = decode_item(p, w, num_pairs, batch);
        if (!it.live) continue;
        for (int t = 0; t < 2; ++t) {
          if (it.n_t[t] == 0) continue;
          tc::mbar_wait(bar(kQEmpty + t), (q_uses[t] & 1u) ^ 1u);
          ++q_uses[t];
          const uint32_t sq = smem + kSmemQ + t * kTileBytes;
          if (lead) {
            tc::mbar_arrive_expect_tx(bar(kQFull + t), kTileBytes);
            tc::tma_load_2d(sq, &tm_q, it.hq[t] * kHeadDim, it.q0 + it.mr[t], bar(kQFull + t));
            tc::tma_load_2d(sq + kHalfBytes, &tm_q, it.hq[t] * kHeadDim + 64, it.q0 + it.mr[t], bar(kQFull + t));
          }
          __syncwarp();
        }
        const int32_t* table =
            p.block_tables ? p.block_tables + static_cast<int64_t>(it.seq) * p.table_stride : nullptr;
        for (int j = 0; j < it.n_tiles; ++j, ++kv_iter) {
          const uint32_t st = kv_iter % kStages;
          tc::mbar_wait(bar(kKVEmpty + st), ((kv_iter / kStages) & 1u) ^ 1u);
          int row;
          if (table) {
            const int t0 = j * TN;
            row = __ldg(table + t0 / p.block_size) * p.block_size + t0 % p.block_size;
          } else {
            row = it.k0 + j * TN;
          }
          const uint32_t sk = smem + kSmemK + st * kTileBytes;
          const uint32_t sv = smem + kSmemV + st * kTileBytes;
          if (lead && (dbg_mode & 4u)) {
            tc::mbar_arrive(bar(kKFull + st));
            tc::mbar_arrive(bar(kVFull + st));
          } else if (lead) {
            tc::mbar_arrive_expect_tx(bar(kKFull + st), kTileBytes);
            tc::tma_load_2d(sk, &tm_k, it.kvh * kHeadDim, row, bar(kKFull + st));
            tc::tma_load_2d(sk + kHalfBytes, &tm_k, it.kvh * kHeadDim + 64, row, bar(kKFull + st));
            tc::mbar_arrive_expect_tx(bar(kVFull + st), kTileBytes);
            tc::tma_load_2d(sv, &tm_v, it.kvh * kHeadDim, row, bar(kVFull + st));
            tc::tma_load_2d(sv + kHalfBytes, &tm_v, it.kvh * kHeadDim + 64, row, bar(kVFull + st));
          }
          __syncwarp();
        }
      }
    } else if (warp == kSoftmaxWarps + 1) {
      // ================================ MMA issuer (warp-wide, issue by one elected lane) ============
      // Measured with clock64 stamps (tools/trace_prefill.py): run by a single divergent lane, every
      // tcgen05.mma cost ~100-140 cycles of R2UR moves plus an ELECT/branch waterfall, i.e. the issue path,
      // not the tensor pipe or the softmax, set the pace.  Warp-uniform control flow keeps descriptors in
      // uniform registers.
      const bool lead = tc::elect_one();
      const uint64_t dq0 = tc::make_smem_desc(smem + kSmemQ, 16, 1024);
      uint32_t kv_base = 0;          // K/V tiles consumed by earlier items
      uint32_t q_uses[2] = {0, 0};   // items that used tile t (phase of q_full / o_empty)
      uint32_t p_cnt[2] = {0, 0};    // P tiles consumed per query tile (phase of p_full)
      int item_k = 0;
      for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
        const Item it = decode_item(p, w, num_pairs, batch);
        if (!it.live) continue;
        const bool tri = kDebug && lead && debug != nullptr && blockIdx.x == 0;
        if (tri) item_stamp(debug, item_k, 5);
        auto issue_qk = [&](int t, int j) {
          const uint32_t kv = kv_base + static_cast<uint32_t>(j);
          const uint32_t st = kv % kStages;
          if (j == 0) tc::mbar_wait(bar(kQFull + t), q_uses[t] & 1u);
          tc::mbar_wait(bar(kKFull + st), (kv / kStages) & 1u);
          tc::fence_after_sync();
          // descriptors differ only in the 14-bit start-address field (units of 16 B): add offsets to the low word
          const uint64_t da = dq0 + static_cast<uint64_t>((t * kTileBytes) >> 4);
          const uint64_t db = dq0 + static_cast<uint64_t>((kSmemK - kSmemQ + st * kTileBytes) >> 4);
          const uint32_t d = tmem + static_cast<uint32_t>(t * TN);
          if (kDebug && (dbg_mode & 24u)) {
            // timing experiments only (results are garbage): bit3 = alternate two accumulators, bit4 = N=256
            if (lead) {
              constexpr int kHalf16 = kHalfBytes >> 4;
              if (dbg_mode & 8u) {
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                  const uint64_t off = static_cast<uint64_t>((ks >> 2) * kHalf16 + (ks & 3) * 2);
                  const uint32_t dd = (ks & 1) ? tmem + 2u * TN + static_cast<uint32_t>(t * kHeadDim) : d;
                  tc::mma_ss(dd, da + off, db + off, kIdescQK, ks > 1 ? 1u : 0u);
                }
              } else {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                  const uint64_t off = static_cast<uint64_t>((ks >> 2) * kHalf16 + (ks & 3) * 2);
                  tc::mma_ss(tmem + static_cast<uint32_t>(t * 2 * TN), da + off, db + off,
                             tc::make_idesc_bf16(TM, 256, false), ks > 0 ? 1u : 0u);
                }
              }
              tc::mma_commit(bar(kSFull + t));
              if (j == it.n_t[t] - 1) tc::mma_commit(bar(kQEmpty + t));
            }
          } else if (lead) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
              // 16 bf16 = 32 B inside the swizzle row; dims 64..127 live in the second SW128 half
              constexpr int kHalf16 = kHalfBytes >> 4;
              const uint64_t off = static_cast<uint64_t>((ks >> 2) * kHalf16 + (ks & 3) * 2);
              tc::mma_ss(d, da + off, db + off, kIdescQK, ks > 0 ? 1u : 0u);
            }
            tc::mma_commit(bar(kSFull + t));
            if (j == it.n_t[t] - 1) tc::mma_commit(bar(kQEmpty + t));  // last read of Q_t for this item
          }
          __syncwarp();
        };
        auto issue_pv = [&](int t, int j) {
          const uint32_t kv = kv_base + static_cast<uint32_t>(j);
          const uint32_t st = kv % kStages;
          tc::mbar_wait(bar(kVFull + st), (kv / kStages) & 1u);
          tc::mbar_wait(bar(kPFull + t), p_cnt[t] & 1u);
          ++p_cnt[t];
          if (j == 0) tc::mbar_wait(bar(kOEmpty + t), (q_uses[t] & 1u) ^ 1u);  // previous item's epilogue drained O_t
          tc::fence_after_sync();
          const uint64_t dv = tc::make_smem_desc(smem + kSmemV + st * kTileBytes, kHalfBytes, 1024);
          const uint32_t d = tmem + 2u * TN + static_cast<uint32_t>(t * kHeadDim);
          const uint32_t a = tmem + static_cast<uint32_t>(t * TN);
          if (lead) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
              // 16 keys = 16 rows of 128 B per k-step (2048 B = 128 x 16 B); the two 64-dim halves are LBO apart
              tc::mma_ts(d, a + static_cast<uint32_t>(ks * 8), dv + static_cast<uint64_t>(ks * 128), kIdescPV,
                         (j > 0 || ks > 0) ? 1u : 0u);
            }
            tc::mma_commit(bar(kOFull + t));
          }
          __syncwarp();
        };
        for (int t = 0; t < 2; ++t)
          if (it.n_t[t] > 0) issue_qk(t, 0);
        if (tri) item_stamp(debug, item_k, 6);
        const bool tr = kDebug && lead && debug != nullptr && blockIdx.x == 0 && w == static_cast<int>(blockIdx.x);
        for (int j = 0; j < it.n_tiles; ++j) {
          if (tr) trace_stamp(debug, 0, j, 0);
          if (j < it.n_t[0]) issue_pv(0, j);
          if (tr) trace_stamp(debug, 0, j, 1);
          if (j + 1 < it.n_t[0]) issue_qk(0, j + 1);
          if (tr) trace_stamp(debug, 0, j, 2);
          if (j < it.n_t[1]) issue_pv(1, j);
          if (tr) trace_stamp(debug, 0, j, 3);
          if (lead) tc::mma_commit(bar(kKVEmpty + (kv_base + j) % kStages));  // K_j, V_j no longer read by any pending MMA
          __syncwarp();
          if (j + 1 < it.n_t[1]) issue_qk(1, j + 1);
          if (tr) trace_stamp(debug, 0, j, 4);
        }
        if (tri) item_stamp(debug, item_k, 7);
        ++item_k;
        kv_base += static_cast<uint32_t>(it.n_tiles);
        for (int t = 0; t < 2; ++t)
          if (it.n_t[t] > 0) ++q_uses[t];
      }
    }
    __syncwarp();
  } else {
    tc::reg_alloc<kRegsSoftmax>();
#include "prefill_tcgen05_softmax.inl"
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == kSoftmaxWarps + 1) tc::tmem_dealloc(tmem, kTmemCols);
}

// ---- host: tensor maps ------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

// [rows][cols] bf16, row stride in elements; box = 128 rows x 64 cols (128 B), SWIZZLE_128B, zero OOB fill.
bool make_map(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t row_stride_elems) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

constexpr int kDefaultEmu = 2;
#ifdef NVLLM_TOOLS_BUILD
float* g_debug = nullptr;
int g_emu = -1;
#endif

}  // namespace

int launch_prefill_tcgen05(const PrefillParams& p, int batch, int max_seqlen_q, int /*max_seqlen_k*/,
                           cudaStream_t stream, bool* handled) {
  *handled = false;
  if (p.total_q <= 0 || p.total_k <= 0) return NVLLM_OK;  // caller did not give TMA bounds: use the mma.sync kernel
  if (p.block_tables && p.block_size % TN != 0) return NVLLM_OK;
  // pair the two query tiles of a work item by GQA head whenever the group allows it (NVLLM_PREFILL_PAIR=rows keeps
  // the round-1 pairing of two consecutive row blocks of one head; read per call so sweeps can flip it)
  PrefillParams pp = p;
  const char* pair_env = getenv("NVLLM_PREFILL_PAIR");
  pp.pair_heads = (p.group % 2 == 0 && !(pair_env && pair_env[0] == 'r')) ? 1 : 0;
  CUtensorMap tq, tk, tv;
  const uint64_t kv_cols = static_cast<uint64_t>(p.num_kv_heads) * kHeadDim;
  if (!make_map(&tq, p.q, static_cast<uint64_t>(p.num_q_heads) * kHeadDim, p.total_q, p.q_row_stride)) return NVLLM_OK;
  if (!make_map(&tk, p.k, kv_cols, p.total_k, p.k_row_stride)) return NVLLM_OK;
  if (!make_map(&tv, p.v, kv_cols, p.total_k, p.v_row_stride)) return NVLLM_OK;
  using Kernel = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const PrefillParams, const int,
                          const int, float*);
  Kernel kernel = prefill_tcgen05_kernel<false, kDefaultEmu>;
  float* debug = nullptr;
#ifdef NVLLM_TOOLS_BUILD
  debug = g_debug;
  switch (g_emu >= 0 ? g_emu : kDefaultEmu) {
    case 0: kernel = debug ? prefill_tcgen05_kernel<true, 0> : prefill_tcgen05_kernel<false, 0>; break;
    case 1: kernel = prefill_tcgen05_kernel<false, 1>; break;
    case 2: kernel = prefill_tcgen05_kernel<false, 2>; break;
    case 3: kernel = prefill_tcgen05_kernel<false, 3>; break;
    case 4: kernel = prefill_tcgen05_kernel<false, 4>; break;
    default: kernel = prefill_tcgen05_kernel<false, 5>; break;
  }
  if (debug && kernel != prefill_tcgen05_kernel<true, 0>) kernel = prefill_tcgen05_kernel<true, kDefaultEmu>;
#endif
  // (the attribute is per function and per device: set it on every launch of a tools build, once in the product)
#ifndef NVLLM_TOOLS_BUILD
  static PerDeviceOnce configured;
  cudaError_t e0 = once_per_device(configured, [kernel] {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  });
#else
  cudaError_t e0 = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
#endif
  if (e0 != cudaSuccess) return record_cuda_error(e0);
  // timeout latch -> mapped host word, once per device (prefill launches are eager: the reference captures decode only);
  // a refusal keeps the device-only latch
  if (tc::tc_error_arm() != cudaSuccess) cudaGetLastError();
  const int num_pairs = pp.pair_heads ? (max_seqlen_q + TM - 1) / TM : (max_seqlen_q + 2 * TM - 1) / (2 * TM);
  const int64_t total = static_cast<int64_t>(num_pairs) * (pp.pair_heads ? pp.num_q_heads / 2 : pp.num_q_heads) * batch;
  if (total > INT32_MAX) return NVLLM_ERR_UNSUPPORTED;
  const int sms = device_sm_count();
  const unsigned grid = static_cast<unsigned>(total < sms ? total : sms);   // persistent: one CTA per SM
  kernel<<<grid, kThreads, kSmemBytes, stream>>>(tq, tk, tv, pp, num_pairs, batch, debug);
  note_launch();
  *handled = true;
  return record_cuda_error(cudaGetLastError());
}

int prefill_tcgen05_error_fetch(unsigned int* code) { return record_cuda_error(tc::tc_error_fetch(code)); }
unsigned int prefill_tcgen05_error_poll() { return tc::tc_error_poll(); }
int prefill_tcgen05_latch_selftest(cudaStream_t stream) { return record_cuda_error(tc::tc_latch_selftest(stream)); }

}  // namespace nvllm

#ifdef NVLLM_TOOLS_BUILD
extern "C" void nvllm_varlen_prefill_set_debug_buffer(void* device_ptr) {
  nvllm::g_debug = static_cast<float*>(device_ptr);
}
extern "C" void nvllm_varlen_prefill_set_exp2_poly(int pairs_of_8) { nvllm::g_emu = pairs_of_8; }
#endif
