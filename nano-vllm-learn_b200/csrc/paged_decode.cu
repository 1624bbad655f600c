// Paged GQA decode attention (one query token per sequence) for sm_100a.
// Replaces flash_attn_with_kvcache at reference nanovllm/layers/attention.py:97-101;
// semantics restated from nanovllm/layers/attention_sdpa.py:122-182.
//
// HBM-bound: every (token, kv-head) K row and V row (256 B each) is read exactly once.
// Algorithmic bytes per launch = sum_b ctx_b * Hkv * 512 + 2 * bs * Hq * 256 + 4*bs + 4*sum_b ceil(ctx_b/256).
//
// Design
//  * grid (num_splits, Hkv, bs): a CTA owns one (sequence, kv head, split).  The split takes an
//    equal share of the LIVE ctx_b read on the device, so launch geometry is capture-time constant
//    while work stays balanced (flash-attn splits the table width instead).
//  * inside the CTA each warp is an independent pipeline: it streams CHUNK-token slices of K and V
//    (CHUNK x 512 B) with 16-byte cp.async into its private STAGES-deep shared-memory ring, indexed
//    through block_tables, with no block-wide barrier in the main loop.
//  * the G query heads sharing the kv head are the (zero-padded) 16 rows of an m16n8k16 bf16 MMA:
//    S = Q.K^T and O += P.V run on the tensor cores with fp32 accumulation; online softmax between
//    them uses warp shuffles only.
//  * warps merge their (m, l, O) in shared memory once; with num_splits > 1 the CTA writes a
//    normalised partial + log2-sum-exp, and the last split CTA of a (sequence, kv head) to arrive
//    (atomic counter in the workspace, self-resetting) merges the splits — no second launch.
#include <cstdlib>

#include "common.cuh"
#include "paged_decode_params.cuh"

namespace nvllm {

struct DecodeParams {
  const __nv_bfloat16* q;
  int64_t q_row_stride;
  const __nv_bfloat16* k_cache;
  const __nv_bfloat16* v_cache;
  int64_t block_stride;  // elements between cache blocks
  int64_t token_stride;  // elements between tokens inside a block
  const int32_t* block_tables;
  int64_t table_stride;
  const int32_t* context_lens;
  __nv_bfloat16* out;  // [bs, Hq, 128]
  float* part_o;       // [bs, Hq, S, 128]
  float* part_lse;     // [bs, Hq, S]   (log2 units)
  int32_t* counters;   // [bs, Hkv] arrival counters of the splits; zero between launches
  int32_t num_q_heads;
  int32_t group;       // G = Hq / Hkv
  int32_t block_size;
  int32_t block_shift; // log2(block_size) when it is a power of two, else -1
  int32_t num_splits;
  int32_t cluster;     // split CTAs per thread-block cluster (1 = no cluster); divides num_splits
  float scale_log2;    // softmax scale * log2(e)
};

// Bytes of the cluster leader's exchange area: one normalised fp32 row (128 values + log2-sum-exp) per member and query row.
constexpr int kXchgRow = kHeadDim + 1;
static inline int cluster_xchg_bytes(int cluster, int group) { return cluster * group * kXchgRow * 4; }

__device__ __forceinline__ uint32_t cluster_map_shared(uint32_t local_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;\n" : "=r"(r) : "r"(local_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;\n" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void cluster_arrive_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}

template <int CHUNK, int WARPS, int STAGES>
struct DecodeCfg {
  static constexpr int kChunk = CHUNK;
  static constexpr int kWarps = WARPS;
  static constexpr int kStages = STAGES;
  static constexpr int kTileBytes = CHUNK * kRowBytes;          // K (or V) bytes of one chunk
  static constexpr int kStageBytes = 2 * kTileBytes;            // K + V
  static constexpr int kWarpBytes = STAGES * kStageBytes;
  static constexpr int kPipeBytes = WARPS * kWarpBytes;
  // merge scratch: per warp 8 rows x 128 fp32 + m[8] + l[8]
  static constexpr int kMergeBytes = WARPS * (8 * kHeadDim * 4 + 64);
  static constexpr int kSmemBytes = kPipeBytes > kMergeBytes ? kPipeBytes : kMergeBytes;
};

#ifdef NVLLM_TOOLS_BUILD
// tools/trace_decode.py: thread 0 of every CTA stamps %globaltimer at the protocol points below into
// trace[cta * 8 + k] (k: 0 entry, 1 first copies issued, 2 first chunk landed, 3 main loop done, 4 partial published,
// 5 exit; cta = (b * Hkv + kvh) * S + split).
static __device__ unsigned long long* g_decode_trace = nullptr;
#define NVLLM_TRACE(k)                                                                                        \
  do {                                                                                                        \
    if (g_decode_trace != nullptr && threadIdx.x == 0) {                                                      \
      unsigned long long t_;                                                                                  \
      asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t_));                                                \
      g_decode_trace[((static_cast<size_t>(blockIdx.z) * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 8 + (k)] = t_; \
    }                                                                                                         \
  } while (0)
#else
#define NVLLM_TRACE(k) do { } while (0)
#endif

template <class Cfg>
__global__ void __launch_bounds__(Cfg::kWarps * 32) paged_decode_kernel(const DecodeParams p) {
  constexpr int CHUNK = Cfg::kChunk;
  constexpr int WARPS = Cfg::kWarps;
  constexpr int STAGES = Cfg::kStages;
  constexpr int NT = CHUNK / 8;    // S n-tiles (8 tokens each)
  constexpr int KS = CHUNK / 16;   // PV k-steps (16 tokens each)
  constexpr int PIECES = CHUNK * 16 / 32;  // 16-byte pieces per lane per tensor per chunk

  extern __shared__ __align__(128) unsigned char smem[];
  NVLLM_TRACE(0);

  const int split = blockIdx.x;
  const int kvh = blockIdx.y;
  const int b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int G = p.group;
  // every member of the cluster has started (its shared memory exists) before anyone pushes into the leader's: phase 0
  // of the cluster barrier is arrived at here and waited for just before the push
  if (p.cluster > 1) cluster_arrive_relaxed();

  const int ctx = __ldg(p.context_lens + b);
  const int n_chunks = (ctx + CHUNK - 1) / CHUNK;
  const int per_split = (n_chunks + p.num_splits - 1) / p.num_splits;
  const int c_begin = split * per_split;
  const int c_end = min(c_begin + per_split, n_chunks);
  // chunks of this warp: c_begin + warp + i * WARPS
  const int n_my = (c_end - c_begin - warp + WARPS - 1) / WARPS;  // may be <= 0

  // ---- Q fragments (A operand, rows >= G and rows 8..15 are zero) -------------------------
  const int qrow = lane >> 2;
  uint32_t qa0[8], qa2[8];
  {
    const bool live = qrow < G;
    const __nv_bfloat16* qp =
        p.q + static_cast<int64_t>(b) * p.q_row_stride + static_cast<int64_t>(kvh * G + (live ? qrow : 0)) * kHeadDim +
        (lane & 3) * 2;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      qa0[ks] = live ? __ldg(reinterpret_cast<const uint32_t*>(qp + ks * 16)) : 0u;
      qa2[ks] = live ? __ldg(reinterpret_cast<const uint32_t*>(qp + ks * 16 + 8)) : 0u;
    }
  }

  float o[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run = -INFINITY;  // running max of raw scores (row = lane>>2)
  float l_run = 0.f;        // per-lane partial of the softmax denominator

  unsigned char* my_smem = smem + warp * Cfg::kWarpBytes;
  const uint32_t my_smem_u32 = smem_u32(my_smem);
  const int32_t* table = p.block_tables + static_cast<int64_t>(b) * p.table_stride;
  const int64_t head_off = static_cast<int64_t>(kvh) * kHeadDim;

  // piece -> (token, 16-byte column) of the chunk: lanes 0..15 cover one 256 B row.
  const int col16 = lane & 15;
  const int tok_lo = lane >> 4;

  // token -> (page slot, row in page): block sizes are powers of two in practice (256); shift/mask instead of a divide
  const int bshift = p.block_shift;
  // Per-lane copy geometry, fixed for the whole kernel: piece j copies token 2j + tok_lo; source addresses advance by a
  // constant byte step, destinations are swizzled offsets whose XOR only depends on (2j + tok_lo) & 7, i.e. on j & 3.
  const char* const k_bytes = reinterpret_cast<const char*>(p.k_cache);
  const int64_t kv_delta = reinterpret_cast<const char*>(p.v_cache) - k_bytes;
  const int64_t step_bytes = 4 * p.token_stride;                       // two tokens further, bf16
  const int64_t lane_bytes = 2 * (static_cast<int64_t>(tok_lo) * p.token_stride + col16 * 8 + head_off);
  uint32_t dsw[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) dsw[j] = swz_off(j * 2 + tok_lo, col16);
  auto issue = [&](int i, int page) {
    const int chunk = c_begin + warp + i * WARPS;
    const int t0 = chunk * CHUNK;
    const int in_page = bshift >= 0 ? (t0 & (p.block_size - 1)) : (t0 % p.block_size);
    const char* kp = k_bytes + lane_bytes +
                     2 * (static_cast<int64_t>(page) * p.block_stride + static_cast<int64_t>(in_page) * p.token_stride);
    const uint32_t sk = my_smem_u32 + (i % STAGES) * Cfg::kStageBytes;
    const uint32_t sv = sk + Cfg::kTileBytes;
    if (t0 + CHUNK <= ctx) {   // full chunk (all but the last one of the sequence): no per-piece predicate
#pragma unroll
      for (int j = 0; j < PIECES; ++j) {
        const uint32_t d = dsw[j & 3] + (j >> 2) * (8 * kRowBytes);
        cp_async16_full(sk + d, kp);
        cp_async16_full(sv + d, kp + kv_delta);
        kp += step_bytes;
      }
    } else {
#pragma unroll
      for (int j = 0; j < PIECES; ++j) {
        const uint32_t nbytes = (t0 + j * 2 + tok_lo < ctx) ? 16u : 0u;   // beyond the context: zero-fill
        const uint32_t d = dsw[j & 3] + (j >> 2) * (8 * kRowBytes);
        cp_async16(sk + d, kp, nbytes);
        cp_async16(sv + d, kp + kv_delta, nbytes);
        kp += step_bytes;
      }
    }
  };
  auto page_of = [&](int i) -> int {
    if (i >= n_my) return 0;
    const int t0 = (c_begin + warp + i * WARPS) * CHUNK;
    return __ldg(table + (bshift >= 0 ? (t0 >> bshift) : (t0 / p.block_size)));
  };

  // ---- prologue: fill STAGES-1 stages -----------------------------------------------------
  int page_next = page_of(0);
#pragma unroll
  for (int i = 0; i < STAGES - 1; ++i) {
    const int page = page_next;
    page_next = page_of(i + 1);
    if (i < n_my) issue(i, page);
    cp_async_commit();
  }

  NVLLM_TRACE(1);
  for (int i = 0; i < n_my; ++i) {
    {
      const int nxt = i + STAGES - 1;
      const int page = page_next;
      page_next = page_of(nxt + 1);
      if (nxt < n_my) issue(nxt, page);
      cp_async_commit();
    }
    cp_async_wait<STAGES - 1>();
    __syncwarp();
    if (i == 0) NVLLM_TRACE(2);

    const uint32_t sk = my_smem_u32 + (i % STAGES) * Cfg::kStageBytes;
    const uint32_t sv = sk + Cfg::kTileBytes;
    const int t0 = (c_begin + warp + i * WARPS) * CHUNK;

    // ---- S = Q . K^T  (16 x CHUNK) ---------------------------------------------------------
    float s[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
    for (int kk2 = 0; kk2 < 4; ++kk2) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        uint32_t r0, r1, r2, r3;
        ldmatrix_x4(r0, r1, r2, r3, sk + swz_off(nt * 8 + (lane & 7), kk2 * 4 + (lane >> 3)));
        mma_bf16_16816(s[nt], qa0[kk2 * 2], 0u, qa2[kk2 * 2], 0u, r0, r1);
        mma_bf16_16816(s[nt], qa0[kk2 * 2 + 1], 0u, qa2[kk2 * 2 + 1], 0u, r2, r3);
      }
    }

    // ---- online softmax on rows 0..7 (this lane: row lane>>2, tokens nt*8 + (lane&3)*2 + {0,1})
    if (t0 + CHUNK > ctx) {   // only the last chunk of the sequence has tokens beyond the context
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int t = t0 + nt * 8 + (lane & 3) * 2;
        if (t >= ctx) s[nt][0] = -INFINITY;
        if (t + 1 >= ctx) s[nt][1] = -INFINITY;
      }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) mx = fmaxf(mx, fmaxf(s[nt][0], s[nt][1]));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m_run, mx);  // finite: every chunk holds at least one live token
    const float m_scaled = m_new * p.scale_log2;
    float psum = 0.f;
    uint32_t pa0[KS], pa2[KS];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const float p00 = fast_exp2(s[2 * ks][0] * p.scale_log2 - m_scaled);
      const float p01 = fast_exp2(s[2 * ks][1] * p.scale_log2 - m_scaled);
      const float p10 = fast_exp2(s[2 * ks + 1][0] * p.scale_log2 - m_scaled);
      const float p11 = fast_exp2(s[2 * ks + 1][1] * p.scale_log2 - m_scaled);
      psum += (p00 + p01) + (p10 + p11);
      pa0[ks] = pack_bf16x2(p00, p01);
      pa2[ks] = pack_bf16x2(p10, p11);
    }
    // the running max settles after a few chunks: rescale O only when some row's max actually moved
    if (__any_sync(0xffffffffu, m_new != m_run)) {
      const float alpha = fast_exp2((m_run - m_new) * p.scale_log2);
      l_run *= alpha;
#pragma unroll
      for (int i2 = 0; i2 < 16; ++i2) {
        o[i2][0] *= alpha;
        o[i2][1] *= alpha;
      }
      m_run = m_new;
    }
    l_run += psum;

    // ---- O += P . V  (16 x 128) ------------------------------------------------------------
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
      for (int c2 = 0; c2 < 8; ++c2) {
        uint32_t r0, r1, r2, r3;
        const int j = lane >> 3;
        ldmatrix_x4_trans(r0, r1, r2, r3, sv + swz_off(ks * 16 + (j & 1) * 8 + (lane & 7), c2 * 2 + (j >> 1)));
        mma_bf16_16816(o[c2 * 2], pa0[ks], 0u, pa2[ks], 0u, r0, r1);
        mma_bf16_16816(o[c2 * 2 + 1], pa0[ks], 0u, pa2[ks], 0u, r2, r3);
      }
    }
    __syncwarp();
  }
  cp_async_wait<0>();
  NVLLM_TRACE(3);

  // ---- merge the warps of this CTA ---------------------------------------------------------
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
  l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
  __syncthreads();  // every warp is done with its pipeline buffers
  float* mo = reinterpret_cast<float*>(smem);                  // [WARPS][8][128]
  float* mm = mo + WARPS * 8 * kHeadDim;                       // [WARPS][8]
  float* ml = mm + WARPS * 8;                                  // [WARPS][8]
  if (qrow < G) {
    float* dst = mo + (warp * 8 + qrow) * kHeadDim + (lane & 3) * 2;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) *reinterpret_cast<float2*>(dst + nt * 8) = make_float2(o[nt][0], o[nt][1]);
    if ((lane & 3) == 0) {
      mm[warp * 8 + qrow] = m_run;
      ml[warp * 8 + qrow] = l_run;
    }
  }
  __syncthreads();

  // With clusters the split CTAs of one (sequence, kv head) merge through the leader's shared memory first: every member
  // pushes its normalised rows (DSMEM stores), one cluster barrier, and the leader alone goes on -- with one partial per
  // CLUSTER for the device-wide merge below, or straight to the output when the cluster holds all the splits.
  const int C = p.cluster;
  const int S = p.num_splits / C;       // partials per (sequence, kv head) in the workspace
  const int slot_of_cta = split / C;
  float* const xchg = reinterpret_cast<float*>(smem + Cfg::kSmemBytes);   // leader: [C][G][129], outside the pipeline rings
  const uint32_t xchg_leader = C > 1 ? cluster_map_shared(smem_u32(xchg), 0) : 0u;
  const int crank = split % C;
  if (C > 1) cluster_wait();
  for (int idx = threadIdx.x; idx < G * kHeadDim; idx += WARPS * 32) {
    const int row = idx >> 7;
    const int d = idx & (kHeadDim - 1);
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) M = fmaxf(M, mm[w * 8 + row]);
    float num = 0.f, den = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int w = 0; w < WARPS; ++w) {
        const float wgt = fast_exp2((mm[w * 8 + row] - M) * p.scale_log2);
        num += wgt * mo[(w * 8 + row) * kHeadDim + d];
        den += wgt * ml[w * 8 + row];
      }
    }
    const float val = den > 0.f ? num / den : 0.f;
    const float lse = den > 0.f ? M * p.scale_log2 + fast_log2(den) : -INFINITY;
    const int h = kvh * G + row;
    if (p.num_splits == 1) {
      p.out[(static_cast<int64_t>(b) * p.num_q_heads + h) * kHeadDim + d] = __float2bfloat16_rn(val);
    } else if (C > 1) {
      const uint32_t dst = xchg_leader + ((crank * G + row) * kXchgRow + d) * 4;
      st_cluster_f32(dst, val);
      if (d == 0) st_cluster_f32(dst + kHeadDim * 4, lse);
    } else {
      const int64_t slot = (static_cast<int64_t>(b) * p.num_q_heads + h) * S + slot_of_cta;
      p.part_o[slot * kHeadDim + d] = val;
      if (d == 0) p.part_lse[slot] = lse;
    }
  }
  if (p.num_splits == 1) return;
  if (C > 1) {
    cluster_sync_all();          // release/acquire at cluster scope: the pushed rows are visible to the leader
    if (crank != 0) return;
    for (int idx = threadIdx.x; idx < G * kHeadDim; idx += WARPS * 32) {
      const int row = idx >> 7;
      const int d = idx & (kHeadDim - 1);
      float L = -INFINITY;
      for (int c = 0; c < C; ++c) L = fmaxf(L, xchg[(c * G + row) * kXchgRow + kHeadDim]);
      float num = 0.f, den = 0.f;
      if (L > -INFINITY) {
#pragma unroll 4
        for (int c = 0; c < C; ++c) {
          const float* src = xchg + (c * G + row) * kXchgRow;
          const float wgt = fast_exp2(src[kHeadDim] - L);
          num += wgt * src[d];
          den += wgt;
        }
      }
      const float val = den > 0.f ? num / den : 0.f;
      const int h = kvh * G + row;
      if (S == 1) {
        p.out[(static_cast<int64_t>(b) * p.num_q_heads + h) * kHeadDim + d] = __float2bfloat16_rn(val);
      } else {
        const int64_t slot = (static_cast<int64_t>(b) * p.num_q_heads + h) * S + slot_of_cta;
        p.part_o[slot * kHeadDim + d] = val;
        if (d == 0) p.part_lse[slot] = den > 0.f ? L + fast_log2(den) : -INFINITY;
      }
    }
    if (S == 1) {
      NVLLM_TRACE(4);
      NVLLM_TRACE(5);
      return;
    }
  }

  // ---- fused combine: the LAST split CTA of this (sequence, kv head) to arrive merges all partials.
  // out = sum_s 2^(lse_s - L) O_s / sum_s 2^(lse_s - L).  Saves the separate combine launch (5-6 us).
  __shared__ int s_is_last;
  __syncthreads();   // every thread's partial stores happen-before thread 0's release below (CTA barrier + cumulativity)
  if (threadIdx.x == 0) {
    int32_t* cnt = p.counters + b * gridDim.y + kvh;
    int old;
    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;\n" : "=r"(old) : "l"(cnt) : "memory");
    s_is_last = (old == S - 1);
    if (s_is_last) *cnt = 0;   // self-resetting: the next launch (or graph replay) starts from zero
  }
  __syncthreads();   // ... and thread 0's acquire happens-before every thread's partial loads
  NVLLM_TRACE(4);
  if (!s_is_last) return;
  // All partial reads are issued with as much memory-level parallelism as the CTA has: thread (group g, row r, float4
  // column c4) owns every `groups`-th split; its first kBatch partial loads are in flight BEFORE the log2-sum-exps are
  // fetched to shared memory, so the whole combine costs about one L2 round trip.  (The pipeline rings are idle by now.)
  {
    constexpr int NTHR = WARPS * 32;
    constexpr int kBatch = 16;
    float* const s_lse = reinterpret_cast<float*>(smem);                    // [G][S], S <= 128
    float4* const s_acc = reinterpret_cast<float4*>(smem + 8 * 128 * 4);    // [groups][G * 32]
    float* const s_den = reinterpret_cast<float*>(smem + 8 * 128 * 4 + NTHR * 16);
    const int64_t row0 = static_cast<int64_t>(b) * p.num_q_heads + kvh * G;
    const int cols = G * 32;
    const int groups = NTHR / cols;     // >= 1 unless a 2-warp variant meets G = 8
    if (groups >= 1) {
      const int g = threadIdx.x / cols, j = threadIdx.x % cols, r = j >> 5, c4 = j & 31;
      const bool mine = g < groups;
      const float4* po = reinterpret_cast<const float4*>(p.part_o + (row0 + r) * S * kHeadDim) + c4;
      float4 v[kBatch];
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        const int sp = g + u * groups;
        v[u] = (mine && sp < S) ? __ldcg(po + static_cast<int64_t>(sp) * (kHeadDim / 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      for (int i = threadIdx.x; i < G * S; i += NTHR) s_lse[i] = __ldcg(p.part_lse + (row0 + i / S) * S + i % S);
      __syncthreads();
      if (mine) {
        float L = -INFINITY;
        for (int sp = 0; sp < S; ++sp) L = fmaxf(L, s_lse[r * S + sp]);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        float den = 0.f;
        if (L > -INFINITY) {   // empty splits hold zeros and -inf: weight 0
#pragma unroll
          for (int u = 0; u < kBatch; ++u) {
            const int sp = g + u * groups;
            const float w = sp < S ? fast_exp2(s_lse[r * S + sp] - L) : 0.f;
            acc.x += w * v[u].x; acc.y += w * v[u].y; acc.z += w * v[u].z; acc.w += w * v[u].w;
            den += w;
          }
          for (int sp0 = g + kBatch * groups; sp0 < S; sp0 += kBatch * groups) {
#pragma unroll
            for (int u = 0; u < kBatch; ++u) {
              const int sp = sp0 + u * groups;
              v[u] = sp < S ? __ldcg(po + static_cast<int64_t>(sp) * (kHeadDim / 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < kBatch; ++u) {
              const int sp = sp0 + u * groups;
              const float w = sp < S ? fast_exp2(s_lse[r * S + sp] - L) : 0.f;
              acc.x += w * v[u].x; acc.y += w * v[u].y; acc.z += w * v[u].z; acc.w += w * v[u].w;
              den += w;
            }
          }
        }
        s_acc[g * cols + j] = acc;
        s_den[g * cols + j] = den;
      }
      __syncthreads();
      if (threadIdx.x < cols) {
        float4 acc = s_acc[threadIdx.x];
        float den = s_den[threadIdx.x];
        for (int gg = 1; gg < groups; ++gg) {
          const float4 t = s_acc[gg * cols + threadIdx.x];
          acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
          den += s_den[gg * cols + threadIdx.x];
        }
        const float inv = den > 0.f ? 1.f / den : 0.f;
        uint2 pk;
        pk.x = pack_bf16x2(acc.x * inv, acc.y * inv);
        pk.y = pack_bf16x2(acc.z * inv, acc.w * inv);
        *reinterpret_cast<uint2*>(p.out + (row0 + r) * kHeadDim + c4 * 4) = pk;
      }
    } else {   // fewer threads than float4 columns (2-warp variants with G = 8): plain strided loop
      for (int i = threadIdx.x; i < G * S; i += NTHR) s_lse[i] = __ldcg(p.part_lse + (row0 + i / S) * S + i % S);
      __syncthreads();
      for (int idx = threadIdx.x; idx < G * kHeadDim; idx += NTHR) {
        const int row = idx >> 7, d = idx & (kHeadDim - 1);
        float L = -INFINITY;
        for (int sp = 0; sp < S; ++sp) L = fmaxf(L, s_lse[row * S + sp]);
        float num = 0.f, den = 0.f;
        if (L > -INFINITY) {
          const float* po = p.part_o + (row0 + row) * S * kHeadDim + d;
#pragma unroll 4
          for (int sp = 0; sp < S; ++sp) {
            const float w = fast_exp2(s_lse[row * S + sp] - L);
            num += w * __ldcg(po + static_cast<int64_t>(sp) * kHeadDim);
            den += w;
          }
        }
        p.out[(row0 + row) * kHeadDim + d] = __float2bfloat16_rn(den > 0.f ? num / den : 0.f);
      }
    }
  }
  NVLLM_TRACE(5);
}

// ---- variants -------------------------------------------------------------------------------
using Cfg0 = DecodeCfg<16, 4, 3>;  //  96 KB: two CTAs per SM
using Cfg1 = DecodeCfg<16, 4, 4>;  // 128 KB
using Cfg2 = DecodeCfg<16, 8, 3>;  // 192 KB
using Cfg3 = DecodeCfg<32, 4, 3>;  // 192 KB
using Cfg4 = DecodeCfg<32, 2, 3>;  //  96 KB: two CTAs per SM
using Cfg5 = DecodeCfg<16, 2, 6>;  //  96 KB: two CTAs per SM
using Cfg6 = DecodeCfg<16, 4, 6>;  // 192 KB
using Cfg7 = DecodeCfg<32, 4, 2>;  // 128 KB
constexpr size_t kCounterBytes = 65536;   // 16384 (sequence, kv head) arrival counters
constexpr int kNumVariants = 8;
// -1 = automatic: <16,8,3> (one 8-warp CTA per SM) when the grid fits one CTA per SM, else <16,4,3>
// (two 4-warp CTAs per SM).  Same warps and bytes in flight per SM; measured in profiles/r01_decode_sweep_*.
static int g_variant = -1;

static int sm_count_cached() { return device_sm_count(); }

// Largest cluster (<= 16 split CTAs of one (sequence, kv head), a divisor of num_splits) whose exchange area fits beside
// the pipeline rings and for which the WHOLE grid is co-resident (cudaOccupancyMaxActiveClusters): a cluster waiting for
// a second wave would idle the SMs its members hold.  1 = no clusters.  `force`: 0 automatic, > 0 that size or none.
static int g_force_cluster = 0;
constexpr int kMaxDecodeSmem = 226 * 1024;   // 227 KB opt-in limit minus the kernel's few static bytes
template <class Cfg>
static int pick_cluster(const DecodeParams& p, int bs, int hkv) {
  if (p.num_splits <= 1 || g_force_cluster == 1) return 1;
  // Automatic mode: the in-cluster merge only pays when one or two (sequence, kv head) items are cut into >= 64 splits
  // (ctx 32768, 4q/1kv tp=8 shard: 12.1 vs 12.8 us at bs 1, 13.9 vs 14.5 at bs 2).  With 4-32 items and 4-32 splits the
  // workspace + last-arriver combine is faster -- 32q/8kv bs 1/2/4: 27.7 / 46.5 / 84.8 us vs 29.6 / 48.9 / 85.8 us in a
  // cluster; 4q/1kv bs 4: 17.9 vs 18.8 (profiles/r02_decode_sweep_c5.jsonl) -- cluster placement delays the ramp of a
  // one-wave grid by more than the merge saves.
  if (g_force_cluster == 0 && p.num_splits < 64) return 1;
  const int64_t ctas = static_cast<int64_t>(p.num_splits) * hkv * bs;
  if (ctas > sm_count_cached()) return 1;
  struct Seen { int dev, c, g, n; };
  static Seen seen[64];
  static int n_seen = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 1;
  // automatic: portable sizes only -- 16 (opt-in size) measured no faster than 8 on B200 and constrains placement
  for (int c = g_force_cluster == 16 ? 16 : 8; c >= 2; c >>= 1) {
    if (g_force_cluster > 1 && c != g_force_cluster) continue;
    if (p.num_splits % c) continue;
    const int smem = Cfg::kSmemBytes + cluster_xchg_bytes(c, p.group);
    if (smem > kMaxDecodeSmem) continue;
    int max_clusters = -1;
    for (int i = 0; i < n_seen; ++i)
      if (seen[i].dev == dev && seen[i].c == c && seen[i].g == p.group) max_clusters = seen[i].n;
    if (max_clusters < 0) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(c * 64, 1, 1);
      cfg.blockDim = dim3(Cfg::kWarps * 32, 1, 1);
      cfg.dynamicSmemBytes = smem;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = c;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, paged_decode_kernel<Cfg>, &cfg) != cudaSuccess) {
        cudaGetLastError();
        n = 0;
      }
      max_clusters = n;
      if (n_seen < 64) seen[n_seen++] = Seen{dev, c, p.group, n};
    }
    if (static_cast<int64_t>(max_clusters) * c >= ctas) return c;
  }
  return 1;
}

template <class Cfg>
static cudaError_t launch_decode(DecodeParams p, int bs, int hkv, cudaStream_t stream) {
  static PerDeviceOnce configured;
  cudaError_t e0 = once_per_device(configured, [] {
    cudaError_t e = cudaFuncSetAttribute(paged_decode_kernel<Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kMaxDecodeSmem);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(paged_decode_kernel<Cfg>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  });
  if (e0 != cudaSuccess) return e0;
  dim3 grid(p.num_splits, hkv, bs);
  p.cluster = pick_cluster<Cfg>(p, bs, hkv);
  if (p.cluster == 1) {
    paged_decode_kernel<Cfg><<<grid, Cfg::kWarps * 32, Cfg::kSmemBytes, stream>>>(p);
    return cudaGetLastError();
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(Cfg::kWarps * 32, 1, 1);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes + cluster_xchg_bytes(p.cluster, p.group);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = p.cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, paged_decode_kernel<Cfg>, p);
}

}  // namespace nvllm

#ifdef NVLLM_TOOLS_BUILD
#include "../../include/nvllm_tools.h"
extern "C" int nvllm_paged_decode_set_trace(void* device_buffer) {
  unsigned long long* p = static_cast<unsigned long long*>(device_buffer);
  return nvllm::record_cuda_error(cudaMemcpyToSymbol(nvllm::g_decode_trace, &p, sizeof(p)));
}
#endif

extern "C" {

size_t nvllm_paged_decode_workspace_bytes(int32_t bs, int32_t num_q_heads, int32_t head_dim, int32_t num_splits) {
  if (bs <= 0 || num_q_heads <= 0 || head_dim <= 0 || num_splits == 1) return 0;
  if (num_splits <= 0) return nvllm::balanced_decode_workspace_bytes(-num_splits);   // balanced grid: batch-independent
  const size_t rows = static_cast<size_t>(bs) * num_q_heads * num_splits;
  // [fixed 64 KiB of per-(sequence, kv head) arrival counters] | partial O | partial LSE.  The counter area
  // sits at a geometry-independent place so one zero-filled workspace can serve calls of any shape.
  return nvllm::kCounterBytes + rows * head_dim * sizeof(float) + rows * sizeof(float);
}

int32_t nvllm_paged_decode_plan(int32_t bs, int32_t num_q_heads, int32_t num_kv_heads, int32_t max_ctx,
                                int32_t sm_count) {
  if (bs <= 0 || num_q_heads <= 0 || num_kv_heads <= 0 || max_ctx <= 0) return 1;
  if (sm_count <= 0) sm_count = nvllm::sm_count_cached();
  // Default: the per-(sequence, kv head, split) grid.  NVLLM_DECODE_PLAN=balanced returns 0 = the persistent balanced grid
  // (paged_decode_balanced.cu), kept as an opt-in: measured on B200 it loses to the fixed grid on every shape of the
  // round-2 sweeps (profiles/r02_decode_sweep.md) -- the SMs stream at different HBM rates, so equal static ranges finish
  // 15-20 % apart, and its segment bookkeeping costs 8-10 us of fixed overhead.
  const char* plan_env = getenv("NVLLM_DECODE_PLAN");   // read per call: callers cache plans, tests flip it
  const int64_t items = static_cast<int64_t>(bs) * num_kv_heads;
  if (plan_env && plan_env[0] == 'b' && bs <= nvllm::kBalancedMaxBatch &&
      num_q_heads <= nvllm::kBalancedMaxGroup * num_kv_heads && max_ctx <= nvllm::kBalancedMaxContext &&
      items * static_cast<int64_t>(sizeof(int32_t)) <= static_cast<int64_t>(nvllm::kCounterBytes) &&
      items * ((max_ctx + 15) / 16) < (int64_t(1) << 31))
    return 0;
  // HBM-bound kernel: what matters is enough CTAs in flight to saturate HBM, not wave quantisation -- measured on B200
  // (profiles/r01_decode_sweep_*.jsonl, profiles/r02_decode_sweep.md), one split is fastest as soon as (sequence,
  // kv-head) items cover ~85 % of the SMs, because every extra split pays the partial write + combine pass.
  if (items * 100 >= static_cast<int64_t>(sm_count) * 85) return 1;
  // Below that: one wave of one 8-warp CTA per SM (floor, not ceil: a few CTAs spilling into a second wave cannot
  // saturate HBM on their own), at most 64 splits (past that the combine of the partials costs more than the extra
  // CTAs stream), >= 256 tokens of the launch-time context bound per split, and a power of two when that keeps >= 3/4
  // of the CTAs: contexts are block multiples, so a power-of-two split count hands every warp the same number of
  // chunks (ctx 32768, 8 items: 16 splits 27.2 us, 18 splits 28.3 us).
  int64_t s = static_cast<int64_t>(sm_count) / items;
  const int64_t by_len = max_ctx / 256 > 1 ? max_ctx / 256 : 1;
  if (s > by_len) s = by_len;
  if (s > 64) s = 64;
  int64_t pow2 = 1;
  while (pow2 * 2 <= s) pow2 *= 2;
  if (pow2 * 4 >= s * 3) s = pow2;
  const int best = static_cast<int>(s < 1 ? 1 : s);
  return best;
}

int nvllm_paged_decode_set_variant(int32_t variant) {
  if (variant < -1 || variant >= nvllm::kNumVariants) return NVLLM_ERR_BAD_ARG;
  nvllm::g_variant = variant;
  return NVLLM_OK;
}
int nvllm_paged_decode_set_tuning(int32_t min_chunks_per_cta, int32_t warps_per_cta) {
  nvllm::balanced_decode_set_min_chunks(min_chunks_per_cta);
  nvllm::balanced_decode_set_warps(warps_per_cta);
  return NVLLM_OK;
}
int32_t nvllm_paged_decode_num_variants(void) { return nvllm::kNumVariants; }
int nvllm_paged_decode_set_cluster(int32_t cluster) {
  if (cluster != 0 && cluster != 1 && cluster != 2 && cluster != 4 && cluster != 8 && cluster != 16) return NVLLM_ERR_BAD_ARG;
  nvllm::g_force_cluster = cluster;
  return NVLLM_OK;
}

int nvllm_paged_decode(void* out, const void* q, int64_t q_row_stride, const void* k_cache, const void* v_cache,
                       int64_t cache_block_stride, int64_t cache_token_stride, const int32_t* block_tables,
                       int64_t table_stride, const int32_t* context_lens, int32_t bs, int32_t num_q_heads,
                       int32_t num_kv_heads, int32_t head_dim, int32_t block_size, float scale, int32_t num_splits,
                       void* workspace, size_t workspace_bytes, void* stream) {
  using namespace nvllm;
  if (bs < 0 || num_q_heads <= 0 || num_kv_heads <= 0 || block_size <= 0 || num_splits < -512)
    return NVLLM_ERR_BAD_ARG;
  if (bs == 0) return NVLLM_OK;
  if (!out || !q || !k_cache || !v_cache || !block_tables || !context_lens) return NVLLM_ERR_BAD_ARG;
  if (head_dim != kHeadDim || num_q_heads % num_kv_heads != 0 || block_size % 32 != 0) return NVLLM_ERR_UNSUPPORTED;
  const int G = num_q_heads / num_kv_heads;
  if (G > 8) return NVLLM_ERR_UNSUPPORTED;
  if (bs > 65535 || num_kv_heads > 65535 || num_splits > 128) return NVLLM_ERR_UNSUPPORTED;
  if (q_row_stride < static_cast<int64_t>(num_q_heads) * head_dim || table_stride < 0) return NVLLM_ERR_BAD_ARG;
  if (!aligned16(q) || !aligned16(out) || !aligned16(k_cache) || !aligned16(v_cache) || (q_row_stride & 7) ||
      (cache_block_stride & 7) || (cache_token_stride & 7))
    return NVLLM_ERR_ALIGNMENT;
  const size_t need = nvllm_paged_decode_workspace_bytes(bs, num_q_heads, head_dim, num_splits);
  if (num_splits != 1 && (!workspace || workspace_bytes < need || !aligned16(workspace))) return NVLLM_ERR_WORKSPACE;

  if (num_splits <= 0) {
    // balanced persistent grid: -num_splits CTAs (0 = one per SM); the chunk index space must fit 31 bits
    if (static_cast<size_t>(bs) * num_kv_heads * sizeof(int32_t) > kCounterBytes) return NVLLM_ERR_UNSUPPORTED;
    if (bs > kBalancedMaxBatch || G > kBalancedMaxGroup || table_stride * block_size > kBalancedMaxContext)
      return NVLLM_ERR_UNSUPPORTED;
    if (static_cast<int64_t>(bs) * num_kv_heads * ((table_stride > 0 ? table_stride : 1) * block_size / 16 + 1) >=
        (int64_t(1) << 31))
      return NVLLM_ERR_UNSUPPORTED;
    BalancedDecodeParams bp;
    bp.q = static_cast<const __nv_bfloat16*>(q);
    bp.q_row_stride = q_row_stride;
    bp.k_cache = static_cast<const __nv_bfloat16*>(k_cache);
    bp.v_cache = static_cast<const __nv_bfloat16*>(v_cache);
    bp.block_stride = cache_block_stride;
    bp.token_stride = cache_token_stride;
    bp.block_tables = block_tables;
    bp.table_stride = table_stride;
    bp.context_lens = context_lens;
    bp.out = static_cast<__nv_bfloat16*>(out);
    bp.part_o = nullptr;
    bp.part_lse = nullptr;
    bp.counters = nullptr;
    bp.bs = bs;
    bp.num_q_heads = num_q_heads;
    bp.num_kv_heads = num_kv_heads;
    bp.group = G;
    bp.block_size = block_size;
    bp.block_shift = -1;
    bp.min_chunks_per_cta = 0;
    bp.scale_log2 = scale * kLog2e;
    const cudaError_t be = launch_balanced_decode(bp, -num_splits, workspace, static_cast<cudaStream_t>(stream));
    note_launch();
    return record_cuda_error(be);
  }

  DecodeParams p;
  p.q = static_cast<const __nv_bfloat16*>(q);
  p.q_row_stride = q_row_stride;
  p.k_cache = static_cast<const __nv_bfloat16*>(k_cache);
  p.v_cache = static_cast<const __nv_bfloat16*>(v_cache);
  p.block_stride = cache_block_stride;
  p.token_stride = cache_token_stride;
  p.block_tables = block_tables;
  p.table_stride = table_stride;
  p.context_lens = context_lens;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.part_o = static_cast<float*>(workspace);
  p.part_lse = nullptr;
  p.counters = nullptr;
  if (num_splits > 1) {
    if (static_cast<size_t>(bs) * num_kv_heads * sizeof(int32_t) > kCounterBytes) return NVLLM_ERR_UNSUPPORTED;
    const size_t rows = static_cast<size_t>(bs) * num_q_heads * num_splits;
    p.counters = static_cast<int32_t*>(workspace);
    p.part_o = reinterpret_cast<float*>(static_cast<char*>(workspace) + kCounterBytes);
    p.part_lse = p.part_o + rows * head_dim;
  }
  p.num_q_heads = num_q_heads;
  p.group = G;
  p.block_size = block_size;
  p.block_shift = -1;
  for (int sh = 0; sh < 31; ++sh)
    if ((1 << sh) == block_size) p.block_shift = sh;
  p.num_splits = num_splits;
  p.cluster = 1;
  p.scale_log2 = scale * kLog2e;

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t e;
  int variant = g_variant;
  if (variant < 0)
    variant = (static_cast<int64_t>(bs) * num_kv_heads * num_splits <= sm_count_cached()) ? 2 : 0;
  switch (variant) {
    case 0: e = launch_decode<Cfg0>(p, bs, num_kv_heads, st); break;
    case 1: e = launch_decode<Cfg1>(p, bs, num_kv_heads, st); break;
    case 2: e = launch_decode<Cfg2>(p, bs, num_kv_heads, st); break;
    case 3: e = launch_decode<Cfg3>(p, bs, num_kv_heads, st); break;
    case 4: e = launch_decode<Cfg4>(p, bs, num_kv_heads, st); break;
    case 5: e = launch_decode<Cfg5>(p, bs, num_kv_heads, st); break;
    case 6: e = launch_decode<Cfg6>(p, bs, num_kv_heads, st); break;
    default: e = launch_decode<Cfg7>(p, bs, num_kv_heads, st); break;
  }
  note_launch();
  return record_cuda_error(e);
}

}  // extern "C"

// ---- nvllm_paged_decode_noalloc: the split-KV decode for callers that cannot allocate -----------------------------
// Declaration, limits and the reasons for it: include/nvllm_attn.h (2b).  No counterpart in the reference.
//
// The split-KV scratch (arrival counters + fp32 partials) sits in the library image instead of a caller buffer.  On the
// host, `g_noalloc_scratch` is only a handle of the symbol, not an address the GPU can use: handing it to a kernel would
// fault (or, where the GPU can reach host memory, scribble over the host-side shadow).  The device address is fetched
// with cudaGetSymbolAddress, once per device, with that device current, and kept in g_noalloc_addr.
namespace nvllm {

static __device__ __align__(256) unsigned char g_noalloc_scratch[NVLLM_PAGED_DECODE_NOALLOC_SCRATCH_BYTES];
static void* g_noalloc_addr[kMaxDevices] = {};

// Device address of the scratch array on the current device.  The look-up happens on the first call per device (and on
// every call for device ordinals >= kMaxDevices, which have no cache slot).
static cudaError_t noalloc_scratch_address(void** addr_out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const bool cached = dev >= 0 && dev < kMaxDevices;
  if (cached && g_noalloc_addr[dev]) {
    *addr_out = g_noalloc_addr[dev];
    return cudaSuccess;
  }
  void* addr = nullptr;
  e = cudaGetSymbolAddress(&addr, g_noalloc_scratch);
  if (e != cudaSuccess) return e;
  if (!addr) return cudaErrorInvalidSymbol;
  if (cached) g_noalloc_addr[dev] = addr;
  *addr_out = addr;
  return cudaSuccess;
}

}  // namespace nvllm

extern "C" int nvllm_paged_decode_noalloc(void* out, const void* q, int64_t q_row_stride, const void* k_cache,
                                          const void* v_cache, int64_t cache_block_stride, int64_t cache_token_stride,
                                          const int32_t* block_tables, int64_t table_stride,
                                          const int32_t* context_lens, int32_t bs, int32_t num_q_heads,
                                          int32_t num_kv_heads, int32_t head_dim, int32_t block_size, float scale,
                                          int32_t num_splits, void* stream) {
  using namespace nvllm;
  if (num_splits < 1) return NVLLM_ERR_BAD_ARG;   // the balanced grid has its own workspace layout: not offered here
  if (bs < 0 || num_q_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0) return NVLLM_ERR_BAD_ARG;
  if (bs == 0) return NVLLM_OK;
  // argument errors that need no GPU are reported before the set-up below touches the CUDA runtime
  if (!out || !q || !k_cache || !v_cache || !block_tables || !context_lens || block_size <= 0) return NVLLM_ERR_BAD_ARG;
  if (head_dim != kHeadDim || num_q_heads % num_kv_heads != 0 || num_q_heads / num_kv_heads > 8 ||
      block_size % 32 != 0 || num_splits > 128)
    return NVLLM_ERR_UNSUPPORTED;
  if (num_splits == 1)   // one split writes the output directly: no scratch, nothing to set up
    return nvllm_paged_decode(out, q, q_row_stride, k_cache, v_cache, cache_block_stride, cache_token_stride,
                              block_tables, table_stride, context_lens, bs, num_q_heads, num_kv_heads, head_dim,
                              block_size, scale, 1, nullptr, 0, stream);
  // Beyond the fixed array: refuse before touching the GPU.  (The counter area alone bounds bs * num_kv_heads at 16384;
  // nvllm_paged_decode checks that one.)
  const size_t need = nvllm_paged_decode_workspace_bytes(bs, num_q_heads, head_dim, num_splits);
  if (need > static_cast<size_t>(NVLLM_PAGED_DECODE_NOALLOC_SCRATCH_BYTES)) return NVLLM_ERR_WORKSPACE;
  const size_t counter_bytes = static_cast<size_t>(bs) * num_kv_heads * sizeof(int32_t);
  if (counter_bytes > kCounterBytes) return NVLLM_ERR_UNSUPPORTED;

  void* scratch = nullptr;
  cudaError_t e = noalloc_scratch_address(&scratch);
  if (e != cudaSuccess) return record_cuda_error(e);
  // The array is shared by every call and keeps the last call's bytes.  Only the arrival counters have to start at
  // zero (a launch that died half-way would leave them non-zero); partials are always written before they are read.
  e = cudaMemsetAsync(scratch, 0, counter_bytes, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return record_cuda_error(e);
  return nvllm_paged_decode(out, q, q_row_stride, k_cache, v_cache, cache_block_stride, cache_token_stride,
                            block_tables, table_stride, context_lens, bs, num_q_heads, num_kv_heads, head_dim,
                            block_size, scale, num_splits, scratch, NVLLM_PAGED_DECODE_NOALLOC_SCRATCH_BYTES, stream);
}
