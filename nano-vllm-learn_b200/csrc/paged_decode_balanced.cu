// Balanced ("stream-K") paged GQA decode for sm_100a: the default form of nvllm_paged_decode.
// Replaces flash_attn_with_kvcache at reference nanovllm/layers/attention.py:97-101 (semantics restated from
// nanovllm/layers/attention_sdpa.py:122-182), like csrc/paged_decode.cu, whose per-(sequence, kv head, split) grid
// it supersedes wherever that grid does not tile the 148 SMs evenly (tensor-parallel shards: bs*Hkv = 128..512 CTAs
// on 148/296 slots; long contexts at small batch; ragged batches).
//
// Work = the concatenation, over items (sequence b, kv head h) in (b, h) order, of the item's CHUNK-token slices of
// K and V; T chunks in all, computed ON THE DEVICE from the live context_lens (so the launch geometry is
// capture-time constant).  A persistent grid of P CTAs (one per SM) cuts [0, T) into P equal contiguous ranges;
// a CTA's range is a few SEGMENTS (its intersection with consecutive items).  Inside a CTA each of the 8 warps is an
// independent cp.async pipeline (private 3-stage shared-memory ring, 16-byte copies, no block-wide barrier) taking
// every 8th chunk of the range, and keeps streaming ACROSS segment boundaries: at the end of a segment a warp parks
// its (m, l, O) in shared memory and moves on; the last warp to arrive merges the 8 states and either
//   * writes the final bf16 rows (the segment is a whole item), or
//   * writes one fp32 partial + log2-sum-exp for the CTA (at most two per CTA: the head of its range continuing an
//     earlier CTA's item, the tail starting an item a later CTA continues) and bumps the item's arrival counter;
//     the CTA that completes an item's count merges its partials at the end of its own range, with all 256 threads.
// So every SM streams the same number of bytes whatever (bs, Hkv, context lengths) are, the fp32 partial traffic is
// <= 2 x 4 KB per CTA, the workspace (<= 1.3 MB) does not depend on the batch, and there is no second launch.
// QK^T and PV run on mma.sync m16n8k16 (the G query heads of a kv head are the zero-padded 16 rows); HBM-bound:
// algorithmic bytes per launch = sum_b ctx_b * Hkv * 512 + 2 * bs * Hq * 256 + 4*bs + 4*sum_b ceil(ctx_b/256).
#include "paged_decode_params.cuh"

namespace nvllm {

namespace {

constexpr int BK_CHUNK = 16;   // tokens per pipeline step of a warp
constexpr int BK_STAGES = 3;
constexpr int BK_TILE_BYTES = BK_CHUNK * kRowBytes;           // K (or V) of one chunk: 4 KB
constexpr int BK_STAGE_BYTES = 2 * BK_TILE_BYTES;             // K + V
constexpr int BK_WARP_BYTES = BK_STAGES * BK_STAGE_BYTES;     // 24 KB
constexpr int BK_CTRL_BYTES = 256;
constexpr int BK_MAX_WARPS = 8;

// WARPS warps per CTA, ROWS query rows (>= G) parked per warp.  <8, 8>: one CTA per SM (225 KB).  <4, 4>: two CTAs
// per SM (105 KB each) for G <= 4 -- two independent CTAs hide each other's segment switches and ramps.
template <int WARPS, int ROWS>
struct BKCfg {
  static constexpr int kWarps = WARPS;
  static constexpr int kRows = ROWS;
  static constexpr int kThreads = WARPS * 32;
  static constexpr int kPipeBytes = WARPS * BK_WARP_BYTES;
  static constexpr int kScratchFloats = WARPS * ROWS * kHeadDim + 2 * WARPS * 8;   // O rows + m + l of every warp
  static constexpr int kSmemBytes = kPipeBytes + kScratchFloats * 4 + BK_CTRL_BYTES;
};

struct Ctrl {             // shared-memory control block
  int arrive;             // warps that parked the current segment
  int merged;             // segments merged so far (monotone; the scratch is free for segment s once merged >= s)
  int found_b, found_excl;
  int warp_sum[BK_MAX_WARPS];
  int pending[2][4];      // deferred cross-CTA combines: {valid, b, kvh, unused}
  int total;
};

struct Cursor {           // one non-empty item in chunk space
  int b, kvh;
  int start;              // global index of the item's first chunk
  int nb;                 // chunks in the item
  int ctx;                // tokens in the item
};

__device__ __forceinline__ int chunks_of(int ctx) { return ctx > 0 ? (ctx + BK_CHUNK - 1) / BK_CHUNK : 0; }

}  // namespace

template <class Cfg>
__global__ void __launch_bounds__(Cfg::kThreads) paged_decode_balanced_kernel(const BalancedDecodeParams p) {
  constexpr int BK_WARPS = Cfg::kWarps;
  constexpr int BK_THREADS = Cfg::kThreads;
  constexpr int BK_ROWS = Cfg::kRows;
  constexpr int BK_PIPE_BYTES = Cfg::kPipeBytes;
  extern __shared__ __align__(128) unsigned char smem[];
  float* const mo = reinterpret_cast<float*>(smem + BK_PIPE_BYTES);    // [WARPS][ROWS][128]
  float* const mm = mo + BK_WARPS * BK_ROWS * kHeadDim;                // [WARPS][8] raw running max
  float* const ml = mm + BK_WARPS * 8;                                 // [WARPS][8] softmax denominator
  Ctrl* const ctrl = reinterpret_cast<Ctrl*>(smem + BK_PIPE_BYTES + Cfg::kScratchFloats * 4);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int P = gridDim.x, c = blockIdx.x;
  const int G = p.group, Hkv = p.num_kv_heads;

  // ---------------------------------------------------------------- 1. T = total chunks (live context lengths)
  if (tid == 0) {
    ctrl->arrive = 0;
    ctrl->merged = 0;
    ctrl->pending[0][0] = 0;
    ctrl->pending[1][0] = 0;
  }
  int local = 0;
  for (int b = tid; b < p.bs; b += BK_THREADS) local += chunks_of(__ldg(p.context_lens + b));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if (lane == 0) ctrl->warp_sum[warp] = local;
  __syncthreads();
  if (tid == 0) {
    int s = 0;
#pragma unroll
    for (int w = 0; w < BK_WARPS; ++w) s += ctrl->warp_sum[w];
    ctrl->total = s;
  }
  // rows without keys (zero-padded rows of a captured batch, model_runner.py:291-293) belong to no range: zeros,
  // written by CTA (b mod P), all threads
  for (int b = c; b < p.bs; b += P) {
    if (__ldg(p.context_lens + b) > 0) continue;
    uint4* dst = reinterpret_cast<uint4*>(p.out + static_cast<int64_t>(b) * p.num_q_heads * kHeadDim);
    for (int i = tid; i < p.num_q_heads * (kHeadDim / 8); i += BK_THREADS) dst[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  __syncthreads();
  const int T = ctrl->total * Hkv;
  if (T <= 0) return;
  int P_eff = (T + p.min_chunks_per_cta - 1) / p.min_chunks_per_cta;
  P_eff = P_eff < 1 ? 1 : (P_eff > P ? P : P_eff);
  if (c >= P_eff) return;
  const int lo = static_cast<int>(static_cast<int64_t>(c) * T / P_eff);
  const int hi = static_cast<int>(static_cast<int64_t>(c + 1) * T / P_eff);
  if (lo >= hi) return;
  // CTA that owns chunk x: the largest c' with floor(c' * T / P_eff) <= x
  auto cta_of = [&](int x) -> int { return static_cast<int>((static_cast<int64_t>(x + 1) * P_eff - 1) / T); };

  // ---------------------------------------------------------------- 2. locate the item that contains chunk `lo`
  {
    int base = 0;
    for (int t0 = 0; t0 < p.bs; t0 += BK_THREADS) {
      const int b = t0 + tid;
      const int nb = b < p.bs ? chunks_of(__ldg(p.context_lens + b)) : 0;
      int v = nb;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
      }
      __syncthreads();   // warp_sum of the previous tile (or of step 1) has been consumed
      if (lane == 31) ctrl->warp_sum[warp] = v;
      __syncthreads();
      int wbase = 0, tile_total = 0;
#pragma unroll
      for (int w = 0; w < BK_WARPS; ++w) {
        const int s = ctrl->warp_sum[w];
        if (w < warp) wbase += s;
        tile_total += s;
      }
      const int incl = base + wbase + v, excl = incl - nb;
      if (nb > 0 && Hkv * excl <= lo && lo < Hkv * incl) {
        ctrl->found_b = b;
        ctrl->found_excl = excl;
      }
      base += tile_total;
      if (static_cast<int64_t>(Hkv) * base > lo) break;   // CTA-uniform
    }
    __syncthreads();
  }

  // ---------------------------------------------------------------- 3. per-warp streaming over the CTA's range
  auto load_ctx = [&](int b) -> int { return b < p.bs ? __ldg(p.context_lens + b) : 0; };
  auto advance = [&](Cursor& cur) {   // next NON-EMPTY item (callers only advance while chunks remain below T)
    cur.start += cur.nb;
    if (++cur.kvh == Hkv) {
      cur.kvh = 0;
      do {
        ++cur.b;
        cur.ctx = load_ctx(cur.b);
        cur.nb = chunks_of(cur.ctx);
      } while (cur.nb == 0 && cur.b < p.bs);
    }
  };
  Cursor cc;   // compute side
  cc.b = ctrl->found_b;
  cc.ctx = load_ctx(cc.b);
  cc.nb = chunks_of(cc.ctx);
  cc.kvh = (lo - Hkv * ctrl->found_excl) / cc.nb;
  cc.start = Hkv * ctrl->found_excl + cc.kvh * cc.nb;
  Cursor ci = cc;   // issue side, runs STAGES-1 chunks ahead

  unsigned char* const my_smem = smem + warp * BK_WARP_BYTES;
  const uint32_t my_smem_u32 = smem_u32(my_smem);
  const int col16 = lane & 15;     // 16-byte column of a 256-byte row
  const int tok_lo = lane >> 4;    // lanes 0-15 copy even tokens of a pair, 16-31 odd ones
  constexpr int PIECES = BK_CHUNK * 16 / 32;

  int xi = lo + warp;       // next chunk to issue
  int st_issue = 0;         // ring stage of the next chunk to issue / to compute (no modulo in the loop)
  int page_next = 0;
  // token -> (page slot, row in page): block sizes are powers of two in practice (256); shift/mask instead of a divide
  const int bshift = p.block_shift;
  auto page_slot = [&](int t0) -> int { return bshift >= 0 ? (t0 >> bshift) : (t0 / p.block_size); };
  auto row_in_page = [&](int t0) -> int { return bshift >= 0 ? (t0 & (p.block_size - 1)) : (t0 % p.block_size); };
  auto lookup = [&](const Cursor& cur, int x) -> int {
    return __ldg(p.block_tables + static_cast<int64_t>(cur.b) * p.table_stride + page_slot((x - cur.start) * BK_CHUNK));
  };
  if (xi < hi) {
    while (xi >= ci.start + ci.nb) advance(ci);
    page_next = lookup(ci, xi);
  }
  // Per-lane copy geometry, fixed for the whole kernel: lane = (16-byte column, even/odd token of a pair); piece j
  // copies token 2j + tok_lo.  Source addresses advance by a constant byte step, destinations are the swizzled
  // offsets below (the XOR only depends on (2j + tok_lo) & 7, i.e. on j & 3).
  const char* const k_bytes = reinterpret_cast<const char*>(p.k_cache);
  const int64_t kv_delta = reinterpret_cast<const char*>(p.v_cache) - k_bytes;
  const int64_t step_bytes = 4 * p.token_stride;                       // two tokens further, bf16
  const int64_t lane_bytes = 2 * (static_cast<int64_t>(tok_lo) * p.token_stride + col16 * 8);
  uint32_t dsw[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) dsw[j] = swz_off(j * 2 + tok_lo, col16);
  auto issue_next = [&]() {
    if (xi < hi) {
      const int t0 = (xi - ci.start) * BK_CHUNK;
      const char* kp = k_bytes + lane_bytes +
                       2 * (static_cast<int64_t>(page_next) * p.block_stride + static_cast<int64_t>(ci.kvh) * kHeadDim +
                            static_cast<int64_t>(row_in_page(t0)) * p.token_stride);
      const uint32_t sk = my_smem_u32 + st_issue * BK_STAGE_BYTES;
      const uint32_t sv = sk + BK_TILE_BYTES;
      if (t0 + BK_CHUNK <= ci.ctx) {   // full chunk (all but the last one of an item): no per-piece predicate
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
          const uint32_t nbytes = (t0 + j * 2 + tok_lo < ci.ctx) ? 16u : 0u;   // beyond the context: zero-fill
          const uint32_t d = dsw[j & 3] + (j >> 2) * (8 * kRowBytes);
          cp_async16(sk + d, kp, nbytes);
          cp_async16(sv + d, kp + kv_delta, nbytes);
          kp += step_bytes;
        }
      }
      xi += BK_WARPS;
      if (xi < hi) {   // position the cursor and fetch the page id one step ahead of its use
        while (xi >= ci.start + ci.nb) advance(ci);
        page_next = lookup(ci, xi);
      }
    }
    st_issue = st_issue == BK_STAGES - 1 ? 0 : st_issue + 1;
    cp_async_commit();
  };
#pragma unroll
  for (int s = 0; s < BK_STAGES - 1; ++s) issue_next();

  const int qrow = lane >> 2;
  int xc = lo + warp;       // next chunk to compute
  int st_comp = 0;
  int seg_idx = 0;
  for (;;) {
    const int seg_x0 = max(lo, cc.start);
    const int seg_x1 = min(hi, cc.start + cc.nb);
    const bool has_work = xc < seg_x1;

    // ---- Q fragments of this item (A operand; rows >= G and rows 8..15 are zero)
    uint32_t qa0[8], qa2[8];
    if (has_work) {
      const bool live = qrow < G;
      const __nv_bfloat16* qp = p.q + static_cast<int64_t>(cc.b) * p.q_row_stride +
                                static_cast<int64_t>(cc.kvh * G + (live ? qrow : 0)) * kHeadDim + (lane & 3) * 2;
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        qa0[ks] = live ? __ldg(reinterpret_cast<const uint32_t*>(qp + ks * 16)) : 0u;
        qa2[ks] = live ? __ldg(reinterpret_cast<const uint32_t*>(qp + ks * 16 + 8)) : 0u;
      }
    } else {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) qa0[ks] = qa2[ks] = 0u;
    }
    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_run = -INFINITY;   // running max of raw scores (row = lane >> 2)
    float l_run = 0.f;         // per-lane partial of the softmax denominator

    while (xc < seg_x1) {
      issue_next();
      cp_async_wait<BK_STAGES - 1>();
      __syncwarp();
      const uint32_t sk = my_smem_u32 + st_comp * BK_STAGE_BYTES;
      const uint32_t sv = sk + BK_TILE_BYTES;
      const int t0 = (xc - cc.start) * BK_CHUNK;
      st_comp = st_comp == BK_STAGES - 1 ? 0 : st_comp + 1;

      // S = Q . K^T  (16 x 16)
      float s[2][4];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int kk2 = 0; kk2 < 4; ++kk2) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          uint32_t r0, r1, r2, r3;
          ldmatrix_x4(r0, r1, r2, r3, sk + swz_off(nt * 8 + (lane & 7), kk2 * 4 + (lane >> 3)));
          mma_bf16_16816(s[nt], qa0[kk2 * 2], 0u, qa2[kk2 * 2], 0u, r0, r1);
          mma_bf16_16816(s[nt], qa0[kk2 * 2 + 1], 0u, qa2[kk2 * 2 + 1], 0u, r2, r3);
        }
      }
      // online softmax on rows 0..7 (this lane: row lane>>2, tokens nt*8 + (lane&3)*2 + {0,1})
      if (t0 + BK_CHUNK > cc.ctx) {   // only the last chunk of an item has tokens beyond the context
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const int t = t0 + nt * 8 + (lane & 3) * 2;
          if (t >= cc.ctx) s[nt][0] = -INFINITY;
          if (t + 1 >= cc.ctx) s[nt][1] = -INFINITY;
        }
      }
      float mx = fmaxf(fmaxf(s[0][0], s[0][1]), fmaxf(s[1][0], s[1][1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run, mx);   // finite: every chunk holds at least one live token
      const float m_scaled = m_new * p.scale_log2;
      const float p00 = fast_exp2(s[0][0] * p.scale_log2 - m_scaled);
      const float p01 = fast_exp2(s[0][1] * p.scale_log2 - m_scaled);
      const float p10 = fast_exp2(s[1][0] * p.scale_log2 - m_scaled);
      const float p11 = fast_exp2(s[1][1] * p.scale_log2 - m_scaled);
      const uint32_t pa0 = pack_bf16x2(p00, p01), pa2 = pack_bf16x2(p10, p11);
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
      l_run += (p00 + p01) + (p10 + p11);
      // O += P . V  (16 x 128)
#pragma unroll
      for (int c2 = 0; c2 < 8; ++c2) {
        uint32_t r0, r1, r2, r3;
        const int j = lane >> 3;
        ldmatrix_x4_trans(r0, r1, r2, r3, sv + swz_off((j & 1) * 8 + (lane & 7), c2 * 2 + (j >> 1)));
        mma_bf16_16816(o[c2 * 2], pa0, 0u, pa2, 0u, r0, r1);
        mma_bf16_16816(o[c2 * 2 + 1], pa0, 0u, pa2, 0u, r2, r3);
      }
      __syncwarp();
      xc += BK_WARPS;
    }

    // ---- park this warp's state of the segment; the last warp to arrive merges the eight ----------------------
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    if (seg_idx > 0) {   // one scratch area: the previous segment must have been merged (it almost always has)
      while (*reinterpret_cast<volatile int*>(&ctrl->merged) < seg_idx) __nanosleep(20);
    }
    __syncwarp();
    if (qrow < G) {
      if (has_work) {
        float* dst = mo + (warp * BK_ROWS + qrow) * kHeadDim + (lane & 3) * 2;
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) *reinterpret_cast<float2*>(dst + nt * 8) = make_float2(o[nt][0], o[nt][1]);
      }
      if ((lane & 3) == 0) {
        mm[warp * 8 + qrow] = has_work ? m_run : -INFINITY;   // -inf: the merge skips this warp's (unwritten) rows
        ml[warp * 8 + qrow] = has_work ? l_run : 0.f;
      }
    }
    __threadfence_block();
    __syncwarp();
    int arrived = 0;
    if (lane == 0) arrived = atomicAdd(&ctrl->arrive, 1);
    arrived = __shfl_sync(0xffffffffu, arrived, 0);
    if (arrived == BK_WARPS - 1) {
      __threadfence_block();
      const bool from_start = seg_x0 == cc.start;
      const bool to_end = seg_x1 == cc.start + cc.nb;
      const bool whole = from_start && to_end;
      const int slot = c * 2 + (from_start ? 1 : 0);   // head partial (continues an earlier CTA's item) : tail partial
      for (int r = 0; r < G; ++r) {
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < BK_WARPS; ++w) M = fmaxf(M, mm[w * 8 + r]);
        float4 num = make_float4(0.f, 0.f, 0.f, 0.f);
        float den = 0.f;
#pragma unroll
        for (int w = 0; w < BK_WARPS; ++w) {
          const float mw = mm[w * 8 + r];
          if (mw > -INFINITY) {
            const float wgt = fast_exp2((mw - M) * p.scale_log2);
            const float4 v = *reinterpret_cast<const float4*>(mo + (w * BK_ROWS + r) * kHeadDim + lane * 4);
            num.x += wgt * v.x; num.y += wgt * v.y; num.z += wgt * v.z; num.w += wgt * v.w;
            den += wgt * ml[w * 8 + r];
          }
        }
        const float inv = den > 0.f ? 1.f / den : 0.f;
        num.x *= inv; num.y *= inv; num.z *= inv; num.w *= inv;
        if (whole) {
          uint2 v;
          v.x = pack_bf16x2(num.x, num.y);
          v.y = pack_bf16x2(num.z, num.w);
          *reinterpret_cast<uint2*>(p.out + (static_cast<int64_t>(cc.b) * p.num_q_heads + cc.kvh * G + r) * kHeadDim +
                                    lane * 4) = v;
        } else {
          *reinterpret_cast<float4*>(p.part_o + (static_cast<int64_t>(slot) * 8 + r) * kHeadDim + lane * 4) = num;
          if (lane == 0) p.part_lse[slot * 8 + r] = den > 0.f ? M * p.scale_log2 + fast_log2(den) : -INFINITY;
        }
      }
      if (!whole) {
        // publish the partial; the CTA that completes the item's count merges it (deferred to the end of its range)
        __threadfence();
        __syncwarp();
        if (lane == 0) {
          const int nparts = cta_of(cc.start + cc.nb - 1) - cta_of(cc.start) + 1;
          int32_t* cnt = p.counters + cc.b * Hkv + cc.kvh;
          const int old = atomicAdd(cnt, 1);
          if (old == nparts - 1) {
            *cnt = 0;   // self-resetting: the next launch (or graph replay) starts from zero
            int* pend = ctrl->pending[from_start ? 1 : 0];
            pend[1] = cc.b;
            pend[2] = cc.kvh;
            pend[3] = cc.start;
            pend[0] = cc.nb;   // > 0 marks the entry valid
          }
        }
      }
      __syncwarp();
      if (lane == 0) {
        ctrl->arrive = 0;
        __threadfence_block();
        *reinterpret_cast<volatile int*>(&ctrl->merged) = seg_idx + 1;
      }
    }
    ++seg_idx;
    if (seg_x1 >= hi) break;
    advance(cc);
  }
  cp_async_wait<0>();

  // ---------------------------------------------------------------- 4. deferred cross-CTA combines (<= 2 per CTA)
  __syncthreads();
  float* const s_lse = reinterpret_cast<float*>(smem);                        // [nparts][G]     (pipeline is idle)
  float4* const s_acc = reinterpret_cast<float4*>(smem + 16384);              // [groups][G*32]
  float* const s_den = reinterpret_cast<float*>(smem + 16384 + BK_THREADS * 16);
  for (int k = 0; k < 2; ++k) {
    const int nb = ctrl->pending[k][0];
    if (nb <= 0) continue;   // CTA-uniform
    __threadfence();         // acquire side of the arrival counter
    const int b = ctrl->pending[k][1], kvh = ctrl->pending[k][2], start = ctrl->pending[k][3];
    const int c_first = cta_of(start), nparts = cta_of(start + nb - 1) - c_first + 1;
    // partial j of the item: j == 0 is the tail slot of c_first, j > 0 the head slot of CTA c_first + j
    auto slot_of = [&](int j) -> int { return (c_first + j) * 2 + (j == 0 ? 1 : 0); };
    for (int i = tid; i < nparts * G; i += BK_THREADS) s_lse[i] = __ldcg(p.part_lse + slot_of(i / G) * 8 + (i % G));
    __syncthreads();
    const int cols = G * 32;                 // float4 columns of the item's G rows
    const int groups = BK_THREADS / cols;    // thread groups splitting the partials between them (G <= 8)
    const int g = tid / cols, j = tid % cols, r = j >> 5, c4 = j & 31;
    if (g < groups) {
      float L = -INFINITY;
      for (int q = 0; q < nparts; ++q) L = fmaxf(L, s_lse[q * G + r]);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float den = 0.f;
      if (L > -INFINITY) {
#pragma unroll 4
        for (int q = g; q < nparts; q += groups) {
          const float w = fast_exp2(s_lse[q * G + r] - L);
          const float4 v =
              __ldcg(reinterpret_cast<const float4*>(p.part_o + (static_cast<int64_t>(slot_of(q)) * 8 + r) * kHeadDim) + c4);
          acc.x += w * v.x; acc.y += w * v.y; acc.z += w * v.z; acc.w += w * v.w;
          den += w;
        }
      }
      s_acc[g * cols + j] = acc;
      s_den[g * cols + j] = den;
    }
    __syncthreads();
    if (tid < cols) {
      float4 acc = s_acc[tid];
      float den = s_den[tid];
      for (int gg = 1; gg < groups; ++gg) {
        const float4 v = s_acc[gg * cols + tid];
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        den += s_den[gg * cols + tid];
      }
      const float inv = den > 0.f ? 1.f / den : 0.f;
      uint2 v;
      v.x = pack_bf16x2(acc.x * inv, acc.y * inv);
      v.y = pack_bf16x2(acc.z * inv, acc.w * inv);
      *reinterpret_cast<uint2*>(p.out + (static_cast<int64_t>(b) * p.num_q_heads + kvh * G + r) * kHeadDim + c4 * 4) = v;
    }
    __syncthreads();   // s_lse / s_acc are reused by the second entry
  }
}

// ---- host side --------------------------------------------------------------------------------------------------
namespace {
constexpr size_t kBalancedCounterBytes = 65536;   // 16384 (sequence, kv head) arrival counters
int g_min_chunks = 0;
int g_force_warps = 0;                            // tuning: 4 | 8 warps per CTA, 0 = automatic
constexpr int kDefaultMinChunks = 64;             // 1024 tokens of one kv head (512 KB of K+V) per CTA at least
constexpr int kMaxBalancedCtas = 512;             // partial slots are laid out for this many CTAs whatever the grid
}  // namespace

size_t balanced_decode_workspace_bytes(int /*ctas*/) {
  return kBalancedCounterBytes + static_cast<size_t>(kMaxBalancedCtas) * 2 * 8 * (kHeadDim + 1) * sizeof(float);
}

void balanced_decode_set_min_chunks(int min_chunks) { g_min_chunks = min_chunks > 0 ? min_chunks : 0; }
void balanced_decode_set_warps(int warps) { g_force_warps = (warps == 4 || warps == 8) ? warps : 0; }

template <class Cfg>
static cudaError_t launch_cfg(const BalancedDecodeParams& p, int ctas, cudaStream_t stream) {
  static PerDeviceOnce configured;
  cudaError_t e = once_per_device(configured, [] {
    return cudaFuncSetAttribute(paged_decode_balanced_kernel<Cfg>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                Cfg::kSmemBytes);
  });
  if (e != cudaSuccess) return e;
  paged_decode_balanced_kernel<Cfg><<<ctas, Cfg::kThreads, Cfg::kSmemBytes, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_balanced_decode(BalancedDecodeParams p, int ctas, void* workspace, cudaStream_t stream) {
  // two 4-warp CTAs per SM when the group fits the smaller scratch (G <= 4: every Qwen3 shape), else one 8-warp CTA
  const bool pair = g_force_warps == 4 || (g_force_warps != 8 && p.group <= 4);
  if (ctas <= 0) ctas = device_sm_count() * (pair && p.group <= 4 ? 2 : 1);
  if (ctas > kMaxBalancedCtas) ctas = kMaxBalancedCtas;
  p.counters = static_cast<int32_t*>(workspace);
  p.part_o = reinterpret_cast<float*>(static_cast<char*>(workspace) + kBalancedCounterBytes);
  p.part_lse = p.part_o + static_cast<size_t>(kMaxBalancedCtas) * 2 * 8 * kHeadDim;
  p.min_chunks_per_cta = g_min_chunks > 0 ? g_min_chunks : kDefaultMinChunks;
  p.block_shift = -1;
  for (int sh = 0; sh < 31; ++sh)
    if ((1 << sh) == p.block_size) p.block_shift = sh;
  if (pair && p.group <= 4) return launch_cfg<BKCfg<4, 4>>(p, ctas, stream);
  return launch_cfg<BKCfg<8, 8>>(p, ctas, stream);
}

}  // namespace nvllm
