// Varlen causal prefill on the 5th-gen tensor cores (sm_100a): tcgen05.mma with TMEM accumulators,
// operands staged by TMA.  Replaces flash_attn_varlen_func (reference nanovllm/layers/attention.py:89-96)
// including the paged-prefix (prefix cache) form: K/V tiles come either from this step's dense k/v or
// from the paged cache through block_tables, both via 2-D tensor maps with 128-byte swizzle.
//
// One CTA = one (sequence, q head, 128-query tile).  256 threads:
//   warps 0-3  softmax warpgroup: one query row per thread (TMEM lane == row)
//   warp  4    TMA producer (one elected lane)
//   warp  5    tcgen05.mma issuer (one elected lane) + TMEM allocator
//   warps 6-7  idle (keep warpgroup alignment)
// Per 128-key tile j:
//   S[j&1] = Q . K_j^T          SS MMA, 8 x (M128 N128 K16), Q and K_j K-major SW128 in smem
//   softmax warps: two passes over their S row (max, then exp2 / row-sum), P (bf16) written back into
//                  the same TMEM columns (P aliases S)
//   O[j&1] = P . V_j            TS MMA: A = P from TMEM, B = V_j MN-major SW128 in smem
//   softmax warps fold O[j&1] into fp32 registers with the online-softmax rescale, one tile late, so the
//   tensor pipe runs QK(j+1) / PV(j) while they work.
// TMEM: S0 S1 O0 O1 = 4 x 128 columns = the whole 512-column TMEM of the SM (one CTA per SM).
#include "prefill_params.cuh"
#include "tcgen05.cuh"

namespace nvllm {

namespace {

constexpr int TM = 128;                      // query rows per CTA
constexpr int TN = 128;                      // keys per tile
constexpr int kHalfBytes = 128 * 128;        // [128 rows][64 bf16] = 16 KB, one SW128 half of a tile
constexpr int kTileBytes = 2 * kHalfBytes;   // 32 KB
constexpr int kStages = 2;                   // K/V ring depth
constexpr int kThreads = 256;
constexpr int kSmemQ = 0;
constexpr int kSmemK = kSmemQ + kTileBytes;
constexpr int kSmemV = kSmemK + kStages * kTileBytes;
constexpr int kSmemBar = kSmemV + kStages * kTileBytes;   // 160 KB
constexpr int kSmemBytes = kSmemBar + 256 + 1024;         // barriers + alignment slack
constexpr uint32_t kTmemCols = 512;

constexpr uint32_t kIdescQK = tc::make_idesc_bf16(TM, TN, /*b_mn_major=*/false);
constexpr uint32_t kIdescPV = tc::make_idesc_bf16(TM, kHeadDim, /*b_mn_major=*/true);

// barrier slots (8 bytes each) inside the barrier block
enum Bar : int {
  kQFull = 0,
  kKFull = 1,               // [kStages]
  kVFull = 1 + kStages,     // [kStages]
  kKVEmpty = 1 + 2 * kStages,  // [kStages]
  kSFull = 1 + 3 * kStages,    // [2]
  kPFull = kSFull + 2,         // [2]
  kOFull = kPFull + 2,         // [2]
  kNumBars = kOFull + 2,
};

__global__ void __launch_bounds__(kThreads, 1)
prefill_tcgen05_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const PrefillParams p, float* __restrict__ debug) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t smem = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SW128 tiles need 1024-byte alignment
  unsigned char* smem_gen = smem_raw + (smem - smem_u32(smem_raw));

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;
  const int mb = static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x);  // heavy (late) tiles first
  const int h = blockIdx.y;
  const int seq = blockIdx.z;

  const int q0 = __ldg(p.cu_q + seq);
  const int Lq = __ldg(p.cu_q + seq + 1) - q0;
  const int k0 = __ldg(p.cu_k + seq);
  const int Lk = __ldg(p.cu_k + seq + 1) - k0;
  const int m0 = mb * TM;
  if (m0 >= Lq) return;  // uniform for the CTA, before any barrier or TMEM allocation
  const int shift = Lk - Lq;
  const int kvh = h / p.group;
  int k_hi = shift + m0 + TM;
  if (k_hi > Lk) k_hi = Lk;
  const int n_tiles = k_hi > 0 ? (k_hi + TN - 1) / TN : 0;

  const uint32_t bar0 = smem + kSmemBar;
  auto bar = [&](int i) -> uint32_t { return bar0 + 8u * static_cast<uint32_t>(i); };
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem_gen + kSmemBar + 8 * kNumBars);

  if (warp == 4 && lane == 0) {
    tc::mbar_init(bar(kQFull), 1);
    for (int s = 0; s < kStages; ++s) {
      tc::mbar_init(bar(kKFull + s), 1);
      tc::mbar_init(bar(kVFull + s), 1);
      tc::mbar_init(bar(kKVEmpty + s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      tc::mbar_init(bar(kSFull + b), 1);
      tc::mbar_init(bar(kPFull + b), 128);
      tc::mbar_init(bar(kOFull + b), 1);
    }
    tc::fence_barrier_init();
  }
  if (warp == 5) {
    tc::tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), kTmemCols);
    tc::tmem_relinquish();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (n_tiles > 0) {
    if (warp == 4) {
      // ================================ TMA producer ================================
      if (lane == 0) {
        tc::tma_prefetch_desc(&tm_q);
        tc::tma_prefetch_desc(&tm_k);
        tc::tma_prefetch_desc(&tm_v);
        tc::mbar_arrive_expect_tx(bar(kQFull), kTileBytes);
        tc::tma_load_2d(smem + kSmemQ, &tm_q, h * kHeadDim, q0 + m0, bar(kQFull));
        tc::tma_load_2d(smem + kSmemQ + kHalfBytes, &tm_q, h * kHeadDim + 64, q0 + m0, bar(kQFull));
        const int32_t* table = p.block_tables ? p.block_tables + static_cast<int64_t>(seq) * p.table_stride : nullptr;
        for (int j = 0; j < n_tiles; ++j) {
          const int st = j % kStages;
          const uint32_t use = static_cast<uint32_t>(j / kStages);
          tc::mbar_wait(bar(kKVEmpty + st), (use & 1u) ^ 1u);
          int row;
          if (table) {
            const int t0 = j * TN;
            row = __ldg(table + t0 / p.block_size) * p.block_size + t0 % p.block_size;
          } else {
            row = k0 + j * TN;
          }
          const uint32_t sk = smem + kSmemK + st * kTileBytes;
          const uint32_t sv = smem + kSmemV + st * kTileBytes;
          tc::mbar_arrive_expect_tx(bar(kKFull + st), kTileBytes);
          tc::tma_load_2d(sk, &tm_k, kvh * kHeadDim, row, bar(kKFull + st));
          tc::tma_load_2d(sk + kHalfBytes, &tm_k, kvh * kHeadDim + 64, row, bar(kKFull + st));
          tc::mbar_arrive_expect_tx(bar(kVFull + st), kTileBytes);
          tc::tma_load_2d(sv, &tm_v, kvh * kHeadDim, row, bar(kVFull + st));
          tc::tma_load_2d(sv + kHalfBytes, &tm_v, kvh * kHeadDim + 64, row, bar(kVFull + st));
        }
      }
      __syncwarp();
    } else if (warp == 5) {
      // ================================ MMA issuer ==================================
      if (lane == 0) {
        auto issue_qk = [&](int j) {
          const int st = j % kStages;
          tc::mbar_wait(bar(kKFull + st), static_cast<uint32_t>(j / kStages) & 1u);
          tc::fence_after_sync();
          const uint32_t sq = smem + kSmemQ;
          const uint32_t sk = smem + kSmemK + st * kTileBytes;
          const uint32_t d = tmem + static_cast<uint32_t>((j & 1) * TN);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint32_t off = (ks >> 2) * kHalfBytes + (ks & 3) * 32;  // 16 bf16 = 32 B inside the swizzle row
            tc::mma_ss(d, tc::make_smem_desc(sq + off, 16, 1024), tc::make_smem_desc(sk + off, 16, 1024), kIdescQK,
                       ks > 0 ? 1u : 0u);
          }
          tc::mma_commit(bar(kSFull + (j & 1)));
        };
        tc::mbar_wait(bar(kQFull), 0);
        issue_qk(0);
        for (int j = 0; j < n_tiles; ++j) {
          if (j + 1 < n_tiles) issue_qk(j + 1);
          const int st = j % kStages;
          const int b = j & 1;
          tc::mbar_wait(bar(kVFull + st), static_cast<uint32_t>(j / kStages) & 1u);
          tc::mbar_wait(bar(kPFull + b), static_cast<uint32_t>(j >> 1) & 1u);
          tc::fence_after_sync();
          const uint32_t sv = smem + kSmemV + st * kTileBytes;
          const uint32_t d = tmem + 2u * TN + static_cast<uint32_t>(b * kHeadDim);
          const uint32_t a = tmem + static_cast<uint32_t>(b * TN);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            // 16 keys = 16 rows of 128 B per k-step; the two 64-dim halves are kHalfBytes apart (LBO)
            tc::mma_ts(d, a + static_cast<uint32_t>(ks * 8), tc::make_smem_desc(sv + ks * 2048, kHalfBytes, 1024),
                       kIdescPV, ks > 0 ? 1u : 0u);
          }
          tc::mma_commit(bar(kOFull + b));
          tc::mma_commit(bar(kKVEmpty + st));
        }
      }
      __syncwarp();
    } else if (warp < 4) {
      // ================================ softmax warpgroup ===========================
      const int row = tid;                       // TMEM lane == query row of the tile
      const int qrow = m0 + row;                 // row inside the sequence
      const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
      const int limit = min(shift + qrow, Lk - 1);   // last visible key of this row
      const float c = p.scale_log2;
      const bool dump = debug != nullptr && mb == 0 && h == 0 && seq == 0;
      float o_acc[kHeadDim];
#pragma unroll
      for (int i = 0; i < kHeadDim; ++i) o_acc[i] = 0.f;
      float m_run = -INFINITY, l_run = 0.f, alpha_prev = 1.f;

      auto fold_o = [&](int j, float alpha) {
        const int b = j & 1;
        tc::mbar_wait(bar(kOFull + b), static_cast<uint32_t>(j >> 1) & 1u);
        tc::fence_after_sync();
        const uint32_t o_addr = tmem + lane_base + 2u * TN + static_cast<uint32_t>(b * kHeadDim);
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t r[32];
          tc::tmem_ld32(o_addr + cc * 32, r);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float v = __uint_as_float(r[i]);
            if (dump && j == 0) debug[TM * TN + row * kHeadDim + cc * 32 + i] = v;
            o_acc[cc * 32 + i] = fmaf(o_acc[cc * 32 + i], alpha, v);
          }
        }
      };

      for (int j = 0; j < n_tiles; ++j) {
        const int b = j & 1;
        const int t_base = j * TN;
        tc::mbar_wait(bar(kSFull + b), static_cast<uint32_t>(j >> 1) & 1u);
        tc::fence_after_sync();
        const uint32_t s_addr = tmem + lane_base + static_cast<uint32_t>(b * TN);
        const bool need_mask = (t_base + TN - 1 > shift + m0) || (t_base + TN > Lk);
        // ---- pass 1: row max
        float mx = -INFINITY;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t r[32];
          tc::tmem_ld32(s_addr + cc * 32, r);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            float s = __uint_as_float(r[i]);
            if (dump && j == 0) debug[row * TN + cc * 32 + i] = s;
            if (need_mask && t_base + cc * 32 + i > limit) s = -INFINITY;
            mx = fmaxf(mx, s);
          }
        }
        const float m_new = fmaxf(m_run, mx);
        const float m_sc = (m_new == -INFINITY) ? 0.f : m_new * c;
        const float alpha = (m_new == -INFINITY) ? 1.f : fast_exp2(m_run * c - m_sc);
        // ---- pass 2: P = exp2(S*c - m*c), row sum, P (bf16) back into the S columns
        float psum = 0.f;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t r[32];
          tc::tmem_ld32(s_addr + cc * 32, r);
          tc::tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float s0 = __uint_as_float(r[2 * i]);
            float s1 = __uint_as_float(r[2 * i + 1]);
            if (need_mask) {
              if (t_base + cc * 32 + 2 * i > limit) s0 = -INFINITY;
              if (t_base + cc * 32 + 2 * i + 1 > limit) s1 = -INFINITY;
            }
            const float p0 = fast_exp2(fmaf(s0, c, -m_sc));
            const float p1 = fast_exp2(fmaf(s1, c, -m_sc));
            psum += p0 + p1;
            pk[i] = pack_bf16x2(p0, p1);
          }
          tc::tmem_st16(s_addr + cc * 16, pk);
        }
        tc::tmem_st_wait();
        tc::fence_before_sync();
        tc::mbar_arrive(bar(kPFull + b));
        l_run = l_run * alpha + psum;
        m_run = m_new;
        // ---- fold the previous tile's O partial (its PV MMA overlapped the passes above)
        if (j > 0) fold_o(j - 1, alpha_prev);
        alpha_prev = alpha;
      }
      fold_o(n_tiles - 1, alpha_prev);

      // ---- epilogue: normalise and store this row (rows beyond the sequence are not stored)
      if (qrow < Lq) {
        const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
        uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<int64_t>(q0 + qrow) * p.num_q_heads + h) * kHeadDim);
#pragma unroll
        for (int i = 0; i < kHeadDim / 8; ++i) {
          uint4 v;
          v.x = pack_bf16x2(o_acc[8 * i + 0] * inv, o_acc[8 * i + 1] * inv);
          v.y = pack_bf16x2(o_acc[8 * i + 2] * inv, o_acc[8 * i + 3] * inv);
          v.z = pack_bf16x2(o_acc[8 * i + 4] * inv, o_acc[8 * i + 5] * inv);
          v.w = pack_bf16x2(o_acc[8 * i + 6] * inv, o_acc[8 * i + 7] * inv);
          dst[i] = v;
        }
      }
    }
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 5) tc::tmem_dealloc(tmem, kTmemCols);
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

float* g_debug = nullptr;

}  // namespace

int launch_prefill_tcgen05(const PrefillParams& p, int batch, int max_seqlen_q, int /*max_seqlen_k*/,
                           cudaStream_t stream, bool* handled) {
  *handled = false;
  if (p.total_q <= 0 || p.total_k <= 0) return NVLLM_OK;  // caller did not give TMA bounds: use the mma.sync kernel
  if (p.block_tables && p.block_size % TN != 0) return NVLLM_OK;
  CUtensorMap tq, tk, tv;
  const uint64_t kv_cols = static_cast<uint64_t>(p.num_kv_heads) * kHeadDim;
  if (!make_map(&tq, p.q, static_cast<uint64_t>(p.num_q_heads) * kHeadDim, p.total_q, p.q_row_stride)) return NVLLM_OK;
  if (!make_map(&tk, p.k, kv_cols, p.total_k, p.k_row_stride)) return NVLLM_OK;
  if (!make_map(&tv, p.v, kv_cols, p.total_k, p.v_row_stride)) return NVLLM_OK;
  static bool configured = false;
  if (!configured) {
    cudaError_t e =
        cudaFuncSetAttribute(prefill_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) return record_cuda_error(e);
    configured = true;
  }
  dim3 grid((max_seqlen_q + TM - 1) / TM, p.num_q_heads, batch);
  prefill_tcgen05_kernel<<<grid, kThreads, kSmemBytes, stream>>>(tq, tk, tv, p, g_debug);
  note_launch();
  *handled = true;
  return record_cuda_error(cudaGetLastError());
}

}  // namespace nvllm

extern "C" void nvllm_varlen_prefill_set_debug_buffer(void* device_ptr) {
  nvllm::g_debug = static_cast<float*>(device_ptr);
}
