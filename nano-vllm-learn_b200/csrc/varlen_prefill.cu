// Varlen causal prefill attention with optional paged prefix (prefix cache) for sm_100a.
// Replaces flash_attn_varlen_func at reference nanovllm/layers/attention.py:89-96.
// Causal mask is aligned bottom-right: query j of a sequence with Lq new tokens and Lk keys attends
// keys 0 .. (Lk - Lq) + j.  The no-prefix case follows nanovllm/layers/attention_sdpa.py:65-119.
//
// Tensor-core bound: FLOPs = 4 * Hq * 128 * sum_i [Lq_i * cached_i + Lq_i (Lq_i + 1) / 2].
//
// This file holds the C-ABI entry point and the warp-level (mma.sync m16n8k16) tile kernel: 64 query rows x
// 64 keys per step, four warps of 16 rows, K/V tiles double-buffered with cp.async.  The default path is the
// tcgen05/TMEM kernel in prefill_tcgen05.cu; this kernel stays as the cross-check (variant 0) and as the
// path taken when no TMA bounds are given or the driver cannot encode tensor maps.
#include <atomic>

#include "common.cuh"
#include "prefill_params.cuh"

namespace nvllm {

constexpr int kBM = 64;  // query rows per CTA
constexpr int kBN = 64;  // keys per tile
constexpr int kPrefillThreads = 128;
constexpr int kQBytes = kBM * kRowBytes;        // 16 KB
constexpr int kKVTileBytes = kBN * kRowBytes;   // 16 KB
constexpr int kPrefillSmem = kQBytes + 4 * kKVTileBytes;  // Q + 2 x (K, V) = 80 KB

__global__ void __launch_bounds__(kPrefillThreads) varlen_prefill_mma_kernel(const PrefillParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int mb = blockIdx.x;
  const int h = blockIdx.y;
  const int seq = blockIdx.z;
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int lane = tid & 31;

  const int q0 = __ldg(p.cu_q + seq);
  const int Lq = __ldg(p.cu_q + seq + 1) - q0;
  const int k0 = __ldg(p.cu_k + seq);
  const int Lk = __ldg(p.cu_k + seq + 1) - k0;
  const int m0 = mb * kBM;
  if (m0 >= Lq) return;
  const int shift = Lk - Lq;  // query row r sees keys t <= shift + r
  const int kvh = h / p.group;
  // keys needed by this query block: 0 .. min(Lk, shift + m0 + kBM) - 1
  int k_hi = shift + m0 + kBM;
  if (k_hi > Lk) k_hi = Lk;
  const int n_tiles = k_hi > 0 ? (k_hi + kBN - 1) / kBN : 0;

  const uint32_t sq = smem_u32(smem);
  const uint32_t skv = sq + kQBytes;

  // ---- async load of the Q tile (rows >= Lq zero-filled) ------------------------------------
  {
    const int col16 = tid & 15;
#pragma unroll
    for (int j = 0; j < kBM * 16 / kPrefillThreads; ++j) {
      const int row = j * (kPrefillThreads / 16) + (tid >> 4);
      const bool live = m0 + row < Lq;
      const __nv_bfloat16* src = p.q + static_cast<int64_t>(q0 + (live ? m0 + row : 0)) * p.q_row_stride +
                                 static_cast<int64_t>(h) * kHeadDim + col16 * 8;
      cp_async16(sq + swz_off(row, col16), src, live ? 16u : 0u);
    }
  }
  const int32_t* table = p.block_tables ? p.block_tables + static_cast<int64_t>(seq) * p.table_stride : nullptr;
  auto issue_kv = [&](int tile) {
    const uint32_t sk = skv + (tile & 1) * 2 * kKVTileBytes;
    const uint32_t sv = sk + kKVTileBytes;
    const int col16 = tid & 15;
    const int t_base = tile * kBN;
    int64_t kbase, vbase;
    int in_page = 0;
    if (table) {
      const int page = __ldg(table + t_base / p.block_size);
      in_page = t_base % p.block_size;
      kbase = static_cast<int64_t>(page) * p.block_stride;
      vbase = kbase;
    } else {
      kbase = static_cast<int64_t>(k0 + t_base) * p.k_row_stride;
      vbase = static_cast<int64_t>(k0 + t_base) * p.v_row_stride;
    }
    const int64_t hoff = static_cast<int64_t>(kvh) * kHeadDim + col16 * 8;
#pragma unroll
    for (int j = 0; j < kBN * 16 / kPrefillThreads; ++j) {
      const int row = j * (kPrefillThreads / 16) + (tid >> 4);
      const bool live = t_base + row < Lk;
      const int r = live ? in_page + row : in_page;
      cp_async16(sk + swz_off(row, col16), p.k + kbase + static_cast<int64_t>(r) * p.k_row_stride + hoff,
                 live ? 16u : 0u);
      cp_async16(sv + swz_off(row, col16), p.v + vbase + static_cast<int64_t>(r) * p.v_row_stride + hoff,
                 live ? 16u : 0u);
    }
  };

  if (n_tiles > 0) issue_kv(0);
  cp_async_commit();  // group 0: Q + tile 0

  float o[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  uint32_t qa[8][4];

  const int r_lo = m0 + warp * 16 + (lane >> 2);  // query row (within the sequence) of c0/c1
  const int r_hi = r_lo + 8;                      // ... of c2/c3

  for (int tile = 0; tile < n_tiles; ++tile) {
    if (tile + 1 < n_tiles) issue_kv(tile + 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();

    if (tile == 0) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const int j = lane >> 3;
        ldmatrix_x4(qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3],
                    sq + swz_off(warp * 16 + (j & 1) * 8 + (lane & 7), ks * 2 + (j >> 1)));
      }
    }
    const uint32_t sk = skv + (tile & 1) * 2 * kKVTileBytes;
    const uint32_t sv = sk + kKVTileBytes;
    const int t_base = tile * kBN;

    // warp-uniform skip of tiles entirely above the diagonal for this warp's 16 rows
    const int warp_row_hi = m0 + warp * 16 + 15;
    if (t_base <= shift + warp_row_hi) {
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int kk2 = 0; kk2 < 4; ++kk2) {
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          uint32_t r0, r1, r2, r3;
          ldmatrix_x4(r0, r1, r2, r3, sk + swz_off(nt * 8 + (lane & 7), kk2 * 4 + (lane >> 3)));
          mma_bf16_16816(s[nt], qa[kk2 * 2][0], qa[kk2 * 2][1], qa[kk2 * 2][2], qa[kk2 * 2][3], r0, r1);
          mma_bf16_16816(s[nt], qa[kk2 * 2 + 1][0], qa[kk2 * 2 + 1][1], qa[kk2 * 2 + 1][2], qa[kk2 * 2 + 1][3], r2,
                         r3);
        }
      }
      // mask (causal bottom-right + end of sequence) when the tile touches either boundary
      const bool need_mask = (t_base + kBN - 1 > shift + m0 + warp * 16) || (t_base + kBN > Lk);
      if (need_mask) {
        const int lim_lo = min(shift + r_lo, Lk - 1);
        const int lim_hi = min(shift + r_hi, Lk - 1);
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const int t = t_base + nt * 8 + (lane & 3) * 2;
          if (t > lim_lo) s[nt][0] = -INFINITY;
          if (t + 1 > lim_lo) s[nt][1] = -INFINITY;
          if (t > lim_hi) s[nt][2] = -INFINITY;
          if (t + 1 > lim_hi) s[nt][3] = -INFINITY;
        }
      }
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
        mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
      }
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
      mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
      mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
      const float mn0 = fmaxf(m_run[0], mx0);
      const float mn1 = fmaxf(m_run[1], mx1);
      // a row with no visible key yet keeps m = -inf; use 0 as the exponent base to avoid inf - inf
      const float ms0 = (mn0 == -INFINITY) ? 0.f : mn0 * p.scale_log2;
      const float ms1 = (mn1 == -INFINITY) ? 0.f : mn1 * p.scale_log2;
      const float a0 = (mn0 == -INFINITY) ? 1.f : fast_exp2(m_run[0] * p.scale_log2 - ms0);
      const float a1 = (mn1 == -INFINITY) ? 1.f : fast_exp2(m_run[1] * p.scale_log2 - ms1);
      m_run[0] = mn0;
      m_run[1] = mn1;
      float ps0 = 0.f, ps1 = 0.f;
      uint32_t pa[4][4];
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        float e[2][4];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          e[u][0] = fast_exp2(s[2 * ks + u][0] * p.scale_log2 - ms0);
          e[u][1] = fast_exp2(s[2 * ks + u][1] * p.scale_log2 - ms0);
          e[u][2] = fast_exp2(s[2 * ks + u][2] * p.scale_log2 - ms1);
          e[u][3] = fast_exp2(s[2 * ks + u][3] * p.scale_log2 - ms1);
          ps0 += e[u][0] + e[u][1];
          ps1 += e[u][2] + e[u][3];
        }
        pa[ks][0] = pack_bf16x2(e[0][0], e[0][1]);
        pa[ks][1] = pack_bf16x2(e[0][2], e[0][3]);
        pa[ks][2] = pack_bf16x2(e[1][0], e[1][1]);
        pa[ks][3] = pack_bf16x2(e[1][2], e[1][3]);
      }
      l_run[0] = l_run[0] * a0 + ps0;
      l_run[1] = l_run[1] * a1 + ps1;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        o[i][0] *= a0;
        o[i][1] *= a0;
        o[i][2] *= a1;
        o[i][3] *= a1;
      }
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
        for (int c2 = 0; c2 < 8; ++c2) {
          uint32_t r0, r1, r2, r3;
          const int j = lane >> 3;
          ldmatrix_x4_trans(r0, r1, r2, r3, sv + swz_off(ks * 16 + (j & 1) * 8 + (lane & 7), c2 * 2 + (j >> 1)));
          mma_bf16_16816(o[c2 * 2], pa[ks][0], pa[ks][1], pa[ks][2], pa[ks][3], r0, r1);
          mma_bf16_16816(o[c2 * 2 + 1], pa[ks][0], pa[ks][1], pa[ks][2], pa[ks][3], r2, r3);
        }
      }
    }
    __syncthreads();  // tile buffer (tile & 1) is refilled by the next iteration's issue
  }
  cp_async_wait<0>();

  // ---- epilogue --------------------------------------------------------------------------------
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    l_run[u] += __shfl_xor_sync(0xffffffffu, l_run[u], 1);
    l_run[u] += __shfl_xor_sync(0xffffffffu, l_run[u], 2);
  }
  const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f;
  const float inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
  if (r_lo < Lq) {
    __nv_bfloat16* dst = p.out + (static_cast<int64_t>(q0 + r_lo) * p.num_q_heads + h) * kHeadDim + (lane & 3) * 2;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt)
      *reinterpret_cast<uint32_t*>(dst + nt * 8) = pack_bf16x2(o[nt][0] * inv0, o[nt][1] * inv0);
  }
  if (r_hi < Lq) {
    __nv_bfloat16* dst = p.out + (static_cast<int64_t>(q0 + r_hi) * p.num_q_heads + h) * kHeadDim + (lane & 3) * 2;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt)
      *reinterpret_cast<uint32_t*>(dst + nt * 8) = pack_bf16x2(o[nt][2] * inv1, o[nt][3] * inv1);
  }
}

static int g_prefill_variant = -1;
static std::atomic<int> g_last_path{-1};          // 1 = tcgen05, 0 = mma.sync
static std::atomic<int64_t> g_path_count[2];
static void note_path(int path) {
  g_last_path.store(path, std::memory_order_relaxed);
  g_path_count[path].fetch_add(1, std::memory_order_relaxed);
}

}  // namespace nvllm

extern "C" {

int32_t nvllm_varlen_prefill_last_path(void) { return nvllm::g_last_path.load(std::memory_order_relaxed); }
int64_t nvllm_varlen_prefill_path_count(int32_t path) {
  return (path == 0 || path == 1) ? nvllm::g_path_count[path].load(std::memory_order_relaxed) : -1;
}
int nvllm_varlen_prefill_error(uint32_t* code_out) {
  if (!code_out) return NVLLM_ERR_BAD_ARG;
  unsigned int v = 0;
  const int rc = nvllm::prefill_tcgen05_error_fetch(&v);
  *code_out = v;
  return rc;
}

int nvllm_varlen_prefill_poll_error(uint32_t* error_out, int32_t clear) {
  if (!error_out) return NVLLM_ERR_BAD_ARG;
  *error_out = nvllm::prefill_tcgen05_error_poll();
  if (*error_out && clear) {   // error path only: clears the device-side latch as well (synchronises)
    unsigned int v = 0;
    return nvllm::prefill_tcgen05_error_fetch(&v);
  }
  return NVLLM_OK;
}

int nvllm_tc_latch_selftest(int32_t unit, void* stream) {
  if (unit == 0) return nvllm::prefill_tcgen05_latch_selftest(static_cast<cudaStream_t>(stream));
  if (unit == 1) return nvllm::rp_tc_latch_selftest(static_cast<cudaStream_t>(stream));
  return NVLLM_ERR_BAD_ARG;
}

int nvllm_varlen_prefill_set_variant(int32_t variant) {
  if (variant < -1 || variant > 1) return NVLLM_ERR_BAD_ARG;
  nvllm::g_prefill_variant = variant;
  return NVLLM_OK;
}

int nvllm_varlen_prefill(void* out, const void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride,
                         const void* v, int64_t v_row_stride, int64_t cache_block_stride,
                         const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k, int32_t batch,
                         int32_t total_q, int32_t total_k, int32_t max_seqlen_q, int32_t max_seqlen_k,
                         const int32_t* block_tables,
                         int64_t table_stride, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                         int32_t block_size, float scale, void* stream) {
  using namespace nvllm;
  if (batch < 0 || max_seqlen_q < 0 || max_seqlen_k < 0 || num_q_heads <= 0 || num_kv_heads <= 0 || total_q < 0 ||
      total_k < 0)
    return NVLLM_ERR_BAD_ARG;
  if (batch == 0 || max_seqlen_q == 0) return NVLLM_OK;
  if (!out || !q || !k || !v || !cu_seqlens_q || !cu_seqlens_k) return NVLLM_ERR_BAD_ARG;
  if (head_dim != kHeadDim || num_q_heads % num_kv_heads != 0) return NVLLM_ERR_UNSUPPORTED;
  if (block_tables && (block_size <= 0 || block_size % kBN != 0)) return NVLLM_ERR_UNSUPPORTED;
  if (batch > 65535 || num_q_heads > 65535) return NVLLM_ERR_UNSUPPORTED;
  if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(out) || (q_row_stride & 7) ||
      (k_row_stride & 7) || (v_row_stride & 7) || (cache_block_stride & 7))
    return NVLLM_ERR_ALIGNMENT;

  PrefillParams p;
  p.pair_heads = 0;   // chosen by the tcgen05 launcher
  p.q = static_cast<const __nv_bfloat16*>(q);
  p.q_row_stride = q_row_stride;
  p.k = static_cast<const __nv_bfloat16*>(k);
  p.v = static_cast<const __nv_bfloat16*>(v);
  p.k_row_stride = k_row_stride;
  p.v_row_stride = v_row_stride;
  p.block_stride = cache_block_stride;
  p.cu_q = cu_seqlens_q;
  p.cu_k = cu_seqlens_k;
  p.block_tables = block_tables;
  p.table_stride = table_stride;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.num_q_heads = num_q_heads;
  p.num_kv_heads = num_kv_heads;
  p.group = num_q_heads / num_kv_heads;
  p.block_size = block_size > 0 ? block_size : 256;
  p.total_q = total_q;
  p.total_k = total_k;
  p.scale_log2 = scale * kLog2e;

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g_prefill_variant != 0) {  // default (-1) and 1: tcgen05/TMEM kernel; 0 forces the mma.sync tile kernel
    bool handled = false;
    int rc = launch_prefill_tcgen05(p, batch, max_seqlen_q, max_seqlen_k, st, &handled);
    if (rc == NVLLM_OK && handled) note_path(1);
    if (rc != NVLLM_OK || handled) return rc;
  }
  static PerDeviceOnce configured;
  cudaError_t e0 = once_per_device(configured, [] {
    return cudaFuncSetAttribute(varlen_prefill_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPrefillSmem);
  });
  if (e0 != cudaSuccess) return record_cuda_error(e0);
  dim3 grid((max_seqlen_q + kBM - 1) / kBM, num_q_heads, batch);
  varlen_prefill_mma_kernel<<<grid, kPrefillThreads, kPrefillSmem, st>>>(p);
  note_launch();
  note_path(0);
  return record_cuda_error(cudaGetLastError());
}

}  // extern "C"
