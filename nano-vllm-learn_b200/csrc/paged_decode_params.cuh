// Parameters and launcher of the balanced paged-decode kernel (paged_decode_balanced.cu), called from the
// nvllm_paged_decode entry point in paged_decode.cu.
#pragma once

#include "common.cuh"

namespace nvllm {

constexpr int kBalancedMaxGroup = 8;     // query heads per kv head
constexpr int kBalancedMaxBatch = 16384; // bounded by the arrival-counter area only
constexpr int kBalancedMaxContext = 1 << 24;

struct BalancedDecodeParams {
  const __nv_bfloat16* q;
  int64_t q_row_stride;
  const __nv_bfloat16* k_cache;
  const __nv_bfloat16* v_cache;
  int64_t block_stride;   // elements between cache blocks
  int64_t token_stride;   // elements between tokens inside a block
  const int32_t* block_tables;
  int64_t table_stride;
  const int32_t* context_lens;
  __nv_bfloat16* out;     // [bs, Hq, 128]
  float* part_o;          // [512][2][8][128]  one head and one tail partial per CTA
  float* part_lse;        // [512][2][8]       (log2 units)
  int32_t* counters;      // [bs * Hkv] arrival counters of the items' partials; zero between launches
  int32_t bs;
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t group;          // G = Hq / Hkv <= 8
  int32_t block_size;
  int32_t block_shift;    // log2(block_size) when it is a power of two, else -1 (filled by the launcher)
  int32_t min_chunks_per_cta;   // fewer CTAs take part when the live work is small (P_eff = ceil(T / this))
  float scale_log2;       // softmax scale * log2(e)
};

// Workspace bytes of the balanced kernel (arrival counters, partial slots): independent of the launch.
size_t balanced_decode_workspace_bytes(int ctas);
// ctas <= 0: fill the SMs of the current device.  Fills the workspace pointers, max_units, block_shift, min_chunks.
cudaError_t launch_balanced_decode(BalancedDecodeParams p, int ctas, void* workspace, cudaStream_t stream);
// tuning knob (sweeps): chunks (16 tokens of one kv head) a CTA should at least own; <= 0 restores the default
void balanced_decode_set_min_chunks(int min_chunks);
// tuning knob (sweeps): 4 = two 4-warp CTAs per SM (G <= 4 only), 8 = one 8-warp CTA per SM, else automatic
void balanced_decode_set_warps(int warps);

}  // namespace nvllm
