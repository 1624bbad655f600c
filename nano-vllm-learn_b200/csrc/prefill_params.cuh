// Parameters shared by the two varlen-prefill kernels (mma.sync tiles and tcgen05/TMEM tiles).
#pragma once

#include "common.cuh"

namespace nvllm {

struct PrefillParams {
  const __nv_bfloat16* q;
  int64_t q_row_stride;
  const __nv_bfloat16* k;
  const __nv_bfloat16* v;
  int64_t k_row_stride;  // dense: row stride; paged: token stride
  int64_t v_row_stride;
  int64_t block_stride;  // paged only
  const int32_t* cu_q;
  const int32_t* cu_k;
  const int32_t* block_tables;  // nullptr -> dense k/v
  int64_t table_stride;
  __nv_bfloat16* out;
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t group;
  int32_t block_size;
  int32_t total_q;  // rows of q / out
  int32_t total_k;  // dense: rows of k/v; paged: num_blocks * block_size
  float scale_log2;
  int32_t pair_heads;  // tcgen05 kernel: the two query tiles of a work item are two q heads of one kv head (else two row blocks)
};

// prefill_tcgen05.cu: returns NVLLM_OK and sets *handled when it launched the tcgen05 kernel.
int launch_prefill_tcgen05(const PrefillParams& p, int batch, int max_seqlen_q, int max_seqlen_k,
                           cudaStream_t stream, bool* handled);

// returns (and clears) the tcgen05 kernel's mbarrier-timeout latch on the current device; synchronises
int prefill_tcgen05_error_fetch(unsigned int* code);
unsigned int prefill_tcgen05_error_poll();                 // mapped host word: no CUDA call
int prefill_tcgen05_latch_selftest(cudaStream_t stream);   // test hook: a wait that times out after 1 ms
int rp_tc_latch_selftest(cudaStream_t stream);             // the same in rowparallel_allreduce.cu's unit

}  // namespace nvllm
