// Device-side prepare_decode: advance the decode metadata of a running batch by one step IN the
// buffers the captured CUDA graph reads, instead of rebuilding them on the host every generated token.
//
// Reference: ModelRunner.prepare_decode, nanovllm/engine/model_runner.py:244-269 -- per sequence
//     input_ids  = seq.last_token                          (the token sampled in the previous step)
//     positions  = len(seq)                                (sic: not len(seq) - 1, :251)
//     context_lens = len(seq)
//     slot       = seq.block_table[-1] * block_size + seq.last_block_num_tokens - 1   (:254-258, sequence.py:53-65)
//     block_tables row = seq.block_table                   (prepare_block_tables, :160-169)
// built as Python lists over `seqs`, five pinned tensors and five host-to-device copies per token.
// Between two decode steps of the SAME batch every sequence grew by exactly one token (Scheduler.postprocess appends
// the sampled id), so: ids <- the sampler's device output, positions += 1, context_lens += 1, slot recomputed from the
// row's own table; the table itself only changes for the rows whose new length opened a block (BlockManager.may_append
// allocates when len % block_size == 1, block_manager.py:134-159) -- those few (row, column, block id) entries travel
// as kernel arguments.  One launch, no host-to-device copy, nothing read back.
#include "common.cuh"

namespace nvllm {
namespace {

constexpr int kMaxPatches = 16;

struct AdvanceParams {
  int64_t* input_ids;
  int64_t* positions;
  int32_t* slot_mapping;
  int32_t* context_lens;
  int32_t* block_tables;
  int64_t table_row_stride;
  const int64_t* sampled;
  int32_t bs, block_size, n_patch;
  int32_t pad_col_begin, pad_col_end;   // table columns that join the live width this step: -1 like the reference's padding
  int32_t patch_row[kMaxPatches], patch_col[kMaxPatches], patch_val[kMaxPatches];
  uint32_t* host_error;                 // mapped pinned host word (may be null): 1 + row whose table entry was missing
};

// A row whose new token falls into a table column that holds no block (-1 padding: the host forgot a patch) must not
// turn into a negative slot and an out-of-range page read.  The row is parked instead -- slot -1 (store_kvcache
// skips it), context length 0 (the decode kernel writes zeros and reads nothing) -- and the mapped host word tells
// the glue, which raises at its next step (nvllm_advance_decode_state_poll_error: one host memory read).
uint32_t* advance_host_error_word() {
  static uint32_t* word = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* q = nullptr;
    if (cudaHostAlloc(&q, 64, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
      for (int i = 0; i < 16; ++i) static_cast<volatile uint32_t*>(q)[i] = 0u;
      word = static_cast<uint32_t*>(q);
    } else {
      cudaGetLastError();
    }
  }
  return word;
}

__global__ void __launch_bounds__(128) advance_decode_state_kernel(const AdvanceParams p) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= p.bs) return;
  int32_t* row = p.block_tables + static_cast<int64_t>(b) * p.table_row_stride;
  // a row's new block id is written by the row's own thread, so no ordering between threads is needed
  for (int c = p.pad_col_begin; c < p.pad_col_end; ++c) row[c] = -1;
#pragma unroll 1
  for (int i = 0; i < p.n_patch; ++i)
    if (p.patch_row[i] == b) row[p.patch_col[i]] = p.patch_val[i];
  const int32_t len = p.context_lens[b] + 1;
  const int32_t last = len - 1;                 // index of the token being appended
  const int32_t blk = row[last / p.block_size];
  p.input_ids[b] = p.sampled[b];
  p.positions[b] = static_cast<int64_t>(len);   // the reference feeds len(seq), model_runner.py:251
  if (blk < 0) {                                // no block behind the new token: park the row and tell the host
    p.context_lens[b] = 0;
    p.slot_mapping[b] = -1;
    if (p.host_error) {
      *reinterpret_cast<volatile uint32_t*>(p.host_error) = 1u + static_cast<uint32_t>(b);
      __threadfence_system();
    }
    return;
  }
  p.context_lens[b] = len;
  p.slot_mapping[b] = blk * p.block_size + last % p.block_size;
}

}  // namespace
}  // namespace nvllm

extern "C" int32_t nvllm_advance_decode_state_max_patches(void) { return nvllm::kMaxPatches; }

extern "C" int nvllm_advance_decode_state_poll_error(uint32_t* error_out, int32_t clear) {
  if (!error_out) return NVLLM_ERR_BAD_ARG;
  uint32_t* w = nvllm::advance_host_error_word();
  *error_out = w ? *reinterpret_cast<volatile uint32_t*>(w) : 0u;
  if (w && clear) *reinterpret_cast<volatile uint32_t*>(w) = 0u;
  return NVLLM_OK;
}

extern "C" int nvllm_advance_decode_state(int64_t* input_ids, int64_t* positions, int32_t* slot_mapping,
                                          int32_t* context_lens, int32_t* block_tables, int64_t table_row_stride,
                                          const int64_t* sampled_tokens, int32_t bs, int32_t block_size,
                                          const int32_t* patch_rows, const int32_t* patch_cols,
                                          const int32_t* patch_vals, int32_t n_patch, int32_t pad_col_begin,
                                          int32_t pad_col_end, void* stream) {
  using namespace nvllm;
  if (bs == 0) return NVLLM_OK;
  if (!input_ids || !positions || !slot_mapping || !context_lens || !block_tables || !sampled_tokens || bs < 0 ||
      block_size <= 0 || table_row_stride <= 0 || n_patch < 0)
    return NVLLM_ERR_BAD_ARG;
  if (n_patch > kMaxPatches) return NVLLM_ERR_UNSUPPORTED;
  if (pad_col_begin < 0 || pad_col_end > table_row_stride || (pad_col_end > pad_col_begin && pad_col_end - pad_col_begin > 64))
    return NVLLM_ERR_BAD_ARG;
  if (n_patch > 0 && (!patch_rows || !patch_cols || !patch_vals)) return NVLLM_ERR_BAD_ARG;
  AdvanceParams p = {};
  p.input_ids = input_ids;
  p.positions = positions;
  p.slot_mapping = slot_mapping;
  p.context_lens = context_lens;
  p.block_tables = block_tables;
  p.table_row_stride = table_row_stride;
  p.sampled = sampled_tokens;
  p.bs = bs;
  p.block_size = block_size;
  p.n_patch = n_patch;
  p.pad_col_begin = pad_col_begin;
  p.pad_col_end = pad_col_end;
  for (int i = 0; i < n_patch; ++i) {
    if (patch_rows[i] < 0 || patch_rows[i] >= bs || patch_cols[i] < 0 || patch_cols[i] >= table_row_stride) return NVLLM_ERR_BAD_ARG;
    p.patch_row[i] = patch_rows[i];
    p.patch_col[i] = patch_cols[i];
    p.patch_val[i] = patch_vals[i];
  }
  p.host_error = nullptr;
  if (uint32_t* hw = advance_host_error_word()) {
    void* dp = nullptr;
    if (cudaHostGetDevicePointer(&dp, hw, 0) == cudaSuccess) p.host_error = static_cast<uint32_t*>(dp);
    else cudaGetLastError();
  }
  advance_decode_state_kernel<<<(bs + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(p);
  note_launch();
  return record_cuda_error(cudaGetLastError());
}
