// C-ABI bookkeeping + the host-side (plain C) index math of the runner glue.
// Declarations and reference citations: include/nvllm_attn.h.
#include <atomic>

#include "common.cuh"

namespace nvllm {

static std::atomic<int64_t> g_launches{0};
static std::atomic<int> g_last_cuda_error{0};

void note_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int record_cuda_error(cudaError_t e) {
  if (e == cudaSuccess) return NVLLM_OK;
  g_last_cuda_error.store(static_cast<int>(e), std::memory_order_relaxed);
  return NVLLM_ERR_CUDA;
}

int device_sm_count() {
  static int cache[kMaxDevices] = {};
  int dev = 0, v = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev >= 0 && dev < kMaxDevices && cache[dev] > 0) return cache[dev];
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
  if (dev >= 0 && dev < kMaxDevices) cache[dev] = v;
  return v;
}

}  // namespace nvllm

extern "C" {

int nvllm_abi_version(void) { return NVLLM_ABI_VERSION; }

const char* nvllm_status_string(int status) {
  switch (status) {
    case NVLLM_OK: return "ok";
    case NVLLM_ERR_BAD_ARG: return "bad argument";
    case NVLLM_ERR_UNSUPPORTED: return "unsupported shape";
    case NVLLM_ERR_ALIGNMENT: return "pointer or stride not 16-byte aligned";
    case NVLLM_ERR_CUDA: return "CUDA launch failed";
    case NVLLM_ERR_WORKSPACE: return "workspace missing or too small";
    default: return "unknown status";
  }
}

int nvllm_last_cuda_error(void) { return nvllm::g_last_cuda_error.load(std::memory_order_relaxed); }

int64_t nvllm_launch_count(void) { return nvllm::g_launches.load(std::memory_order_relaxed); }
void nvllm_reset_launch_count(void) { nvllm::g_launches.store(0, std::memory_order_relaxed); }

// reference: nanovllm/engine/model_runner.py:254-258 with
// Sequence.num_blocks / last_block_num_tokens (nanovllm/engine/sequence.py:57-65)
int nvllm_prepare_decode_slots(const int32_t* num_tokens, const int32_t* last_block, int32_t n, int32_t block_size,
                               int32_t* slot_mapping_out) {
  if (n < 0 || block_size <= 0) return NVLLM_ERR_BAD_ARG;
  if (n > 0 && (!num_tokens || !last_block || !slot_mapping_out)) return NVLLM_ERR_BAD_ARG;
  for (int32_t i = 0; i < n; ++i) {
    const int64_t len = num_tokens[i];
    if (len <= 0) return NVLLM_ERR_BAD_ARG;
    const int64_t num_blocks = (len + block_size - 1) / block_size;
    const int64_t last_block_num_tokens = len - (num_blocks - 1) * block_size;
    slot_mapping_out[i] = static_cast<int32_t>(static_cast<int64_t>(last_block[i]) * block_size +
                                               last_block_num_tokens - 1);
  }
  return NVLLM_OK;
}

// reference: nanovllm/engine/model_runner.py:212-221
int64_t nvllm_prepare_prefill_slots(const int32_t* block_table, int32_t num_table_entries, int32_t num_tokens,
                                    int32_t num_cached_tokens, int32_t block_size, int32_t* slots_out,
                                    int64_t capacity) {
  if (block_size <= 0 || num_tokens < 0 || num_cached_tokens < 0 || num_table_entries < 0 || capacity < 0)
    return -NVLLM_ERR_BAD_ARG;
  if (num_table_entries == 0) return 0;  // warmup sequences carry no block table (model_runner.py:209-210)
  if (!block_table || (!slots_out && capacity > 0)) return -NVLLM_ERR_BAD_ARG;
  const int64_t num_blocks = (static_cast<int64_t>(num_tokens) + block_size - 1) / block_size;
  const int64_t num_cached_blocks = num_cached_tokens / block_size;
  if (num_blocks > num_table_entries) return -NVLLM_ERR_BAD_ARG;
  const int64_t last_block_num_tokens = num_tokens - (num_blocks - 1) * block_size;
  int64_t n = 0;
  for (int64_t i = num_cached_blocks; i < num_blocks; ++i) {
    const int64_t start = static_cast<int64_t>(block_table[i]) * block_size;
    const int64_t count = (i != num_blocks - 1) ? block_size : last_block_num_tokens;
    if (n + count > capacity) return -NVLLM_ERR_WORKSPACE;
    for (int64_t j = 0; j < count; ++j) slots_out[n + j] = static_cast<int32_t>(start + j);
    n += count;
  }
  return n;
}

}  // extern "C"
