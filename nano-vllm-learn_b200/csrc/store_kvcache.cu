// store_kvcache: slot-mapping scatter of this step's K/V rows into the paged cache.
// Replaces the Triton store_kvcache_kernel (reference nanovllm/layers/attention.py:19-41):
// one 16-byte vector of K and the matching vector of V per thread, 64-bit offsets,
// negative slots skipped (superset of attention_triton.py:30-31).
//
// Roofline: pure HBM data movement, N * (4 * row_bytes + 4) algorithmic bytes per launch.
// In decode (N <= 512 rows of 2 KiB) it is launch-latency bound.
#include "common.cuh"

namespace nvllm {

__device__ __forceinline__ int4 ld_stream16(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// vecs_per_row = row_bytes / 16.  Thread t of the flat grid moves vector (t % vecs_per_row)
// of token (t / vecs_per_row) for both K and V: consecutive lanes touch consecutive 16-byte
// pieces of one row, so a warp reads and writes 512 contiguous bytes per tensor.
__global__ void __launch_bounds__(256) store_kvcache_kernel(const char* __restrict__ key, int64_t key_row_bytes,
                                                            const char* __restrict__ value,
                                                            int64_t value_row_bytes, char* __restrict__ k_cache,
                                                            char* __restrict__ v_cache,
                                                            const int32_t* __restrict__ slot_mapping,
                                                            int64_t total_vecs, int32_t vecs_per_row) {
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total_vecs; t += stride) {
    const int64_t token = t / vecs_per_row;
    const int64_t vec = t - token * vecs_per_row;
    const int64_t slot = __ldg(slot_mapping + token);
    if (slot < 0) continue;
    const int4 kv = ld_stream16(key + token * key_row_bytes + vec * 16);
    const int4 vv = ld_stream16(value + token * value_row_bytes + vec * 16);
    const int64_t dst = (slot * vecs_per_row + vec) * 16;
    *reinterpret_cast<int4*>(k_cache + dst) = kv;
    *reinterpret_cast<int4*>(v_cache + dst) = vv;
  }
}

}  // namespace nvllm

extern "C" int nvllm_store_kvcache(const void* key, int64_t key_row_stride, const void* value,
                                   int64_t value_row_stride, void* k_cache, void* v_cache,
                                   const int32_t* slot_mapping, int32_t n_tokens, int32_t num_kv_heads,
                                   int32_t head_dim, int32_t elem_bytes, void* stream) {
  using namespace nvllm;
  if (n_tokens < 0 || num_kv_heads <= 0 || head_dim <= 0 || elem_bytes <= 0) return NVLLM_ERR_BAD_ARG;
  if (n_tokens == 0) return NVLLM_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping) return NVLLM_ERR_BAD_ARG;
  const int64_t row_bytes = static_cast<int64_t>(num_kv_heads) * head_dim * elem_bytes;
  if (key_row_stride < static_cast<int64_t>(num_kv_heads) * head_dim ||
      value_row_stride < static_cast<int64_t>(num_kv_heads) * head_dim)
    return NVLLM_ERR_BAD_ARG;
  const int64_t key_row_bytes = key_row_stride * elem_bytes;
  const int64_t value_row_bytes = value_row_stride * elem_bytes;
  if ((row_bytes & 15) || (key_row_bytes & 15) || (value_row_bytes & 15) || !aligned16(key) || !aligned16(value) ||
      !aligned16(k_cache) || !aligned16(v_cache))
    return NVLLM_ERR_ALIGNMENT;

  const int32_t vecs_per_row = static_cast<int32_t>(row_bytes / 16);
  const int64_t total_vecs = static_cast<int64_t>(n_tokens) * vecs_per_row;
  const int threads = 256;
  // 148 SMs x 8 resident 256-thread CTAs; a grid-stride loop covers larger prefills.
  const int64_t max_blocks = 148 * 8;
  int64_t blocks = (total_vecs + threads - 1) / threads;
  if (blocks > max_blocks) blocks = max_blocks;
  store_kvcache_kernel<<<static_cast<unsigned>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const char*>(key), key_row_bytes, static_cast<const char*>(value), value_row_bytes,
      static_cast<char*>(k_cache), static_cast<char*>(v_cache), slot_mapping, total_vecs, vecs_per_row);
  note_launch();
  return record_cuda_error(cudaGetLastError());
}
