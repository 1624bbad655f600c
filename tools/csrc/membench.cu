// Read-bandwidth probe used to bound the decode kernel (tools/membench.py): streams a buffer with either the
// decode kernel's access shape (256-byte (token, kv-head) rows at a token stride of Hkv * 256 bytes, one
// head per CTA) or fully contiguous 512-byte warp reads.  Not on the product path.
#include "../../nano-vllm-learn_b200/csrc/common.cuh"
#include "../../include/nvllm_tools.h"

namespace nvllm {

__global__ void __launch_bounds__(256) stream_read_kernel(const int4* __restrict__ buf, int64_t n_vec16,
                                                          int pattern, int hkv, int* __restrict__ sink) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  int acc = 0;
  if (pattern == 0) {
    // contiguous: warp w reads vectors [i*32, i*32+32) for i = w, w + nwarps, ...
    for (int64_t i = warp; i * 32 + 31 < n_vec16; i += nwarps) {
      int4 v;
      asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];\n"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                   : "l"(buf + i * 32 + lane));
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
  } else {
    // paged-decode shape: the buffer is [tokens][hkv][16 vec16]; a warp owns head (w % hkv) and reads
    // two consecutive tokens' 256-byte rows per instruction (lanes 0-15 / 16-31), token stride hkv*256 B
    const int64_t tokens = n_vec16 / (16 * hkv);
    const int head = static_cast<int>(warp % hkv);
    const int64_t wslot = warp / hkv, wslots = nwarps / hkv;
    for (int64_t t = wslot * 2; t + 1 < tokens; t += wslots * 2) {
      const int64_t tok = t + (lane >> 4);
      int4 v;
      asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];\n"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                   : "l"(buf + (tok * hkv + head) * 16 + (lane & 15)));
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
    }
  }
  if (acc == 0x7fffffff) *sink = acc;  // keep the loads alive
}

}  // namespace nvllm

extern "C" int nvllm_debug_stream_read(const void* buf, int64_t bytes, int32_t pattern, int32_t hkv, int32_t blocks,
                                       void* sink, void* stream) {
  using namespace nvllm;
  if (!buf || !sink || bytes <= 0 || hkv <= 0 || blocks <= 0) return NVLLM_ERR_BAD_ARG;
  stream_read_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const int4*>(buf), bytes / 16, pattern, hkv, static_cast<int*>(sink));
  return record_cuda_error(cudaGetLastError());
}
