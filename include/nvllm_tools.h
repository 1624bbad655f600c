/*
 * nvllm_tools.h — measurement and debug aids, NOT part of the product ABI.
 *
 * Built by `nano-vllm-learn_b200/build.py --tools` into lib/libnvllm_tools.so: the product sources
 * compiled with -DNVLLM_TOOLS_BUILD (which adds the traced instantiation of the tcgen05 prefill
 * kernel) plus tools/csrc/*.cu.  The tools library exports everything include/nvllm_attn.h
 * declares and, in addition, the two entry points below; tools/*.py load it instead of the
 * product library.  Nothing in the product package or the tests of the product path uses it.
 */
#ifndef NVLLM_TOOLS_H_
#define NVLLM_TOOLS_H_

#include "nvllm_attn.h"

#ifdef __cplusplus
extern "C" {
#endif

/* tools/debug_tcgen05.py, tools/trace_prefill.py: when non-NULL, CTA (q-tile 0, head 0, sequence 0)
 * of the tcgen05 prefill kernel dumps its first raw S tile (128x128 fp32) followed by its first raw
 * O partial (128x128 fp32) to this device buffer, and CTA 0 stamps clock64() at its protocol points
 * behind them (>= 147456 bytes in all). */
NVLLM_API void nvllm_varlen_prefill_set_debug_buffer(void* device_ptr);

/* tools/sweep_prefill.py: of every 8 column pairs of an unmasked S tile, how many take the FMA-pipe cubic
 * instead of MUFU ex2 in the tcgen05 prefill kernel (0..5; < 0 restores the product default). */
NVLLM_API void nvllm_varlen_prefill_set_exp2_poly(int pairs_of_8);

/* tools/trace_decode.py: when non-NULL, thread 0 of every CTA of the fixed-grid paged-decode kernel stamps
 * %globaltimer (ns) at six protocol points into buffer[cta * 8 + k]: 0 entry, 1 first copies issued, 2 first chunk
 * landed, 3 main loop done, 4 partial published, 5 exit (combining CTA only); cta = (b * Hkv + kvh) * S + split. */
NVLLM_API int nvllm_paged_decode_set_trace(void* device_buffer);

/* tools/membench.py: streams `bytes` of `buf` with contiguous warp reads (pattern 0) or with the
 * decode kernel's 256-byte-row / token-stride access shape (pattern 1, hkv heads). */
NVLLM_API int nvllm_debug_stream_read(const void* buf, int64_t bytes, int32_t pattern,
                            int32_t hkv, int32_t blocks, void* sink, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NVLLM_TOOLS_H_ */
