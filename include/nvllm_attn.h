/*
 * nvllm_attn.h — C ABI of the B200 (sm_100a) paged-attention hot path.
 *
 * This is the drop-in boundary for nano-vllm's `Attention.forward(q, k, v)` /
 * `Context` contract.  Every entry point takes plain device pointers, sizes,
 * strides (in ELEMENTS unless the name ends in _bytes) and a CUDA stream
 * handle (cudaStream_t passed as void*), and returns an int status
 * (NVLLM_OK == 0).  No torch types, no exceptions, no allocation per call:
 * the caller owns every buffer including the output and the split-KV
 * workspace, so every call is CUDA-graph capturable (this sentence is about
 * the entry points that take a workspace; it says nothing about
 * nvllm_paged_decode_noalloc, section (2b)).
 *
 * Reference interfaces replaced (paths relative to the reference repo):
 *   nvllm_store_kvcache     <- nanovllm/layers/attention.py:19-55
 *                              (store_kvcache_kernel + store_kvcache)
 *   nvllm_paged_decode      <- nanovllm/layers/attention.py:97-101
 *                              (flash_attn_with_kvcache(q, k_cache, v_cache,
 *                               cache_seqlens, block_table, softmax_scale))
 *                              semantics: nanovllm/layers/attention_sdpa.py:122-182
 *   nvllm_varlen_prefill    <- nanovllm/layers/attention.py:89-96
 *                              (flash_attn_varlen_func(..., causal=True,
 *                               block_table=...)); no-prefix semantics:
 *                              nanovllm/layers/attention_sdpa.py:65-119
 *   nvllm_sample_tokens     <- nanovllm/layers/sampler.py:10-18 (Sampler.forward)
 *   nvllm_prepare_decode / nvllm_prepare_prefill_slots
 *                           <- nanovllm/engine/model_runner.py:244-269 and
 *                              :212-221 (index math of the runner glue)
 *
 * All tensors are bf16 (2-byte elements) unless stated; index tensors are
 * int32 exactly as the reference's Context carries them
 * (nanovllm/utils/context.py:5-14).
 */
#ifndef NVLLM_ATTN_H_
#define NVLLM_ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NVLLM_ABI_VERSION 2

#if defined(__GNUC__)
#define NVLLM_API __attribute__((visibility("default")))
#else
#define NVLLM_API
#endif

enum nvllm_status {
  NVLLM_OK = 0,
  NVLLM_ERR_BAD_ARG = 1,      /* null pointer, negative size, bad stride      */
  NVLLM_ERR_UNSUPPORTED = 2,  /* shape the kernels are not built for          */
  NVLLM_ERR_ALIGNMENT = 3,    /* pointer/stride not 16-byte aligned           */
  NVLLM_ERR_CUDA = 4,         /* launch failed; see nvllm_last_cuda_error()   */
  NVLLM_ERR_WORKSPACE = 5     /* workspace missing or too small               */
};

/* ABI version of the loaded library (== NVLLM_ABI_VERSION of its build). */
NVLLM_API int nvllm_abi_version(void);

/* Static string for a status code (never NULL). */
NVLLM_API const char* nvllm_status_string(int status);

/* cudaError_t of the last failed launch seen by this library (0 if none). */
NVLLM_API int nvllm_last_cuda_error(void);

/* Number of kernel launches this library has issued since load (or since the
 * last nvllm_reset_launch_count).  bench.py reports it as `gpu_launches`. */
NVLLM_API int64_t nvllm_launch_count(void);
NVLLM_API void nvllm_reset_launch_count(void);

/* ------------------------------------------------------------------------
 * (1) store_kvcache — slot-mapping scatter into the paged K/V cache.
 *     reference: nanovllm/layers/attention.py:19-55
 *
 *   for i in [0, n_tokens):  s = slot_mapping[i]
 *     k_cache[s * row_elems + j] = key  [i * key_row_stride   + j]
 *     v_cache[s * row_elems + j] = value[i * value_row_stride + j]   j < row_elems
 *   row_elems = num_kv_heads * head_dim.
 *
 *   s == -1 (any negative slot) is skipped: superset of the flash-backend
 *   copy (no guard) and the triton-backend copy (guard,
 *   attention_triton.py:30-31).  Slots are honoured literally otherwise,
 *   including duplicated slot 0 of zero-padded CUDA-graph rows
 *   (model_runner.py:291-293).  Offsets are 64-bit.
 *
 *   elem_bytes is 2 (bf16/fp16); row_elems*elem_bytes, both row strides in
 *   bytes and all four base pointers must be multiples of 16 (16-byte vector
 *   loads/stores), else NVLLM_ERR_ALIGNMENT.
 * ---------------------------------------------------------------------- */
NVLLM_API int nvllm_store_kvcache(const void* key, int64_t key_row_stride,
                        const void* value, int64_t value_row_stride,
                        void* k_cache, void* v_cache,
                        const int32_t* slot_mapping,
                        int32_t n_tokens, int32_t num_kv_heads,
                        int32_t head_dim, int32_t elem_bytes, void* stream);

/* ------------------------------------------------------------------------
 * (2) paged GQA decode (one query token per sequence).
 *     reference: nanovllm/layers/attention.py:97-101
 *
 *   out[b, h, :] = softmax(scale * q[b, h, :] . K_b[0:ctx_b, h / G, :]^T)
 *                  . V_b[0:ctx_b, h / G, :],        G = num_q_heads / num_kv_heads
 *   token t of sequence b lives at
 *   cache[block_tables[b * table_stride + t / block_size], t % block_size, kvh, :].
 *   ctx_b = context_lens[b] is read ON THE DEVICE; launch geometry depends
 *   only on (bs, heads, num_splits), so the call is graph-capturable.
 *   ctx_b == 0 -> the output row is zeros.  Table entries at or beyond
 *   ceil(ctx_b / block_size) are never read (they may be -1 or 0 padding).
 *
 *   q   [bs, num_q_heads, head_dim], row stride q_row_stride, head stride head_dim
 *   out [bs, num_q_heads, head_dim], contiguous
 *   k_cache/v_cache [num_blocks, block_size, num_kv_heads, head_dim]:
 *       cache_block_stride between blocks, cache_token_stride between tokens,
 *       head stride head_dim, unit inner stride.
 *   num_splits >= 1: split-KV grid (num_splits, num_kv_heads, bs), at most 128
 *   splits.  Each split takes an equal share of the LIVE ctx_b (not of the
 *   table width); the last split CTA of each (sequence, kv head) to arrive
 *   merges the partials inside the same launch.  This is what
 *   nvllm_paged_decode_plan returns by default.
 *   num_splits <= 0: BALANCED persistent grid (opt-in; measured slower than
 *   the split-KV grid on B200): -num_splits CTAs, 0 =
 *   fill the SMs (two 4-warp CTAs per SM for G <= 4, else one 8-warp CTA; at
 *   most 512).  The kernel concatenates the 16-token K/V slices of all
 *   (sequence, kv head) items, reads the LIVE context_lens and gives every
 *   CTA an equal contiguous share; a CTA whose share cuts an item writes one
 *   fp32 partial and the CTA completing the item merges them in the same
 *   launch.  Needs bs * num_kv_heads <= 16384 and table_stride * block_size
 *   <= 2^24 (else NVLLM_ERR_UNSUPPORTED).
 *   Every form but num_splits == 1 needs a workspace of
 *   nvllm_paged_decode_workspace_bytes(), which must be ZERO-FILLED ONCE before
 *   its first use (its first 64 KiB hold self-resetting arrival counters, then
 *   the partials) and must not be shared by launches that may run concurrently.
 *   Supported: head_dim == 128, G in {1,2,4,8}, block_size % 32 == 0.
 * ---------------------------------------------------------------------- */
NVLLM_API int nvllm_paged_decode(void* out, const void* q, int64_t q_row_stride,
                       const void* k_cache, const void* v_cache,
                       int64_t cache_block_stride, int64_t cache_token_stride,
                       const int32_t* block_tables, int64_t table_stride,
                       const int32_t* context_lens,
                       int32_t bs, int32_t num_q_heads, int32_t num_kv_heads,
                       int32_t head_dim, int32_t block_size, float scale,
                       int32_t num_splits, void* workspace,
                       size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------
 * (2b) paged GQA decode for callers that have no way to allocate device memory.
 *      No counterpart in the reference (there torch's caching allocator hands
 *      flash-attn its split-KV scratch); this exists for integrations that sit
 *      below torch.
 *
 *   Who needs it: nvllm_paged_decode with num_splits > 1 wants a zero-filled
 *   workspace from the caller.  A host bound to this library over plain FFI (a
 *   C, Go or Rust serving loop, or a plugin slot that only forwards the
 *   tensors the engine already owns) has no torch allocator to ask, and a raw
 *   cudaMalloc is not an option for it either: by the time decode runs the
 *   engine has sized the KV cache to take what was left of HBM
 *   (ModelRunner.allocate_kv_cache, gpu_memory_utilization), so the request
 *   can fail, and cudaMalloc / cudaFree synchronise the whole device, which a
 *   per-token serving loop cannot afford.
 *
 *   What it does: the same computation, the same kernel and the same arguments
 *   as nvllm_paged_decode with the split-KV grid (num_splits in 1..128), minus
 *   the two workspace arguments.  The arrival counters and fp32 partials live
 *   in a fixed array of NVLLM_PAGED_DECODE_NOALLOC_SCRATCH_BYTES (2 MiB) that
 *   is part of the library image: every GPU on which this process has loaded
 *   the library holds those 2 MiB until the process exits, whether or not this
 *   function is ever called.  The function itself calls no cudaMalloc,
 *   cudaFree or any other allocating routine.
 *
 *   Limit: nvllm_paged_decode_workspace_bytes(bs, num_q_heads, head_dim,
 *   num_splits) must not exceed NVLLM_PAGED_DECODE_NOALLOC_SCRATCH_BYTES, i.e.
 *   64 KiB + bs * num_q_heads * num_splits * 516 B <= 2 MiB (3937 partial rows;
 *   a one-wave grid on 148 SMs needs at most 148 * 8 = 1184).  A larger
 *   geometry returns NVLLM_ERR_WORKSPACE before anything is launched; use
 *   fewer splits or nvllm_paged_decode with your own buffer.  num_splits <= 0
 *   (the balanced grid) is NVLLM_ERR_BAD_ARG here.
 *
 *   First call on each device: with that device current, the function looks
 *   up the device address of the array once (cudaGetSymbolAddress; this may
 *   also make the CUDA runtime load the library's device code) and remembers
 *   it per device.  Make that first call during start-up, outside any stream
 *   capture.  Later calls only enqueue work on `stream`.
 *
 *   Sharing: all calls on one device use the one array, and it keeps whatever
 *   the previous call left in it.  Each call therefore begins by clearing the
 *   arrival counters it is about to use (a cudaMemsetAsync on `stream`, ahead
 *   of the kernel); the partials need no clearing, every one read was written
 *   by the same launch.  Two calls that can run on the GPU at the same time
 *   (different streams, or different host threads) would trample each other's
 *   counters and partials: order them on one stream or with events.
 * ---------------------------------------------------------------------- */
#define NVLLM_PAGED_DECODE_NOALLOC_SCRATCH_BYTES (2u << 20)
NVLLM_API int nvllm_paged_decode_noalloc(void* out, const void* q, int64_t q_row_stride,
                       const void* k_cache, const void* v_cache,
                       int64_t cache_block_stride, int64_t cache_token_stride,
                       const int32_t* block_tables, int64_t table_stride,
                       const int32_t* context_lens,
                       int32_t bs, int32_t num_q_heads, int32_t num_kv_heads,
                       int32_t head_dim, int32_t block_size, float scale,
                       int32_t num_splits, void* stream);

/* Bytes of workspace for the given geometry: 0 when num_splits == 1; for
 * num_splits <= 0 (balanced grid) it depends on the CTA count only. */
NVLLM_API size_t nvllm_paged_decode_workspace_bytes(int32_t bs, int32_t num_q_heads,
                                          int32_t head_dim, int32_t num_splits);

/* Host-side split heuristic (pure function, no CUDA calls besides an SM-count
 * query when sm_count <= 0): picks num_splits from the launch-time-known
 * geometry only — max_ctx is the table width * block_size under CUDA graphs.
 * Returns a split count >= 1: 1 once the (sequence, kv head) items cover ~85 % of
 * the SMs, else about one CTA per SM (<= 64 splits, >= 256 tokens per split, a
 * power of two when that keeps >= 3/4 of the CTAs).  NVLLM_DECODE_PLAN=balanced
 * in the environment makes it return 0 (the balanced persistent grid) wherever
 * that kernel applies, for A/B runs. */
NVLLM_API int32_t nvllm_paged_decode_plan(int32_t bs, int32_t num_q_heads, int32_t num_kv_heads,
                                int32_t max_ctx, int32_t sm_count);

/* Tuning hook: selects the decode tile variant (warps/stages/chunk) used by
 * subsequent nvllm_paged_decode calls in this process; -1 restores the
 * default.  Returns NVLLM_ERR_BAD_ARG for an unknown id. */
NVLLM_API int nvllm_paged_decode_set_variant(int32_t variant);
NVLLM_API int32_t nvllm_paged_decode_num_variants(void);
/* Tuning hooks of the balanced grid: 16-token chunks of one kv head a CTA owns at
 * least (fewer CTAs take part when the live work is small; <= 0 = default), and
 * warps per CTA (4 = two CTAs per SM, needs G <= 4; 8 = one CTA per SM; else automatic). */
NVLLM_API int nvllm_paged_decode_set_tuning(int32_t min_chunks_per_cta, int32_t warps_per_cta);
/* Tuning hook of the split-KV grid: when the whole grid is one wave, the split CTAs
 * of a (sequence, kv head) form thread-block clusters and merge through the leader's
 * shared memory (one workspace partial per cluster, none when the cluster holds all
 * splits).  0 = automatic (largest of 8/4/2 that divides num_splits and is
 * co-resident), 1 = never, 2/4/8/16 = that size or none. */
NVLLM_API int nvllm_paged_decode_set_cluster(int32_t cluster);

/* ------------------------------------------------------------------------
 * (3) varlen causal prefill with optional paged prefix (prefix cache).
 *     reference: nanovllm/layers/attention.py:89-96
 *
 *   For sequence i: Lq_i = cu_seqlens_q[i+1]-cu_seqlens_q[i] new tokens,
 *   Lk_i = cu_seqlens_k[i+1]-cu_seqlens_k[i] keys; query j attends keys
 *   0 .. (Lk_i - Lq_i) + j  (causal mask aligned bottom-right, flash-attn
 *   semantics).  Lq_i == 0 produces no rows.
 *
 *   block_tables == NULL: k/v are this step's dense [sum Lk, num_kv_heads,
 *       head_dim] with row strides k_row_stride / v_row_stride.
 *   block_tables != NULL: k/v are the paged caches
 *       [num_blocks, block_size, num_kv_heads, head_dim]; k_row_stride /
 *       v_row_stride are the TOKEN strides and cache_block_stride the block
 *       stride; key t of sequence i is read from
 *       block_tables[i * table_stride + t / block_size].
 *   q   [sum Lq, num_q_heads, head_dim] with row stride q_row_stride
 *   out [sum Lq, num_q_heads, head_dim] contiguous
 *   total_q = rows of q/out; total_k = rows of dense k/v, or
 *   num_blocks * block_size for the paged caches (TMA bounds).
 *   max_seqlen_q bounds the grid (as in the reference call).
 *   Supported: head_dim == 128, num_q_heads % num_kv_heads == 0.
 * ---------------------------------------------------------------------- */
NVLLM_API int nvllm_varlen_prefill(void* out, const void* q, int64_t q_row_stride,
                         const void* k, int64_t k_row_stride,
                         const void* v, int64_t v_row_stride,
                         int64_t cache_block_stride,
                         const int32_t* cu_seqlens_q,
                         const int32_t* cu_seqlens_k,
                         int32_t batch, int32_t total_q, int32_t total_k,
                         int32_t max_seqlen_q, int32_t max_seqlen_k,
                         const int32_t* block_tables, int64_t table_stride,
                         int32_t num_q_heads, int32_t num_kv_heads,
                         int32_t head_dim, int32_t block_size, float scale,
                         void* stream);

/* Tuning hook: 0 = mma.sync tile kernel, 1 = tcgen05/TMEM tile kernel,
 * -1 = default (tcgen05). */
NVLLM_API int nvllm_varlen_prefill_set_variant(int32_t variant);

/* Which kernel served the most recent nvllm_varlen_prefill call of this process:
 * 1 = tcgen05/TMEM tile kernel, 0 = mma.sync tile kernel, -1 = none launched yet.
 * The tcgen05 kernel needs TMA bounds (total_q, total_k > 0), the driver's
 * tensor-map encoder and, for paged K/V, block_size % 128 == 0; everything else
 * runs the mma.sync kernel.  nvllm_varlen_prefill_path_count(path) counts the
 * launches per path since load, so a caller can assert what actually ran. */
NVLLM_API int32_t nvllm_varlen_prefill_last_path(void);
NVLLM_API int64_t nvllm_varlen_prefill_path_count(int32_t path);

/* The tcgen05 kernel bounds every mbarrier wait (2 s): on a timeout it latches
 * an error word and drains instead of trapping.  Returns (and clears) that word
 * for the current device in *code_out; 0 = no error.  Synchronises the device:
 * call it outside graph capture (tests, end of a benchmark, engine shutdown). */
NVLLM_API int nvllm_varlen_prefill_error(uint32_t* code_out);
/* The same word without a CUDA call: the timed-out waiter also stores its code
 * (bit 31 set) into a mapped pinned host word, and this reads it -- cheap enough
 * to run before every prefill launch, which is what ops.flash_attn_varlen_func
 * does so that a timeout surfaces as an exception on the product path instead
 * of silently spoiling every later launch.  clear != 0 and a non-zero word: also
 * clears the device-side latch (that branch synchronises the device).  The host
 * word is per process: with several devices in one process it does not say
 * which device timed out, and clear re-arms the current device only.
 * The tcgen05 tile of nvllm_rowparallel_linear_allreduce has a latch of its own;
 * it reports through nvllm_rowparallel_poll_error. */
NVLLM_API int nvllm_varlen_prefill_poll_error(uint32_t* error_out, int32_t clear);
/* Test hook: enqueues a one-thread kernel that waits 1 ms on an mbarrier nobody
 * completes, in the prefill unit (unit 0) or the RowParallelLinear unit (unit 1),
 * so the tests can watch a timeout travel to the host and be cleared again. */
NVLLM_API int nvllm_tc_latch_selftest(int32_t unit, void* stream);

/* ------------------------------------------------------------------------
 * (3b) fused pre-attention step (SURVEY §8f "next", rank 1): per-head q/k
 *     RMSNorm + NeoX RoPE + store_kvcache in one launch, reading the fused
 *     qkv projection output [n_tokens, (Hq + 2 Hkv) * head_dim] once.
 *     reference: nanovllm/layers/layernorm.py:17-27 (RMSNorm.rms_forward per
 *     head), nanovllm/layers/rotary_embedding.py:6-55, nanovllm/layers/
 *     attention.py:19-55, as sequenced by nanovllm/models/qwen3.py:106-117.
 *
 *   q_out [n, Hq*D] receives the normalised, rotated queries; k_out (may be
 *   NULL) the same for keys (needed by prefill without a prefix hit);
 *   k_cache/v_cache (both NULL = no cache bound, the warmup case) receive K
 *   (normalised, rotated) and V rows at slot_mapping[i] (negative = skip).
 *   cos_sin_cache is the reference's fp32 [max_position, D] table
 *   (cos | sin halves); positions are int64 as the reference feeds them.
 *   Supported: head_dim == 128.
 * ---------------------------------------------------------------------- */
NVLLM_API int nvllm_qk_norm_rope_store(const void* qkv, int64_t qkv_row_stride,
                             const void* q_weight, const void* k_weight,
                             float eps, const float* cos_sin_cache,
                             const int64_t* positions,
                             void* q_out, int64_t q_out_stride,
                             void* k_out, int64_t k_out_stride,
                             void* k_cache, void* v_cache,
                             const int32_t* slot_mapping,
                             int32_t n_tokens, int32_t num_q_heads,
                             int32_t num_kv_heads, int32_t head_dim,
                             void* stream);

/* ------------------------------------------------------------------------
 * (4) runner glue index math on the HOST (plain C, no CUDA): the integer
 *     arithmetic of prepare_decode / prepare_prefill, so the Python glue can
 *     fill persistent pinned staging buffers without per-step list building.
 *     reference: nanovllm/engine/model_runner.py:244-269 and :212-221,
 *                nanovllm/engine/sequence.py:53-65
 * ---------------------------------------------------------------------- */

/* prepare_decode for one sequence: given num_tokens (= len(seq)) and the id
 * of its LAST block, the slot of the token being appended is
 *   last_block * block_size + last_block_num_tokens - 1,
 *   last_block_num_tokens = num_tokens - (ceil(num_tokens/block_size)-1)*block_size.
 * Vectorised over n sequences. */
NVLLM_API int nvllm_prepare_decode_slots(const int32_t* num_tokens,
                               const int32_t* last_block, int32_t n,
                               int32_t block_size, int32_t* slot_mapping_out);

/* prepare_prefill slot ranges for one sequence: appends, for block index i in
 * [num_cached_blocks, num_blocks), range(bt[i]*bs, bt[i]*bs + (bs or
 * last_block_num_tokens)).  Returns the number of slots written, or a
 * negative status on bad arguments / insufficient capacity. */
NVLLM_API int64_t nvllm_prepare_prefill_slots(const int32_t* block_table,
                                    int32_t num_table_entries,
                                    int32_t num_tokens,
                                    int32_t num_cached_tokens,
                                    int32_t block_size, int32_t* slots_out,
                                    int64_t capacity);

/* ------------------------------------------------------------------------
 * (4b) prepare_decode ON THE DEVICE: advance the decode
 *      metadata of an unchanged running batch by one generated token, in
 *      place, in the buffers the captured CUDA graph reads.
 *      reference: nanovllm/engine/model_runner.py:244-269 (prepare_decode:
 *                 Python lists over seqs + 5 pinned tensors + 5 H2D copies per
 *                 token), :160-169 (prepare_block_tables),
 *                 engine/block_manager.py:134-159 (may_append opens a block
 *                 when len % block_size == 1)
 *
 *   for b < bs:  block_tables[patch_rows[i], patch_cols[i]] = patch_vals[i]
 *                input_ids[b]    = sampled_tokens[b]   (sampler output, device)
 *                context_lens[b] += 1;  positions[b] = context_lens[b]  (sic)
 *                slot_mapping[b] = block_tables[b, (len-1)/block_size]
 *                                  * block_size + (len-1) % block_size
 *   patch_* are HOST arrays of n_patch <= nvllm_advance_decode_state_max_patches()
 *   entries (the blocks may_append opened for this step), passed to the kernel
 *   by value: no host-to-device copy.  Everything else is device memory.
 *   Columns [pad_col_begin, pad_col_end) of every row are set to -1 first:
 *   when a new block widens the live table, the other rows carry the
 *   reference's -1 padding there (prepare_block_tables, :163-166).
 * ---------------------------------------------------------------------- */
NVLLM_API int nvllm_advance_decode_state(int64_t* input_ids, int64_t* positions,
                               int32_t* slot_mapping, int32_t* context_lens,
                               int32_t* block_tables, int64_t table_row_stride,
                               const int64_t* sampled_tokens, int32_t bs,
                               int32_t block_size, const int32_t* patch_rows,
                               const int32_t* patch_cols, const int32_t* patch_vals,
                               int32_t n_patch, int32_t pad_col_begin,
                               int32_t pad_col_end, void* stream);
NVLLM_API int32_t nvllm_advance_decode_state_max_patches(void);
/* A row whose new token lands in a table column without a block (negative entry: a patch was not delivered) is
 * parked by the kernel -- slot -1, context length 0, so nothing is stored and no page is read for it -- and 1 + the row
 * index is written to a mapped pinned host word.  This reads that word (plain memory read, no CUDA call; 0 = none);
 * clear != 0 resets it.  RunnerGlue polls it at every decode step and raises. */
NVLLM_API int nvllm_advance_decode_state_poll_error(uint32_t* error_out, int32_t clear);

/* ------------------------------------------------------------------------
 * (5) fused token sampler (SURVEY §8f "next", rank 4): one launch for the
 *     whole of Sampler.forward.
 *     reference: nanovllm/layers/sampler.py:10-18, called from
 *                nanovllm/engine/model_runner.py:311
 *
 *   greedy_b  = argmax_i float(logits[b, i])   (ties -> lowest index, NaN
 *               greatest: torch.argmax ordering)
 *   sampled_b = argmax_i softmax(logits[b,:] / T_b)[i] / (E[b,i] + 1e-10)
 *             = argmax_i (logits[b,i] / T_b - ln(E[b,i] + 1e-10))
 *   tokens_out[b] = T_b == 0 ? greedy_b : sampled_b            (int64)
 *
 *   logits [bs, vocab], logits_elem_bytes 2 (bf16) or 4 (fp32), row stride in
 *   elements; 16-byte aligned rows take the vector path, others a scalar one.
 *   noise != NULL: E is read from this fp32 [bs, vocab] tensor (the parity
 *   tests inject the reference's own exponential_ draw).  noise == NULL:
 *   E[b,i] = -ln(1 - m), m = (w >> 9) / 2^23, w = word i % 4 of the Philox4x32-10
 *   block with key = seed and counter = (i / 4, b, offset).  Elements whose
 *   logit cannot reach the running best key even with the largest possible
 *   noise term (-ln(1e-10)) are skipped without drawing their noise.
 *   Rows holding NaN or +inf logits with T != 0 are outside the contract (the
 *   reference's softmax turns such a row into NaN and answers token 0).
 *   workspace: nvllm_sample_workspace_bytes(bs) bytes, 16-byte aligned,
 *   ZERO-FILLED ONCE before first use (self-resetting arrival counters), not
 *   shared by concurrent launches.  bs <= 4096.
 * ---------------------------------------------------------------------- */
NVLLM_API int nvllm_sample_tokens(int64_t* tokens_out, const void* logits,
                        int32_t logits_elem_bytes, int64_t logits_row_stride,
                        const float* temperatures, const float* noise,
                        int64_t noise_row_stride, int32_t bs, int32_t vocab,
                        uint64_t seed, uint64_t offset, void* workspace,
                        size_t workspace_bytes, void* stream);
NVLLM_API size_t nvllm_sample_workspace_bytes(int32_t bs);

/* ------------------------------------------------------------------------
 * (5b) vocabulary-parallel sampling (second half): the
 *      sampler on each rank's OWN logits shard + a 16-byte-per-row exchange
 *      over NVLink peer memory, instead of gathering [bs, vocab/tp] logits
 *      from every rank to rank 0 over NCCL and concatenating them.
 *      reference: nanovllm/layers/embed_head.py:66-70 (dist.gather +
 *                 torch.cat in ParallelLMHead.forward), then
 *                 nanovllm/layers/sampler.py:10-18 on rank 0 and the
 *                 device->host read at engine/model_runner.py:311.
 *
 *   logits_shard [bs, vocab_shard] holds columns [vocab_offset, vocab_offset
 *   + vocab_shard) of the full logits (vocab_offset % 4 == 0).  Each rank
 *   reduces its shard to one (key, global index) pair per row -- Philox
 *   counters use the GLOBAL element index, so the noise of an element does
 *   not depend on how the vocabulary is split -- pushes the pair to every
 *   peer as two 8-byte {payload, call epoch} words and takes the best of the
 *   `world` pairs in torch.argmax order.  EVERY rank receives tokens_out
 *   [bs], bit-identical to nvllm_sample_tokens on the concatenated logits
 *   with the same (seed, offset) / the concatenated noise.  tokens_out_host
 *   (NULL or the device mapping of pinned host memory; with UVA that is the
 *   host pointer itself) receives a second copy, so the host reads the ids
 *   after a stream synchronise without a device-to-host copy.  noise_shard, if given, is this rank's
 *   [bs, vocab_shard] slice.  temperatures: required on rank 0, which also
 *   hands them to the peers through the exchange buffer; a rank > 0 may pass
 *   NULL and then uses rank 0's (the reference's worker ranks never learn the
 *   temperatures: engine/sequence.py:77-83 does not pickle them,
 *   model_runner.py:309 prepares them on rank 0 only).
 *   Communicator: one buffer of nvllm_sample_comm_bytes(world, max_bs) bytes
 *   per rank from nvllm_tp_comm_alloc (zeroed), peers mapped as in (6);
 *   comm_ptrs / comm_bytes / rank / world as in (6), the same comm_bytes on
 *   every rank.  All ranks issue the same sequence of calls with the same bs;
 *   calls on one communicator are stream-ordered (not concurrent), graph-
 *   capturable, and need no host synchronisation between ranks.  A peer that
 *   never arrives: timeout (default 20 s, nvllm_sample_sharded_set_timeout_ms)
 *   -> error latch readable with nvllm_sample_sharded_poll_error (mapped host
 *   word: 0 = none, else 1 + row) and nvllm_tp_comm_status.
 * ---------------------------------------------------------------------- */
NVLLM_API size_t nvllm_sample_comm_bytes(int32_t world, int32_t max_bs);
NVLLM_API int nvllm_sample_tokens_sharded(int64_t* tokens_out, int64_t* tokens_out_host,
                                const void* logits_shard,
                                int32_t logits_elem_bytes, int64_t logits_row_stride,
                                const float* temperatures, const float* noise_shard,
                                int64_t noise_row_stride, int32_t bs, int32_t vocab_shard,
                                int32_t vocab_offset, uint64_t seed, uint64_t offset,
                                void* workspace, size_t workspace_bytes,
                                void* const* comm_ptrs, size_t comm_bytes,
                                int32_t rank, int32_t world, void* stream);
NVLLM_API int nvllm_sample_sharded_poll_error(uint32_t* error_out, int32_t clear);
NVLLM_API int nvllm_sample_sharded_set_timeout_ms(int64_t timeout_ms);

/* ------------------------------------------------------------------------
 * (6) fused RowParallelLinear (SURVEY §8f "next", rank 2): the o_proj /
 *     down_proj GEMM and the tensor-parallel all-reduce that follows it, in
 *     ONE launch per rank over NVLink peer memory.
 *     reference: nanovllm/layers/linear.py:158-190 (RowParallelLinear:
 *                y = F.linear(x, W_r, bias if rank 0); dist.all_reduce(y)),
 *                call sites nanovllm/models/qwen3.py:118 (o_proj), :147
 *                (down_proj); process group: engine/model_runner.py:31.
 *
 *   y[n, out] = bf16( sum_r sum_k x_r[n, k] * W_r[out, k]  (+ bias[out]) )
 *   with fp32 accumulation and a single rounding; every rank receives the
 *   bit-identical y.  x_r [n_tokens, in_features_per_rank] and W_r
 *   [out_features, in_features_per_rank] are this rank's bf16 shards (row
 *   strides in elements, rows 16-byte aligned), bias bf16 [out_features] or
 *   NULL (pass the same values on every rank: it is added once, after the
 *   reduction).  1 <= n_tokens <= 512 (decode and short prefill; larger
 *   batches are GEMM-bound and keep the reference's cuBLAS + NCCL path),
 *   in_features_per_rank % 64 == 0, out_features % 64 == 0, 2 <= world <= 8.
 *
 *   Communicator: every rank owns ONE buffer of nvllm_rowparallel_comm_bytes()
 *   bytes, obtained from nvllm_tp_comm_alloc (cudaMalloc + zeroed control
 *   area + a 64-byte CUDA IPC handle to hand to the peer processes, e.g. via
 *   torch.distributed.all_gather_object) and mapped by the peers with
 *   nvllm_tp_comm_open; within one process driving several GPUs,
 *   nvllm_tp_comm_enable_peer + the raw pointers are enough.  comm_ptrs is a
 *   HOST array of `world` device pointers — rank r's buffer as mapped in the
 *   calling process, comm_ptrs[rank] being the local one.  All ranks must
 *   issue the same sequence of calls with the same shapes (as tensor
 *   parallelism does); calls are stream-ordered, graph-capturable, allocate
 *   nothing and need no host synchronisation between ranks: a call epoch in
 *   the buffer's control area advances on the device.  A peer that never
 *   arrives makes the kernel give up after the configured timeout (default
 *   20 s; nvllm_rowparallel_set_plan) and latch an error word: on the device
 *   (nvllm_tp_comm_status, synchronous) and in mapped host memory
 *   (nvllm_rowparallel_poll_error, free to poll).
 * ---------------------------------------------------------------------- */
NVLLM_API size_t nvllm_rowparallel_comm_bytes(int32_t world, int32_t max_tokens,
                                    int32_t in_features_per_rank,
                                    int32_t out_features);
/* The parts nvllm_rowparallel_comm_bytes adds up: header (control words + flags), and per epoch parity the
 * "regular" area (fp32 inboxes + two-shot result area, growing from the start of the parity) and the
 * flag-in-data inbox (anchored at its end).  comm_bytes = header + 2 * (regular + flag_in_data).  A buffer shared
 * by several shapes must hold header + 2 * (max regular + max flag_in_data) over those shapes, so that no
 * call's fp32/bf16 data can land in words another call polls as {payload, epoch}. */
NVLLM_API int nvllm_rowparallel_comm_layout(int32_t world, int32_t max_tokens,
                                  int32_t in_features_per_rank, int32_t out_features,
                                  size_t* header_bytes_out, size_t* regular_bytes_out,
                                  size_t* flag_in_data_bytes_out);
NVLLM_API int nvllm_tp_comm_alloc(size_t bytes, void** dev_ptr, void* ipc_handle_out /* 64 bytes or NULL */);
NVLLM_API int nvllm_tp_comm_open(const void* ipc_handle, void** peer_ptr);
NVLLM_API int nvllm_tp_comm_close(void* peer_ptr);
NVLLM_API int nvllm_tp_comm_free(void* dev_ptr);
/* zero the control area (epoch, flags) again; every rank must do it between the same two calls */
NVLLM_API int nvllm_tp_comm_reset(void* dev_ptr, void* stream);
NVLLM_API int nvllm_tp_comm_enable_peer(int32_t peer_device);
/* synchronous read-back of the call epoch and the latched error word (0 = none) */
NVLLM_API int nvllm_tp_comm_status(const void* dev_ptr, uint32_t* epoch_out, uint32_t* error_out);
/* Non-blocking: the timeout latch as a mapped pinned HOST word (0 = none, else 1 + the tile whose peer never
 * arrived in some launch of this process); plain memory read, no CUDA call, usable after every launch or graph
 * replay.  clear != 0 resets it.  The results of a call that timed out are undefined.
 * When no peer timeout is pending it reports the mbarrier-timeout latch of the kernel's tcgen05 tile instead (bit 31
 * set; see nvllm_varlen_prefill_poll_error); clearing that one also re-arms the device-side latch (synchronises). */
NVLLM_API int nvllm_rowparallel_poll_error(uint32_t* error_out, int32_t clear);
NVLLM_API int nvllm_rowparallel_linear_allreduce(void* y, int64_t y_row_stride,
                                       const void* x, int64_t x_row_stride,
                                       const void* weight, int64_t w_row_stride,
                                       const void* bias, int32_t n_tokens,
                                       int32_t in_features_per_rank,
                                       int32_t out_features,
                                       void* const* comm_ptrs, size_t comm_bytes,
                                       int32_t rank, int32_t world, void* stream);
/* The launch plan is a function of the shape alone, so every rank agrees:
 * feature tile (32|64 output features per CTA), k split (= thread-block
 * cluster size; the launcher may halve it to fit the device's co-resident
 * cluster capacity), exchange mode
 * (3 = one-shot with flag-in-data words {payload, call epoch}: partials
 *      pushed to every rank, no fence or separate flag; pushes <= 1 MiB;
 *  0 = one-shot with release/acquire flags; pushes <= 4 MiB;
 *  1 = two-shot: partials pushed to the tile's owner, owner broadcasts bf16;
 *  2 = profiling only, never planned: the rank's own partial goes straight
 *      to y, nothing is exchanged). */
NVLLM_API int nvllm_rowparallel_plan(int32_t n_tokens, int32_t in_features_per_rank,
                           int32_t out_features, int32_t world,
                           int32_t* feature_tile_out, int32_t* ksplit_out,
                           int32_t* mode_out);
/* tuning / tests: force plan fields (0, 0, -1 = automatic) and the peer timeout (<= 0 keeps it) */
NVLLM_API int nvllm_rowparallel_set_plan(int32_t feature_tile, int32_t ksplit,
                               int32_t mode, int64_t timeout_ms);
/* Token count from which the GEMM half runs on the tcgen05 tensor cores (swap-AB tile: 128 output features x 64|128
 * tokens, accumulator in TMEM) instead of the mma.sync tile; needs out_features % 128 == 0.  <= 0 = never.  Must be the
 * same on every rank (the plan is part of the exchange layout).  Env NVLLM_RP_TC_MIN_TOKENS sets the initial value. */
NVLLM_API int nvllm_rowparallel_set_tc_min_tokens(int32_t min_tokens);

/* Measurement and debug aids (read-bandwidth probe, prefill trace buffer) are not
 * part of this ABI: include/nvllm_tools.h, built into lib/libnvllm_tools.so. */

/* ------------------------------------------------------------------------
 * (7) fused VocabParallelEmbedding (the third collective
 *     of a tensor-parallel step): embedding lookup + the exchange of the
 *     looked-up rows in ONE launch per rank over NVLink peer memory.
 *     reference: nanovllm/layers/embed_head.py:35-43 (VocabParallelEmbedding.
 *                forward: masked F.embedding on the local vocabulary shard,
 *                rows of foreign tokens zeroed, dist.all_reduce of [n, hidden])
 *
 *   y[t, :] = W[token_ids[t], :]  on every rank, where rank r holds rows
 *   [vocab_start, vocab_start + vocab_per_rank) of W as `weight`
 *   [vocab_per_rank, hidden] bf16.  The owner of a token writes the row to
 *   its y and pushes it to every peer as 8-byte {bf16 x 2, call epoch} words;
 *   the others poll their inbox.  Bit-exact copy of the owner's row (the
 *   reference's sum adds world-1 zero rows to it).  n_tokens <= 512 (decode and
 *   short prefill; longer batches keep the reference's NCCL path), hidden % 4
 *   == 0, rows 8-byte aligned, 2 <= world <= 8.  Ids outside every shard are
 *   outside the contract (the reference would produce a zero row).
 *   Communicator: nvllm_vocab_embedding_comm_bytes(world, max_tokens, hidden)
 *   bytes per rank from nvllm_tp_comm_alloc, mapped as in (6); every rank
 *   passes the same max_tokens / hidden / comm_bytes.  Same calling rules as
 *   (6): identical call sequence on all ranks, stream-ordered, graph-
 *   capturable, no host synchronisation; timeout -> error latch
 *   (nvllm_vocab_embedding_poll_error: 0 = none, else 1 + token row).
 * ---------------------------------------------------------------------- */
NVLLM_API size_t nvllm_vocab_embedding_comm_bytes(int32_t world, int32_t max_tokens, int32_t hidden);
NVLLM_API int nvllm_vocab_parallel_embedding(void* y, int64_t y_row_stride,
                                   const int64_t* token_ids, const void* weight,
                                   int64_t w_row_stride, int32_t vocab_start,
                                   int32_t vocab_per_rank, int32_t n_tokens,
                                   int32_t hidden, int32_t max_tokens,
                                   void* const* comm_ptrs, size_t comm_bytes,
                                   int32_t rank, int32_t world, void* stream);
NVLLM_API int nvllm_vocab_embedding_poll_error(uint32_t* error_out, int32_t clear);
NVLLM_API int nvllm_vocab_embedding_set_timeout_ms(int64_t timeout_ms);

#ifdef __cplusplus
}
#endif
#endif /* NVLLM_ATTN_H_ */
