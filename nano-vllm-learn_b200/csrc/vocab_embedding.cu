// Fused VocabParallelEmbedding for tensor parallelism (the third collective of a decode step).
//
// Reference: VocabParallelEmbedding.forward, nanovllm/layers/embed_head.py:35-43 -- every rank looks up ALL token ids
// in its own vocabulary shard with out-of-shard ids masked to row 0, zeroes the rows it does not own, and an NCCL
// all-reduce of the [n, hidden] result adds the one real row of each token to world-1 zero rows.
// Here the lookup and the exchange are ONE launch per rank over NVLink peer memory: the rank that OWNS a token's row
// writes it to its own output and pushes it to every peer's inbox as 8-byte {2 x bf16, call epoch} words
// (flag-in-data: no fence, no separate flag, no barrier); the other ranks poll the words of that row and copy the
// payload out.  hidden*2 bytes per token and peer cross NVLink once (x2 for the epoch words) instead of an all-reduce
// of mostly zeros; no arithmetic happens, so the result is the owner's row bit for bit.
// Run-ahead protection: a rank that owns every token of a call waits for nobody, so besides the two epoch parities of
// the inbox each rank publishes "I have entered call e" to its peers, and an owner only overwrites the parity of call
// e-2 once every peer has entered call e-1 (i.e. has finished reading call e-2).  In a lock-stepped engine the check
// is one volatile load that always passes.
#include <string.h>

#include "common.cuh"

namespace nvllm {
namespace {

constexpr int kEmbThreads = 128;
constexpr int kEmbMaxTokens = 512;
constexpr size_t kEmbHeader = 4096;          // control words + entered[world] flags
constexpr size_t kEmbMinBytes = 299008;      // nvllm_tp_comm_alloc's minimum

struct EmbControl {
  uint32_t epoch, done, error, pad;
  uint32_t entered[8];    // entered[r]: the latest call rank r has started (written by rank r over NVLink)
};

struct EmbParams {
  __nv_bfloat16* y;
  int64_t y_stride;
  const int64_t* ids;
  const __nv_bfloat16* weight;
  int64_t w_stride;
  int32_t vocab_start, vocab_per_rank, n, hidden;
  int32_t world, rank, max_tokens;
  char* comm[8];
  unsigned long long timeout_ns;
  uint32_t* host_error;
};

__device__ __forceinline__ unsigned long long emb_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;\n" : "=l"(t));
  return t;
}
__device__ __forceinline__ void emb_st_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint2 emb_ld_v2(const void* p) {
  uint2 r;
  asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];\n" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint32_t emb_ld_u32(const void* p) {
  uint32_t r;
  asm volatile("ld.volatile.global.u32 %0, [%1];\n" : "=r"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void emb_latch(const EmbParams& p, EmbControl* ctl, int row) {
  atomicCAS(&ctl->error, 0u, 1u + static_cast<uint32_t>(row));
  if (p.host_error) *reinterpret_cast<volatile uint32_t*>(p.host_error) = 1u + static_cast<uint32_t>(row);
}

// One CTA per token.  Inbox slot of a token: hidden/4 16-byte units, each {bf16x2, epoch, bf16x2, epoch}.
__global__ void __launch_bounds__(kEmbThreads) vocab_embedding_kernel(const EmbParams p) {
  const int t = blockIdx.x;
  const int tid = threadIdx.x;
  EmbControl* ctl = reinterpret_cast<EmbControl*>(p.comm[p.rank]);
  const uint32_t epoch = emb_ld_u32(&ctl->epoch) + 1u;
  const int units = p.hidden / 4;
  const size_t row_bytes = static_cast<size_t>(p.hidden) * 4;
  const size_t slot = kEmbHeader + (static_cast<size_t>(epoch & 1u) * p.max_tokens + t) * row_bytes;
  // tell the peers this rank has started call `epoch` (and so has finished reading call epoch-1)
  if (t == 0 && tid < p.world && tid != p.rank)
    *reinterpret_cast<volatile uint32_t*>(&reinterpret_cast<EmbControl*>(p.comm[tid])->entered[p.rank]) = epoch;

  const int64_t id = p.ids[t];
  const int64_t local = id - p.vocab_start;
  __nv_bfloat16* yrow = p.y + static_cast<int64_t>(t) * p.y_stride;
  if (id < 0 || id >= static_cast<int64_t>(p.world) * p.vocab_per_rank) {
    // no rank owns this id, so nobody would ever send its row: every rank sees the same id and writes the zeros the
    // reference produces (all masks false -> all_reduce of zero rows, embed_head.py:35-43) without waiting (ADVICE r2:
    // this used to spin for the whole timeout, at every replay of a captured graph)
    for (int u = tid; u < units; u += kEmbThreads) *reinterpret_cast<uint2*>(yrow + 4 * u) = make_uint2(0u, 0u);
  } else if (local >= 0 && local < p.vocab_per_rank) {
    // owner: before overwriting the parity used by call epoch-2, every peer must have entered call epoch-1
    __shared__ int ok;
    if (tid == 0) {
      ok = 1;
      const unsigned long long t0 = emb_timer_ns();
      for (int r = 0; r < p.world; ++r) {
        if (r == p.rank) continue;
        for (uint32_t spins = 1; static_cast<int32_t>(emb_ld_u32(&ctl->entered[r]) - (epoch - 1u)) < 0; ++spins) {
          if ((spins & 1023u) == 0 && emb_timer_ns() - t0 > p.timeout_ns) {
            emb_latch(p, ctl, t);
            break;
          }
        }
      }
    }
    __syncthreads();
    const uint2* wrow = reinterpret_cast<const uint2*>(p.weight + local * p.w_stride);
    for (int u = tid; u < units; u += kEmbThreads) {
      const uint2 v = __ldg(wrow + u);
      *reinterpret_cast<uint2*>(yrow + 4 * u) = v;
      for (int r = 0; r < p.world; ++r)
        if (r != p.rank) emb_st_v4(p.comm[r] + slot + static_cast<size_t>(u) * 16, v.x, epoch, v.y, epoch);
    }
  } else {
    const char* src = p.comm[p.rank] + slot;
    const unsigned long long t0 = emb_timer_ns();
    for (int u = tid; u < units; u += kEmbThreads) {
      uint2 a, b;
      for (uint32_t spins = 1;; ++spins) {
        a = emb_ld_v2(src + static_cast<size_t>(u) * 16);
        b = emb_ld_v2(src + static_cast<size_t>(u) * 16 + 8);
        if (a.y == epoch && b.y == epoch) break;
        if ((spins & 1023u) == 0 && emb_timer_ns() - t0 > p.timeout_ns) {
          emb_latch(p, ctl, t);
          a.x = b.x = 0u;
          break;
        }
      }
      *reinterpret_cast<uint2*>(yrow + 4 * u) = make_uint2(a.x, b.x);
    }
  }
  // the last CTA of the launch advances the call epoch
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(&ctl->done, 1u) == static_cast<uint32_t>(p.n) - 1u) {
      ctl->done = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(&ctl->epoch) = epoch;
    }
  }
}

unsigned long long g_emb_timeout_ns = 20ull * 1000 * 1000 * 1000;

uint32_t* emb_host_error_word() {
  static uint32_t* word = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* q = nullptr;
    if (cudaHostAlloc(&q, 64, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
      memset(q, 0, 64);
      word = static_cast<uint32_t*>(q);
    } else {
      cudaGetLastError();
    }
  }
  return word;
}

}  // namespace
}  // namespace nvllm

extern "C" size_t nvllm_vocab_embedding_comm_bytes(int32_t world, int32_t max_tokens, int32_t hidden) {
  if (world < 2 || world > 8 || max_tokens <= 0 || max_tokens > nvllm::kEmbMaxTokens || hidden <= 0 || hidden % 4) return 0;
  const size_t need = nvllm::kEmbHeader + 2 * static_cast<size_t>(max_tokens) * hidden * 4;
  return need < nvllm::kEmbMinBytes ? nvllm::kEmbMinBytes : need;
}

extern "C" int nvllm_vocab_parallel_embedding(void* y, int64_t y_row_stride, const int64_t* token_ids, const void* weight,
                                              int64_t w_row_stride, int32_t vocab_start, int32_t vocab_per_rank,
                                              int32_t n_tokens, int32_t hidden, int32_t max_tokens,
                                              void* const* comm_ptrs, size_t comm_bytes, int32_t rank, int32_t world,
                                              void* stream) {
  using namespace nvllm;
  if (n_tokens == 0) return NVLLM_OK;
  if (!y || !token_ids || !weight || !comm_ptrs || n_tokens < 0 || hidden <= 0 || vocab_per_rank <= 0 || vocab_start < 0 ||
      world < 2 || world > 8 || rank < 0 || rank >= world || y_row_stride < hidden || w_row_stride < hidden)
    return NVLLM_ERR_BAD_ARG;
  if (hidden % 4 || n_tokens > kEmbMaxTokens || max_tokens < n_tokens || max_tokens > kEmbMaxTokens) return NVLLM_ERR_UNSUPPORTED;
  if ((reinterpret_cast<uintptr_t>(y) & 7) || (reinterpret_cast<uintptr_t>(weight) & 7) || (y_row_stride * 2) % 8 ||
      (w_row_stride * 2) % 8)
    return NVLLM_ERR_ALIGNMENT;
  // every rank must pass the same (max_tokens, hidden, comm_bytes): the slot stride derives from them
  if (comm_bytes < kEmbHeader + 2 * static_cast<size_t>(max_tokens) * hidden * 4) return NVLLM_ERR_WORKSPACE;
  EmbParams p = {};
  p.y = static_cast<__nv_bfloat16*>(y);
  p.y_stride = y_row_stride;
  p.ids = token_ids;
  p.weight = static_cast<const __nv_bfloat16*>(weight);
  p.w_stride = w_row_stride;
  p.vocab_start = vocab_start;
  p.vocab_per_rank = vocab_per_rank;
  p.n = n_tokens;
  p.hidden = hidden;
  p.world = world;
  p.rank = rank;
  p.max_tokens = max_tokens;
  for (int r = 0; r < world; ++r) {
    if (!comm_ptrs[r] || !aligned16(comm_ptrs[r])) return NVLLM_ERR_BAD_ARG;
    p.comm[r] = static_cast<char*>(comm_ptrs[r]);
  }
  p.timeout_ns = g_emb_timeout_ns;
  if (uint32_t* hw = emb_host_error_word()) {
    void* dp = nullptr;
    if (cudaHostGetDevicePointer(&dp, hw, 0) == cudaSuccess) p.host_error = static_cast<uint32_t*>(dp);
    else cudaGetLastError();
  }
  vocab_embedding_kernel<<<n_tokens, kEmbThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  note_launch();
  return record_cuda_error(cudaGetLastError());
}

extern "C" int nvllm_vocab_embedding_poll_error(uint32_t* error_out, int32_t clear) {
  if (!error_out) return NVLLM_ERR_BAD_ARG;
  uint32_t* w = nvllm::emb_host_error_word();
  *error_out = w ? *reinterpret_cast<volatile uint32_t*>(w) : 0u;
  if (w && clear) *reinterpret_cast<volatile uint32_t*>(w) = 0u;
  return NVLLM_OK;
}

extern "C" int nvllm_vocab_embedding_set_timeout_ms(int64_t timeout_ms) {
  if (timeout_ms <= 0) return NVLLM_ERR_BAD_ARG;
  nvllm::g_emb_timeout_ns = static_cast<unsigned long long>(timeout_ms) * 1000000ull;
  return NVLLM_OK;
}
