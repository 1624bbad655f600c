// Fused token sampler (SURVEY §8f "next", rank 4).  One launch replaces the eight eager kernels of the
// reference Sampler.forward (nanovllm/layers/sampler.py:10-18): fp32 cast, greedy argmax, divide by the
// temperature, softmax, exponential noise, divide, argmax, where(temperature == 0, greedy, sampled).
//
// HBM bound: every logit is read exactly once (bf16 or fp32), nothing of size [bs, vocab] is written.
//   greedy_b  = argmax_i float(l[b,i])                       ties -> lowest index, NaN counts as greatest
//   sampled_b = argmax_i softmax(l[b,:]/T_b)[i] / (E_i + 1e-10),   E_i ~ Exp(1)
//             = argmax_i ( l[b,i]/T_b - ln(E_i + 1e-10) )     (ln is monotone and the softmax normaliser is a
//                                                              positive per-row constant, so no max/sum pass)
//   token_b   = T_b == 0 ? greedy_b : sampled_b
// E_i comes either from a caller-supplied fp32 tensor (parity tests inject the reference's own noise) or from
// Philox4x32-10 keyed by (seed, offset, row, element group).
// Grid (chunks, bs): a CTA reduces one chunk of one row to a (value, index) pair per criterion; the last CTA
// of a row to arrive (self-resetting counter in the workspace, as in paged_decode.cu) merges the chunk
// partials and writes the int64 token id.
//
// Vocabulary-parallel mode (nvllm_sample_tokens_sharded; SURVEY §8f rank 4, second half): under tensor parallelism
// every rank holds logits[:, r*V/tp : (r+1)*V/tp] (ParallelLMHead, nanovllm/layers/embed_head.py:61-65).  The
// reference gathers all shards on rank 0 over NCCL and concatenates them (embed_head.py:66-70: bs*V*2 bytes into
// one GPU per step) before Sampler.forward runs there.  Here each rank reduces ITS shard to one (key, global index)
// pair per row with the very same walk (Philox counters use the global element index, so the draw does not depend on
// the sharding), and the row's finishing warp swaps pairs with the other ranks through NVLink peer memory as 8-byte
// {payload, call epoch} words -- flag-in-data, no fence, no separate flag -- then takes the best of `world` pairs in
// torch.argmax order.  16*(world-1) bytes per row cross NVLink instead of V*2*(world-1)/world; every rank ends up
// with the token ids (the next step's input_ids never leave the devices), bit-identical to the unsharded kernel on
// the concatenated logits.
#include <limits.h>
#include <string.h>

#include "common.cuh"

namespace nvllm {

namespace {

constexpr int kSampleThreads = 256;
constexpr int kMaxChunks = 64;
constexpr size_t kCounterBytes = 16384;   // int32 arrival counter per row (bs <= 4096)
constexpr float kNoiseEps = 1e-10f;       // sampler.py:16

struct Cand {
  float v;
  int i;
};

// torch's argmax ordering (ATen GreaterOrNan): NaN beats everything, ties go to the lower index.
__device__ __forceinline__ bool better(float av, int ai, float bv, int bi) {
  if (av != av) return (bv != bv) ? ai < bi : true;
  if (bv != bv) return false;
  return (av == bv) ? ai < bi : av > bv;
}
__device__ __forceinline__ void take(Cand& a, float v, int i) {
  if (better(v, i, a.v, a.i)) {
    a.v = v;
    a.i = i;
  }
}
__device__ __forceinline__ void warp_best(Cand& a) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float v = __shfl_xor_sync(0xffffffffu, a.v, o);
    const int i = __shfl_xor_sync(0xffffffffu, a.i, o);
    take(a, v, i);
  }
}

// Philox4x32-10 (Salmon et al., SC'11): counter c[4], key k[2] -> 4 x 32 random bits.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}
// 23 random bits -> m uniform on [0, 1) (exact multiples of 2^-23), E = -ln(1 - m) ~ Exp(1) on [0, 15.95].
// Small m goes through the series m + m^2/2 + m^3/3 so that E keeps RELATIVE accuracy where the second
// logarithm, -ln(E + eps), is steep (lg2.approx alone has ~2^-22 absolute error near 1).
__device__ __forceinline__ float exp1_from_bits(uint32_t w) {
  const float m = __uint_as_float((w >> 9) | 0x3F800000u) - 1.0f;
  const float series = m * fmaf(m, fmaf(m, 0.33333334f, 0.5f), 1.0f);
  const float lg = -0.6931471805599453f * fast_log2(1.0f - m);
  return m < 0.015625f ? series : lg;
}

struct SampleParams {
  const void* logits;
  int64_t row_stride;        // elements
  const float* temperatures;
  const float* noise;        // optional [bs, vocab] fp32 exponential samples
  int64_t noise_row_stride;
  int64_t* tokens;
  int64_t* tokens_host;      // optional second copy (mapped pinned host memory), vocabulary-parallel mode
  int* counters;
  float4* partials;          // [bs][chunks] = (greedy value, greedy index, sample key, sample index)
  int vocab, chunk_elems, chunks;
  uint32_t seed_lo, seed_hi, off_lo, off_hi;
  // vocabulary-parallel mode (world > 1)
  int vocab_offset;          // global index of this shard's column 0 (multiple of 4)
  int world, rank, bs, max_bs;
  char* comm[8];             // every rank's exchange buffer as mapped here; comm[rank] is the local one
  unsigned long long timeout_ns;
  uint32_t* host_error;      // mapped pinned host word (may be null)
};

// ---- exchange buffer of the vocabulary-parallel mode ------------------------------------------------------------
// [0, 4096): control words {epoch = calls completed, rows finished in the running call, error latch};
// then [epoch parity][source rank][row] 16-byte slots {key bits, epoch, index, epoch}, then [epoch parity][row] 8-byte
// words {temperature bits, epoch} written by rank 0.  A rank can be at most one
// call ahead of a peer (it needs every peer's pair of call e to finish call e), so two parities suffice; the buffer
// is zeroed at allocation and epochs start at 1, so a stale word never passes for a current one.
constexpr size_t kXchgHeader = 4096;
constexpr size_t kXchgMinBytes = 299008;   // nvllm_tp_comm_alloc's minimum (it also serves the fused all-reduce)
struct XchgControl {
  uint32_t epoch, rows_done, error;   // same word order as the fused all-reduce's control area (nvllm_tp_comm_status)
};

__device__ __forceinline__ unsigned long long sampler_globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;\n" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_volatile_v2u(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};\n" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ uint2 ld_volatile_v2u(const void* p) {
  uint2 r;
  asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];\n" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const void* p) {
  uint32_t r;
  asm volatile("ld.volatile.global.u32 %0, [%1];\n" : "=r"(r) : "l"(p) : "memory");
  return r;
}

template <typename T>
__device__ __forceinline__ float to_f32(T x);
template <>
__device__ __forceinline__ float to_f32<float>(float x) { return x; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }

// One 16-byte vector (8 bf16 / 4 fp32) of a row, raw; lanes at or beyond `end` read as -inf.
template <typename T, bool kAligned>
__device__ __forceinline__ uint4 load_raw(const T* lrow, int i0, int end) {
  constexpr int kVec = 16 / sizeof(T);
  uint4 raw;
  if (kAligned && i0 + kVec <= end) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w)
                 : "l"(lrow + i0));
  } else {
    T* e = reinterpret_cast<T*>(&raw);
#pragma unroll
    for (int k = 0; k < kVec; ++k) e[k] = (i0 + k < end) ? lrow[i0 + k] : static_cast<T>(-INFINITY);
  }
  return raw;
}
template <typename T>
__device__ __forceinline__ void unpack(const uint4& raw, float (&x)[16 / sizeof(T)]) {
  const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
  for (int k = 0; k < 16 / sizeof(T); ++k) x[k] = to_f32<T>(e[k]);
}

__device__ __forceinline__ float max_nan_f32(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;\n" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ uint32_t max_nan_bf16x2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.NaN.bf16x2 %0, %1, %2;\n" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// Largest element of a vector, NaN if it holds one (torch.argmax ranks NaN above everything).
template <typename T>
__device__ __forceinline__ float vec_max(const uint4& raw);
template <>
__device__ __forceinline__ float vec_max<float>(const uint4& r) {
  return max_nan_f32(max_nan_f32(__uint_as_float(r.x), __uint_as_float(r.y)),
                     max_nan_f32(__uint_as_float(r.z), __uint_as_float(r.w)));
}
template <>
__device__ __forceinline__ float vec_max<__nv_bfloat16>(const uint4& r) {
  const uint32_t m = max_nan_bf16x2(max_nan_bf16x2(r.x, r.y), max_nan_bf16x2(r.z, r.w));
  return max_nan_f32(__uint_as_float(m << 16), __uint_as_float(m & 0xffff0000u));
}

// An element can only win the sampled argmax if logit/T + Gmax beats the best key so far, Gmax = -ln(eps) =
// 23.03 being the largest noise term there is.  In the logit domain: logit > (best - Gmax) * T.  Slack covers
// the roundings; NaN (temperature < 0, or a NaN/inf best) disables skipping because !(x <= NaN) holds.
__device__ __forceinline__ float skip_threshold(float best_key, float temp) {
  if (!(temp > 0.f)) return __int_as_float(0x7fc00000);
  const float k = best_key - 23.1f - fabsf(best_key) * 9.5367431640625e-7f;
  const float t = k * temp;
  return t - fabsf(t) * 9.5367431640625e-7f;
}

// Vocabulary-parallel merge of one row, run by the first warp of the row's finishing CTA.  Lane r talks to rank r:
// it pushes this rank's (key, global index) into rank r's inbox slot [parity][this rank][row] and polls the local
// slot [parity][r][row] until both 8-byte words carry this call's epoch.  Returns the best of all ranks in every lane.
__device__ __forceinline__ Cand exchange_row(const SampleParams& p, int row, Cand mine, int lane) {
  XchgControl* ctl = reinterpret_cast<XchgControl*>(p.comm[p.rank]);
  const uint32_t epoch = ld_volatile_u32(&ctl->epoch) + 1u;
  const size_t parity_off = kXchgHeader + static_cast<size_t>(epoch & 1u) * p.world * p.max_bs * 16;
  Cand got = Cand{-INFINITY, INT_MAX};
  if (lane == p.rank) {
    got = mine;
  } else if (lane < p.world) {
    char* dst = p.comm[lane] + parity_off + (static_cast<size_t>(p.rank) * p.max_bs + row) * 16;
    st_volatile_v2u(dst, __float_as_uint(mine.v), epoch);
    st_volatile_v2u(dst + 8, static_cast<uint32_t>(mine.i), epoch);
    const char* src = p.comm[p.rank] + parity_off + (static_cast<size_t>(lane) * p.max_bs + row) * 16;
    const unsigned long long t0 = sampler_globaltimer_ns();
    uint2 a, b;
    for (uint32_t spins = 1;; ++spins) {
      a = ld_volatile_v2u(src);
      b = ld_volatile_v2u(src + 8);
      if (a.y == epoch && b.y == epoch) break;
      if ((spins & 1023u) == 0 && sampler_globaltimer_ns() - t0 > p.timeout_ns) {
        // the peer never arrived: latch (device word + mapped host word) and go on with what is here
        atomicCAS(&ctl->error, 0u, 1u + static_cast<uint32_t>(row));
        if (p.host_error) *reinterpret_cast<volatile uint32_t*>(p.host_error) = 1u + static_cast<uint32_t>(row);
        a.x = __float_as_uint(-INFINITY);
        b.x = static_cast<uint32_t>(INT_MAX);
        break;
      }
    }
    got = Cand{__uint_as_float(a.x), static_cast<int>(b.x)};
  }
  warp_best(got);
  return got;
}

constexpr int kUnroll = 4;   // 16-byte loads in flight per thread

// kVec elements per thread per load: 8 bf16 or 4 fp32.
template <typename T, bool kAligned>
__global__ void __launch_bounds__(kSampleThreads, 4) sample_tokens_kernel(const SampleParams p) {
  constexpr int kVec = 16 / sizeof(T);
  constexpr int kStride = kSampleThreads * kVec;
  const int row = blockIdx.y;
  const int chunk = blockIdx.x;
  const int tid = threadIdx.x;
  const T* lrow = static_cast<const T*>(p.logits) + static_cast<int64_t>(row) * p.row_stride;
  float temp;
  if (p.world > 1) {
    // Only rank 0 is sure to know the temperatures (the engine's worker ranks never see them: Sequence.__getstate__,
    // nanovllm/engine/sequence.py:77-83, does not pickle them).  Rank 0 hands each row's value to the peers as one
    // {temperature bits, call epoch} word; a rank called without temperatures waits for that word.
    XchgControl* ctl = reinterpret_cast<XchgControl*>(p.comm[p.rank]);
    const uint32_t epoch = ld_volatile_u32(&ctl->epoch) + 1u;
    const size_t toff = kXchgHeader + 2 * static_cast<size_t>(p.world) * p.max_bs * 16 +
                        (static_cast<size_t>(epoch & 1u) * p.max_bs + row) * 8;
    if (p.temperatures) {
      temp = __ldg(p.temperatures + row);
      if (p.rank == 0 && chunk == 0 && tid > 0 && tid < p.world) st_volatile_v2u(p.comm[tid] + toff, __float_as_uint(temp), epoch);
    } else {
      const char* src = p.comm[p.rank] + toff;
      const unsigned long long t0 = sampler_globaltimer_ns();
      uint2 w;
      for (uint32_t spins = 1;; ++spins) {
        w = ld_volatile_v2u(src);
        if (w.y == epoch) break;
        if ((spins & 1023u) == 0 && sampler_globaltimer_ns() - t0 > p.timeout_ns) {
          atomicCAS(&ctl->error, 0u, 1u + static_cast<uint32_t>(row));
          if (p.host_error) *reinterpret_cast<volatile uint32_t*>(p.host_error) = 1u + static_cast<uint32_t>(row);
          w.x = 0u;   // fall back to greedy
          break;
        }
      }
      temp = __uint_as_float(w.x);
    }
  } else {
    temp = __ldg(p.temperatures + row);
  }
  const bool sample = temp != 0.f;
  const float inv_t = __frcp_rn(temp);
  const float* nrow = p.noise ? p.noise + static_cast<int64_t>(row) * p.noise_row_stride : nullptr;
  const int begin = chunk * p.chunk_elems;
  const int end = min(p.vocab, begin + p.chunk_elems);
  const uint2 key = make_uint2(p.seed_lo, p.seed_hi);

  // a row needs only one of the two criteria: greedy (T == 0) or sampled
  Cand g = {-INFINITY, INT_MAX}, s = {-INFINITY, INT_MAX};
  if (!sample) {
    // Greedy: ~1 instruction per element.  Track the best VECTOR (its maximum and where it starts); vectors are
    // visited in rising index order, so a later vector replaces the best one only when strictly greater (or the
    // first to hold a NaN).  The winning vector is re-read once at the end to find the element.
    float best = -INFINITY;
    int best_i0 = -1;
    for (int base = begin + tid * kVec; base < end; base += kUnroll * kStride) {
      uint4 raw[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) raw[u] = load_raw<T, kAligned>(lrow, base + u * kStride, end);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const float m = vec_max<T>(raw[u]);
        const bool upd = (m > best) || ((m != m) && (best == best));
        best = upd ? m : best;
        best_i0 = upd ? base + u * kStride : best_i0;
      }
    }
    if (best_i0 >= 0) {
      float x[kVec];
      unpack<T>(load_raw<T, kAligned>(lrow, best_i0, end), x);
#pragma unroll
      for (int k = kVec - 1; k >= 0; --k)
        if (best_i0 + k < end && (x[k] == best || (best != best && x[k] != x[k]))) g = Cand{x[k], best_i0 + k};
    }
  } else {
    float thr = skip_threshold(-INFINITY, temp);
    for (int base = begin + tid * kVec; base < end; base += kUnroll * kStride) {
      uint4 raw[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) raw[u] = load_raw<T, kAligned>(lrow, base + u * kStride, end);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int i0 = base + u * kStride;
        if (i0 >= end || vec_max<T>(raw[u]) <= thr) continue;   // nothing in this vector can win
        float x[kVec], e[kVec];
        unpack<T>(raw[u], x);
        if (nrow) {
#pragma unroll
          for (int k = 0; k < kVec; ++k) e[k] = (i0 + k < end) ? __ldg(nrow + i0 + k) : 1.f;
        } else {
          // one Philox block per 4 consecutive elements: counter = (element group, row, offset)
#pragma unroll
          for (int k = 0; k < kVec; k += 4) {
            const uint4 r = philox4x32_10(
                make_uint4(static_cast<uint32_t>((p.vocab_offset + i0 + k) >> 2), static_cast<uint32_t>(row), p.off_lo,
                           p.off_hi), key);
            e[k + 0] = exp1_from_bits(r.x);
            e[k + 1] = exp1_from_bits(r.y);
            e[k + 2] = exp1_from_bits(r.z);
            e[k + 3] = exp1_from_bits(r.w);
          }
        }
#pragma unroll
        for (int k = 0; k < kVec; ++k) {
          // logit / T - ln(E + eps); the quotient as a product with the correctly rounded reciprocal.  Indices
          // rise along the walk, so strictly-greater keeps the lowest index among equal keys.
          const float kv = fmaf(-0.6931471805599453f, fast_log2(e[k] + kNoiseEps), x[k] * inv_t);
          const bool upd = (i0 + k < end) && (kv > s.v);
          s.v = upd ? kv : s.v;
          s.i = upd ? i0 + k : s.i;
        }
        thr = skip_threshold(s.v, temp);
      }
    }
  }

  // ---- CTA reduction
  __shared__ Cand sg[kSampleThreads / 32], ss[kSampleThreads / 32];
  __shared__ bool is_last;
  warp_best(g);
  warp_best(s);
  if ((tid & 31) == 0) {
    sg[tid >> 5] = g;
    ss[tid >> 5] = s;
  }
  __syncthreads();
  if (tid < 32) {
    g = tid < kSampleThreads / 32 ? sg[tid] : Cand{-INFINITY, INT_MAX};
    s = tid < kSampleThreads / 32 ? ss[tid] : Cand{-INFINITY, INT_MAX};
    warp_best(g);
    warp_best(s);
    if (tid == 0) {
      p.partials[static_cast<int64_t>(row) * p.chunks + chunk] =
          make_float4(g.v, __int_as_float(g.i), s.v, __int_as_float(s.i));
      __threadfence();
      const int arrived = atomicAdd(p.counters + row, 1);
      is_last = arrived == p.chunks - 1;
      if (is_last) p.counters[row] = 0;   // self-resetting: the next launch starts from zero again
    }
  }
  __syncthreads();
  if (!is_last || tid >= 32) return;
  __threadfence();
  g = Cand{-INFINITY, INT_MAX};
  s = Cand{-INFINITY, INT_MAX};
  for (int c = tid; c < p.chunks; c += 32) {
    const float4 q = __ldcg(p.partials + static_cast<int64_t>(row) * p.chunks + c);
    take(g, q.x, __float_as_int(q.y));
    take(s, q.z, __float_as_int(q.w));
  }
  warp_best(g);
  warp_best(s);
  Cand f = sample ? s : g;   // every lane holds the row's winner (xor butterfly); .i is the index inside this shard
  if (p.world > 1) {
    if (f.i != INT_MAX) f.i += p.vocab_offset;
    f = exchange_row(p, row, f, tid);
  }
  if (tid == 0) {
    int tok = f.i;
    if (tok == INT_MAX) tok = 0;   // a row whose every key is -inf: torch.argmax answers 0
    p.tokens[row] = static_cast<int64_t>(tok);
    if (p.tokens_host) p.tokens_host[row] = static_cast<int64_t>(tok);
    if (p.world > 1) {
      // the call is complete on this rank once every row has swapped: the last row advances the call epoch
      XchgControl* ctl = reinterpret_cast<XchgControl*>(p.comm[p.rank]);
      const uint32_t epoch = ld_volatile_u32(&ctl->epoch) + 1u;
      __threadfence();
      if (atomicAdd(&ctl->rows_done, 1u) == static_cast<uint32_t>(p.bs) - 1u) {
        ctl->rows_done = 0;
        __threadfence();
        *reinterpret_cast<volatile uint32_t*>(&ctl->epoch) = epoch;
      }
    }
  }
}

}  // namespace

}  // namespace nvllm

extern "C" size_t nvllm_sample_workspace_bytes(int32_t bs) {
  if (bs <= 0) return 0;
  return nvllm::kCounterBytes + static_cast<size_t>(bs) * nvllm::kMaxChunks * sizeof(float4);
}

namespace nvllm {
namespace {

unsigned long long g_sample_timeout_ns = 20ull * 1000 * 1000 * 1000;

// mapped pinned host word: the timeout latch of the vocabulary-parallel exchange, readable without a CUDA call
uint32_t* sampler_host_error_word() {
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

int launch_sampler(int64_t* tokens_out, int64_t* tokens_host, const void* logits, int32_t logits_elem_bytes, int64_t logits_row_stride,
                   const float* temperatures, const float* noise, int64_t noise_row_stride, int32_t bs, int32_t vocab,
                   int32_t vocab_offset, uint64_t seed, uint64_t offset, void* workspace, size_t workspace_bytes,
                   void* const* comm_ptrs, size_t comm_bytes, int32_t rank, int32_t world, void* stream) {
  if (bs == 0) return NVLLM_OK;
  if (!tokens_out || !logits || bs < 0 || vocab <= 0 || logits_row_stride < vocab) return NVLLM_ERR_BAD_ARG;
  if (!temperatures && !(world > 1 && rank > 0)) return NVLLM_ERR_BAD_ARG;   // only worker ranks may wait for rank 0's
  if (logits_elem_bytes != 2 && logits_elem_bytes != 4) return NVLLM_ERR_UNSUPPORTED;
  if (static_cast<size_t>(bs) * sizeof(int) > kCounterBytes) return NVLLM_ERR_UNSUPPORTED;
  if (!workspace || workspace_bytes < nvllm_sample_workspace_bytes(bs) || !aligned16(workspace)) return NVLLM_ERR_WORKSPACE;

  SampleParams p = {};
  p.world = 1;
  if (world > 1) {
    if (world > 8 || rank < 0 || rank >= world || !comm_ptrs || vocab_offset < 0) return NVLLM_ERR_BAD_ARG;
    if (vocab_offset % 4 != 0) return NVLLM_ERR_UNSUPPORTED;   // Philox blocks cover 4 consecutive GLOBAL elements
    if (static_cast<int64_t>(vocab_offset) + vocab > INT_MAX - 1) return NVLLM_ERR_UNSUPPORTED;
    // rows per parity this buffer holds: every rank must pass the same comm_bytes (the slot stride derives from it)
    if (comm_bytes < kXchgHeader) return NVLLM_ERR_WORKSPACE;
    const size_t max_bs = (comm_bytes - kXchgHeader) / (2 * static_cast<size_t>(world) * 16 + 16);
    if (max_bs < static_cast<size_t>(bs)) return NVLLM_ERR_WORKSPACE;
    p.max_bs = static_cast<int>(max_bs > 4096 ? 4096 : max_bs);
    for (int r = 0; r < world; ++r) {
      if (!comm_ptrs[r] || !aligned16(comm_ptrs[r])) return NVLLM_ERR_BAD_ARG;
      p.comm[r] = static_cast<char*>(comm_ptrs[r]);
    }
    p.world = world;
    p.rank = rank;
    p.vocab_offset = vocab_offset;
    p.timeout_ns = g_sample_timeout_ns;
    p.host_error = nullptr;
    if (uint32_t* hw = sampler_host_error_word()) {
      void* dp = nullptr;
      if (cudaHostGetDevicePointer(&dp, hw, 0) == cudaSuccess) p.host_error = static_cast<uint32_t*>(dp);
      else cudaGetLastError();
    }
  }
  p.bs = bs;

  const int sms = device_sm_count();
  const int vec = 16 / logits_elem_bytes;
  // one resident wave (__launch_bounds__ guarantees four CTAs per SM) split across the batch, never 1.x waves;
  // a chunk is a whole number of 16-byte vectors
  int chunks = (4 * sms) / bs;
  chunks = chunks < 1 ? 1 : (chunks > kMaxChunks ? kMaxChunks : chunks);
  int chunk_elems = (vocab + chunks - 1) / chunks;
  chunk_elems = (chunk_elems + vec - 1) / vec * vec;
  chunks = (vocab + chunk_elems - 1) / chunk_elems;

  p.logits = logits;
  p.row_stride = logits_row_stride;
  p.temperatures = temperatures;
  p.noise = noise;
  p.noise_row_stride = noise_row_stride;
  p.tokens = tokens_out;
  p.tokens_host = tokens_host;
  p.counters = static_cast<int*>(workspace);
  p.partials = reinterpret_cast<float4*>(static_cast<char*>(workspace) + kCounterBytes);
  p.vocab = vocab;
  p.chunk_elems = chunk_elems;
  p.chunks = chunks;
  p.seed_lo = static_cast<uint32_t>(seed);
  p.seed_hi = static_cast<uint32_t>(seed >> 32);
  p.off_lo = static_cast<uint32_t>(offset);
  p.off_hi = static_cast<uint32_t>(offset >> 32);

  const bool aligned = aligned16(logits) && (logits_row_stride * logits_elem_bytes) % 16 == 0;
  const dim3 grid(static_cast<unsigned>(chunks), static_cast<unsigned>(bs));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (logits_elem_bytes == 2) {
    if (aligned) sample_tokens_kernel<__nv_bfloat16, true><<<grid, kSampleThreads, 0, st>>>(p);
    else sample_tokens_kernel<__nv_bfloat16, false><<<grid, kSampleThreads, 0, st>>>(p);
  } else {
    if (aligned) sample_tokens_kernel<float, true><<<grid, kSampleThreads, 0, st>>>(p);
    else sample_tokens_kernel<float, false><<<grid, kSampleThreads, 0, st>>>(p);
  }
  note_launch();
  return record_cuda_error(cudaGetLastError());
}

}  // namespace
}  // namespace nvllm

extern "C" int nvllm_sample_tokens(int64_t* tokens_out, const void* logits, int32_t logits_elem_bytes,
                                   int64_t logits_row_stride, const float* temperatures, const float* noise,
                                   int64_t noise_row_stride, int32_t bs, int32_t vocab, uint64_t seed,
                                   uint64_t offset, void* workspace, size_t workspace_bytes, void* stream) {
  return nvllm::launch_sampler(tokens_out, nullptr, logits, logits_elem_bytes, logits_row_stride, temperatures, noise,
                               noise_row_stride, bs, vocab, 0, seed, offset, workspace, workspace_bytes, nullptr, 0, 0, 1,
                               stream);
}

extern "C" size_t nvllm_sample_comm_bytes(int32_t world, int32_t max_bs) {
  if (world < 2 || world > 8 || max_bs <= 0 || max_bs > 4096) return 0;
  const size_t need = nvllm::kXchgHeader + (2 * static_cast<size_t>(world) * 16 + 16) * max_bs;
  return need < nvllm::kXchgMinBytes ? nvllm::kXchgMinBytes : need;
}

extern "C" int nvllm_sample_tokens_sharded(int64_t* tokens_out, int64_t* tokens_out_host, const void* logits_shard,
                                           int32_t logits_elem_bytes,
                                           int64_t logits_row_stride, const float* temperatures,
                                           const float* noise_shard, int64_t noise_row_stride, int32_t bs,
                                           int32_t vocab_shard, int32_t vocab_offset, uint64_t seed, uint64_t offset,
                                           void* workspace, size_t workspace_bytes, void* const* comm_ptrs,
                                           size_t comm_bytes, int32_t rank, int32_t world, void* stream) {
  if (world < 2) return NVLLM_ERR_BAD_ARG;
  return nvllm::launch_sampler(tokens_out, tokens_out_host, logits_shard, logits_elem_bytes, logits_row_stride, temperatures, noise_shard,
                               noise_row_stride, bs, vocab_shard, vocab_offset, seed, offset, workspace, workspace_bytes,
                               comm_ptrs, comm_bytes, rank, world, stream);
}

extern "C" int nvllm_sample_sharded_poll_error(uint32_t* error_out, int32_t clear) {
  if (!error_out) return NVLLM_ERR_BAD_ARG;
  uint32_t* w = nvllm::sampler_host_error_word();
  *error_out = w ? *reinterpret_cast<volatile uint32_t*>(w) : 0u;
  if (w && clear) *reinterpret_cast<volatile uint32_t*>(w) = 0u;
  return NVLLM_OK;
}

extern "C" int nvllm_sample_sharded_set_timeout_ms(int64_t timeout_ms) {
  if (timeout_ms <= 0) return NVLLM_ERR_BAD_ARG;
  nvllm::g_sample_timeout_ns = static_cast<unsigned long long>(timeout_ms) * 1000000ull;
  return NVLLM_OK;
}
