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
#include <limits.h>

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
// 24 random bits -> uniform in (0, 1): never 0 (ln stays finite) and never 1
__device__ __forceinline__ float u01(uint32_t x) { return (static_cast<float>(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

struct SampleParams {
  const void* logits;
  int64_t row_stride;        // elements
  const float* temperatures;
  const float* noise;        // optional [bs, vocab] fp32 exponential samples
  int64_t noise_row_stride;
  int64_t* tokens;
  int* counters;
  float4* partials;          // [bs][chunks] = (greedy value, greedy index, sample key, sample index)
  int vocab, chunk_elems, chunks;
  uint32_t seed_lo, seed_hi, off_lo, off_hi;
};

template <typename T>
__device__ __forceinline__ float to_f32(T x);
template <>
__device__ __forceinline__ float to_f32<float>(float x) { return x; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }

// kVec elements per thread per trip: 8 bf16 or 4 fp32 = one 16-byte load.
template <typename T, bool kAligned>
__global__ void __launch_bounds__(kSampleThreads) sample_tokens_kernel(const SampleParams p) {
  constexpr int kVec = 16 / sizeof(T);
  const int row = blockIdx.y;
  const int chunk = blockIdx.x;
  const int tid = threadIdx.x;
  const T* lrow = static_cast<const T*>(p.logits) + static_cast<int64_t>(row) * p.row_stride;
  const float temp = __ldg(p.temperatures + row);
  const bool sample = temp != 0.f;
  const float* nrow = p.noise ? p.noise + static_cast<int64_t>(row) * p.noise_row_stride : nullptr;
  const int begin = chunk * p.chunk_elems;
  const int end = min(p.vocab, begin + p.chunk_elems);

  Cand g = {-INFINITY, INT_MAX}, s = {-INFINITY, INT_MAX};
  for (int i0 = begin + tid * kVec; i0 < end; i0 += kSampleThreads * kVec) {
    float x[kVec];
    if (kAligned && i0 + kVec <= end) {
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(lrow + i0));
      const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int k = 0; k < kVec; ++k) x[k] = to_f32<T>(e[k]);
    } else {
#pragma unroll
      for (int k = 0; k < kVec; ++k) x[k] = (i0 + k < end) ? to_f32<T>(lrow[i0 + k]) : -INFINITY;
    }
#pragma unroll
    for (int k = 0; k < kVec; ++k)
      if (i0 + k < end) take(g, x[k], i0 + k);
    if (sample) {
      float e[kVec];
      if (nrow) {
#pragma unroll
        for (int k = 0; k < kVec; ++k) e[k] = (i0 + k < end) ? __ldg(nrow + i0 + k) : 1.f;
      } else {
        // one Philox block per 4 consecutive elements: counter = (element group, row, offset)
#pragma unroll
        for (int k = 0; k < kVec; k += 4) {
          const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>((i0 + k) >> 2), static_cast<uint32_t>(row),
                                                   p.off_lo, p.off_hi),
                                        make_uint2(p.seed_lo, p.seed_hi));
          e[k + 0] = -logf(u01(r.x));
          e[k + 1] = -logf(u01(r.y));
          e[k + 2] = -logf(u01(r.z));
          e[k + 3] = -logf(u01(r.w));
        }
      }
#pragma unroll
      for (int k = 0; k < kVec; ++k)
        if (i0 + k < end) take(s, __fdiv_rn(x[k], temp) - logf(e[k] + kNoiseEps), i0 + k);   // IEEE divide, as logits.div_(T)
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
  if (tid == 0) {
    int tok = sample ? s.i : g.i;
    if (tok == INT_MAX) tok = 0;   // a row of -inf sample keys (all-NaN keys cannot lose to the sentinel)
    p.tokens[row] = static_cast<int64_t>(tok);
  }
}

}  // namespace

}  // namespace nvllm

extern "C" size_t nvllm_sample_workspace_bytes(int32_t bs) {
  if (bs <= 0) return 0;
  return nvllm::kCounterBytes + static_cast<size_t>(bs) * nvllm::kMaxChunks * sizeof(float4);
}

extern "C" int nvllm_sample_tokens(int64_t* tokens_out, const void* logits, int32_t logits_elem_bytes,
                                   int64_t logits_row_stride, const float* temperatures, const float* noise,
                                   int64_t noise_row_stride, int32_t bs, int32_t vocab, uint64_t seed,
                                   uint64_t offset, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace nvllm;
  if (bs == 0) return NVLLM_OK;
  if (!tokens_out || !logits || !temperatures || bs < 0 || vocab <= 0 || logits_row_stride < vocab) return NVLLM_ERR_BAD_ARG;
  if (logits_elem_bytes != 2 && logits_elem_bytes != 4) return NVLLM_ERR_UNSUPPORTED;
  if (static_cast<size_t>(bs) * sizeof(int) > kCounterBytes) return NVLLM_ERR_UNSUPPORTED;
  if (!workspace || workspace_bytes < nvllm_sample_workspace_bytes(bs) || !aligned16(workspace)) return NVLLM_ERR_WORKSPACE;

  static int sms = 0;
  if (sms == 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    sms = v;
  }
  const int vec = 16 / logits_elem_bytes;
  // about four CTAs per SM across the batch; a chunk is a whole number of 16-byte vectors per thread trip
  int chunks = (4 * sms + bs - 1) / bs;
  chunks = chunks < 1 ? 1 : (chunks > kMaxChunks ? kMaxChunks : chunks);
  int chunk_elems = (vocab + chunks - 1) / chunks;
  chunk_elems = (chunk_elems + vec - 1) / vec * vec;
  chunks = (vocab + chunk_elems - 1) / chunk_elems;

  SampleParams p;
  p.logits = logits;
  p.row_stride = logits_row_stride;
  p.temperatures = temperatures;
  p.noise = noise;
  p.noise_row_stride = noise_row_stride;
  p.tokens = tokens_out;
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
