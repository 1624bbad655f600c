// Fused pre-attention step: per-head q/k RMSNorm + NeoX RoPE + store_kvcache in ONE
// kernel, reading the fused qkv projection output once.  Replaces, for the cuda backend, the sequence
//   q_norm / k_norm   reference nanovllm/layers/layernorm.py:17-27   (RMSNorm.rms_forward per head)
//   rotary_emb        reference nanovllm/layers/rotary_embedding.py:6-55
//   store_kvcache     reference nanovllm/layers/attention.py:19-55
// called from Qwen3Attention.forward (nanovllm/models/qwen3.py:106-117): four launch-bound kernels per
// layer in decode (N <= 512 rows) become one.
//
// Arithmetic follows the reference op by op: x -> fp32, var = mean(x^2), x *= rsqrt(var + eps),
// round to bf16, multiply by the bf16 weight (rounded to bf16), then RoPE in fp32 on the two halves
// (y1 = x1*cos - x2*sin, y2 = x2*cos + x1*sin, products and sum rounded separately), round to bf16.
//
// HBM-bound, trivially: N * ((Hq + 2 Hkv) * 256 read + (Hq + 3 Hkv) * 256 written [+ Hkv * 256 dense K]) bytes.
// One warp per (token, head); a lane owns 4 consecutive elements (8-byte vectors); the RoPE partner
// (element i +- 64) lives in lane +- 16.
#include "common.cuh"

namespace nvllm {

struct PreAttnParams {
  const __nv_bfloat16* qkv;
  int64_t qkv_row_stride;
  const __nv_bfloat16* q_weight;
  const __nv_bfloat16* k_weight;
  const float* cos_sin;        // [max_pos, 128]: cos[0:64] | sin[0:64]
  const int64_t* positions;    // [N]
  const int32_t* slots;        // [N] or nullptr
  __nv_bfloat16* q_out;        // [N, Hq*128]
  int64_t q_out_stride;
  __nv_bfloat16* k_out;        // [N, Hkv*128] or nullptr
  int64_t k_out_stride;
  __nv_bfloat16* k_cache;      // flat [slots, Hkv*128] or nullptr
  __nv_bfloat16* v_cache;
  int32_t n, hq, hkv;
  float eps;
};

__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__global__ void __launch_bounds__(256) qk_norm_rope_store_kernel(const PreAttnParams p) {
  const int lane = threadIdx.x & 31;
  const int heads = p.hq + 2 * p.hkv;
  const int64_t units = static_cast<int64_t>(p.n) * heads;
  const int64_t warp0 = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t u = warp0; u < units; u += nwarps) {
    const int tok = static_cast<int>(u / heads);
    const int head = static_cast<int>(u - static_cast<int64_t>(tok) * heads);
    const __nv_bfloat16* src = p.qkv + static_cast<int64_t>(tok) * p.qkv_row_stride + head * kHeadDim + lane * 4;
    const uint2 raw = *reinterpret_cast<const uint2*>(src);
    const int slot = p.slots ? __ldg(p.slots + tok) : -1;
    const bool to_cache = p.k_cache != nullptr && slot >= 0;
    if (head >= p.hq + p.hkv) {  // V head: plain copy into the cache
      if (to_cache) {
        const int vh = head - p.hq - p.hkv;
        *reinterpret_cast<uint2*>(p.v_cache + (static_cast<int64_t>(slot) * p.hkv + vh) * kHeadDim + lane * 4) = raw;
      }
      continue;
    }
    const bool is_q = head < p.hq;
    const __nv_bfloat162 lo = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
    const __nv_bfloat162 hi = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
    float x[4] = {__low2float(lo), __high2float(lo), __low2float(hi), __high2float(hi)};
    float ss = x[0] * x[0] + x[1] * x[1] + x[2] * x[2] + x[3] * x[3];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float r = rsqrtf(ss * (1.f / kHeadDim) + p.eps);
    const uint2 wraw = __ldg(reinterpret_cast<const uint2*>((is_q ? p.q_weight : p.k_weight) + lane * 4));
    const __nv_bfloat162 wlo = *reinterpret_cast<const __nv_bfloat162*>(&wraw.x);
    const __nv_bfloat162 whi = *reinterpret_cast<const __nv_bfloat162*>(&wraw.y);
    const float w[4] = {__low2float(wlo), __high2float(wlo), __low2float(whi), __high2float(whi)};
    float y[4], partner[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      y[i] = bf16_round(bf16_round(x[i] * r) * w[i]);  // .to(bf16).mul_(weight)
      partner[i] = __shfl_xor_sync(0xffffffffu, y[i], 16);
    }
    const int64_t pos = __ldg(p.positions + tok);
    const float4 cs = __ldg(reinterpret_cast<const float4*>(p.cos_sin + pos * kHeadDim + (lane & 15) * 4));
    const float4 sn = __ldg(reinterpret_cast<const float4*>(p.cos_sin + pos * kHeadDim + 64 + (lane & 15) * 4));
    const float c[4] = {cs.x, cs.y, cs.z, cs.w};
    const float s[4] = {sn.x, sn.y, sn.z, sn.w};
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      // lanes 0-15 hold x1 (first half), lanes 16-31 hold x2
      o[i] = lane < 16 ? __fsub_rn(__fmul_rn(y[i], c[i]), __fmul_rn(partner[i], s[i]))
                       : __fadd_rn(__fmul_rn(y[i], c[i]), __fmul_rn(partner[i], s[i]));
    }
    uint2 outv;
    outv.x = pack_bf16x2(o[0], o[1]);
    outv.y = pack_bf16x2(o[2], o[3]);
    if (is_q) {
      *reinterpret_cast<uint2*>(p.q_out + static_cast<int64_t>(tok) * p.q_out_stride + head * kHeadDim + lane * 4) = outv;
    } else {
      const int kh = head - p.hq;
      if (p.k_out)
        *reinterpret_cast<uint2*>(p.k_out + static_cast<int64_t>(tok) * p.k_out_stride + kh * kHeadDim + lane * 4) = outv;
      if (to_cache)
        *reinterpret_cast<uint2*>(p.k_cache + (static_cast<int64_t>(slot) * p.hkv + kh) * kHeadDim + lane * 4) = outv;
    }
  }
}

}  // namespace nvllm

extern "C" int nvllm_qk_norm_rope_store(const void* qkv, int64_t qkv_row_stride, const void* q_weight,
                                        const void* k_weight, float eps, const float* cos_sin_cache,
                                        const int64_t* positions, void* q_out, int64_t q_out_stride, void* k_out,
                                        int64_t k_out_stride, void* k_cache, void* v_cache,
                                        const int32_t* slot_mapping, int32_t n_tokens, int32_t num_q_heads,
                                        int32_t num_kv_heads, int32_t head_dim, void* stream) {
  using namespace nvllm;
  if (n_tokens < 0 || num_q_heads <= 0 || num_kv_heads <= 0) return NVLLM_ERR_BAD_ARG;
  if (n_tokens == 0) return NVLLM_OK;
  if (!qkv || !q_weight || !k_weight || !cos_sin_cache || !positions || !q_out) return NVLLM_ERR_BAD_ARG;
  if ((k_cache == nullptr) != (v_cache == nullptr)) return NVLLM_ERR_BAD_ARG;
  if (k_cache && !slot_mapping) return NVLLM_ERR_BAD_ARG;
  if (head_dim != kHeadDim) return NVLLM_ERR_UNSUPPORTED;
  const int64_t row = static_cast<int64_t>(num_q_heads + 2 * num_kv_heads) * head_dim;
  if (qkv_row_stride < row || q_out_stride < static_cast<int64_t>(num_q_heads) * head_dim ||
      (k_out && k_out_stride < static_cast<int64_t>(num_kv_heads) * head_dim))
    return NVLLM_ERR_BAD_ARG;
  if (!aligned16(qkv) || !aligned16(q_out) || !aligned16(q_weight) || !aligned16(k_weight) ||
      !aligned16(cos_sin_cache) || (k_out && !aligned16(k_out)) || (k_cache && (!aligned16(k_cache) || !aligned16(v_cache))) ||
      (qkv_row_stride & 3) || (q_out_stride & 3) || (k_out_stride & 3))
    return NVLLM_ERR_ALIGNMENT;
  PreAttnParams p;
  p.qkv = static_cast<const __nv_bfloat16*>(qkv);
  p.qkv_row_stride = qkv_row_stride;
  p.q_weight = static_cast<const __nv_bfloat16*>(q_weight);
  p.k_weight = static_cast<const __nv_bfloat16*>(k_weight);
  p.cos_sin = cos_sin_cache;
  p.positions = positions;
  p.slots = slot_mapping;
  p.q_out = static_cast<__nv_bfloat16*>(q_out);
  p.q_out_stride = q_out_stride;
  p.k_out = static_cast<__nv_bfloat16*>(k_out);
  p.k_out_stride = k_out_stride;
  p.k_cache = static_cast<__nv_bfloat16*>(k_cache);
  p.v_cache = static_cast<__nv_bfloat16*>(v_cache);
  p.n = n_tokens;
  p.hq = num_q_heads;
  p.hkv = num_kv_heads;
  p.eps = eps;
  const int64_t units = static_cast<int64_t>(n_tokens) * (num_q_heads + 2 * num_kv_heads);
  int64_t blocks = (units + 7) / 8;           // 8 warps per 256-thread block
  if (blocks > 148 * 8) blocks = 148 * 8;     // grid-stride beyond 8 resident blocks per SM
  qk_norm_rope_store_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  note_launch();
  return record_cuda_error(cudaGetLastError());
}
