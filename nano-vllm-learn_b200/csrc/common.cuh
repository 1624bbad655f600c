// Shared device helpers for the sm_100a paged-attention kernels.
// Everything here is inline PTX or plain CUDA; no third-party headers.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/nvllm_attn.h"

namespace nvllm {

constexpr int kHeadDim = 128;             // the only head_dim the kernels are built for
constexpr int kRowBytes = kHeadDim * 2;   // one (token, kv-head) row of bf16
constexpr float kLog2e = 1.4426950408889634f;

// ---- bookkeeping shared by all launchers (defined in api.cu) ---------------
void note_launch(int n = 1);
int record_cuda_error(cudaError_t e);  // returns NVLLM_ERR_CUDA (or OK)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- per-device host state ---------------------------------------------------
// One process may drive several GPUs (TPComm.single_process, the rowparallel tests), and both the SM count and
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) are per device: nothing below is cached per process.
constexpr int kMaxDevices = 64;
int device_sm_count();                 // SM count of the CURRENT device (cached per device; 148 if the query fails)
struct PerDeviceOnce {
  bool done[kMaxDevices] = {};
};
// Runs `configure()` the first time it is called on each device; devices >= kMaxDevices configure every time.
template <class F>
inline cudaError_t once_per_device(PerDeviceOnce& state, F configure) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < kMaxDevices && state.done[dev]) return cudaSuccess;
  e = configure();
  if (e == cudaSuccess && dev >= 0 && dev < kMaxDevices) state.done[dev] = true;
  return e;
}

// ---- small PTX wrappers -----------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 16-byte async copy global->shared, bypassing L1; src_bytes in {0,16}: 0 zero-fills.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(src_bytes)
               : "memory");
}
// same, always 16 bytes (no source-size operand: one register and one predicate less per copy)
__device__ __forceinline__ void cp_async16_full(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3,
                                                  uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_log2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;\n" : "=f"(y) : "f"(x));
  return y;
}

// pack two floats to a bf16x2 register: lo -> bits[0:16), hi -> bits[16:32)
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// Byte offset of 16-byte chunk `chunk` (0..15) of row `row` in a [rows][256 B]
// tile whose chunks are XOR-swizzled by the row's low 3 bits: an 8-row x 16 B
// ldmatrix phase then touches 8 distinct 16-byte bank groups.
__device__ __forceinline__ uint32_t swz_off(int row, int chunk) {
  return static_cast<uint32_t>(row * kRowBytes + ((chunk ^ (row & 7)) << 4));
}

}  // namespace nvllm
