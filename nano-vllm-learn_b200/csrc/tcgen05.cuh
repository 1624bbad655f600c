// Inline-PTX wrappers for the Blackwell (sm_100a) tensor-core path: mbarrier, TMA tile loads,
// TMEM allocation, tcgen05.mma / commit / ld / st, and the shared-memory + instruction descriptors.
// No CUTLASS/CuTe: everything here is the raw ISA.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace nvllm {
namespace tc {

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded spin: a protocol bug must neither hang the GPU box nor kill the CUDA context.  After the timeout the
// waiter latches an error word and returns as if the barrier had completed; every other wait of the grid sees the
// latch at its next check and returns too, so the launch drains quickly with undefined results.  The latch is one
// per translation unit (prefill, RowParallelLinear tile).  The waiter also stores the code into a mapped, pinned
// HOST word of the same unit (tc_error_arm), so the product path notices at its next launch with one plain memory
// read (tc_error_poll) -- no CUDA call, no synchronisation -- raises, and clears both words (tc_error_fetch): a
// timeout spoils the launches already queued behind it, never the ones after the host's next call.
static __device__ unsigned int g_tc_error_latch = 0;
static __device__ unsigned int* g_tc_error_host = nullptr;     // device view of the unit's host word (null until armed)
constexpr unsigned long long kMbarTimeoutNs = 2000000000ull;   // 2 s; a legitimate wait is microseconds
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity, unsigned long long timeout_ns) {
  uint32_t spins = 0;
  unsigned long long t0 = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0xFFFu) != 0u) continue;
    // slow path, once per 4096 failed polls
    if (*reinterpret_cast<volatile unsigned int*>(&g_tc_error_latch) != 0u) return;
    const unsigned long long now = global_timer_ns();
    if (t0 == 0) t0 = now;
    if (now - t0 > timeout_ns) {
      const unsigned int code = 0x80000000u | (bar & 0xFFFFu) | ((parity & 1u) << 16);
      if (atomicCAS(&g_tc_error_latch, 0u, code) == 0u) {
        unsigned int* host = *reinterpret_cast<unsigned int* volatile*>(&g_tc_error_host);
        if (host) {
          *reinterpret_cast<volatile unsigned int*>(host) = code;
          __threadfence_system();
        }
      }
      return;
    }
  }
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) { mbar_wait_bounded(bar, parity, kMbarTimeoutNs); }

// ---- host side of the latch.  `static`: every translation unit that includes this header owns its own copy of the
// functions, bound to its own g_tc_error_latch / g_tc_error_host.
// The unit's host word: 64 bytes of mapped pinned memory, allocated at first use, never freed.  nullptr if refused.
static inline unsigned int* tc_error_host_word() {
  static unsigned int* word = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    if (cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
      for (int i = 0; i < 16; ++i) static_cast<volatile unsigned int*>(p)[i] = 0u;
      word = static_cast<unsigned int*>(p);
    } else {
      cudaGetLastError();
    }
  }
  return word;
}
// Points the CURRENT device's g_tc_error_host at the host word; once per device (cudaMemcpyToSymbol: call it from
// set-up code or an eager launch, not while a stream is being captured).  A refusal leaves the device-only latch.
static inline cudaError_t tc_error_arm() {
  static bool armed[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && armed[dev]) return cudaSuccess;
  unsigned int* hw = tc_error_host_word();
  if (hw) {
    void* dp = nullptr;
    e = cudaHostGetDevicePointer(&dp, hw, 0);
    if (e != cudaSuccess) return e;
    unsigned int* dpw = static_cast<unsigned int*>(dp);
    e = cudaMemcpyToSymbol(g_tc_error_host, &dpw, sizeof(dpw));
    if (e != cudaSuccess) return e;
  }
  if (dev >= 0 && dev < 64) armed[dev] = true;
  return cudaSuccess;
}
// One host memory read: the code of a timeout seen by any launch of this unit since the last fetch, 0 = none.
static inline unsigned int tc_error_poll() {
  unsigned int* hw = tc_error_host_word();
  return hw ? *reinterpret_cast<volatile unsigned int*>(hw) : 0u;
}
// Returns and clears the latch of THIS translation unit on the current device, and the host word (synchronises).
static inline cudaError_t tc_error_fetch(unsigned int* out) {
  unsigned int v = 0;
  cudaError_t e = cudaMemcpyFromSymbol(&v, g_tc_error_latch, sizeof(v));
  if (e != cudaSuccess) return e;
  unsigned int* hw = tc_error_host_word();
  const unsigned int hv = hw ? *reinterpret_cast<volatile unsigned int*>(hw) : 0u;
  *out = v ? v : hv;
  if (v != 0u) {
    const unsigned int zero = 0u;
    e = cudaMemcpyToSymbol(g_tc_error_latch, &zero, sizeof(zero));
  }
  if (hw && hv != 0u) *reinterpret_cast<volatile unsigned int*>(hw) = 0u;
  return e;
}
// Self-test of the latch plumbing: one thread waits 1 ms on a barrier nobody completes.
static __global__ void tc_latch_selftest_kernel() {
  __shared__ __align__(8) unsigned long long bar;
  const uint32_t b = static_cast<uint32_t>(__cvta_generic_to_shared(&bar));
  mbar_init(b, 1);
  fence_barrier_init();
  mbar_wait_bounded(b, 0, 1000000ull);
}
static inline cudaError_t tc_latch_selftest(cudaStream_t stream) {
  cudaError_t e = tc_error_arm();
  if (e != cudaSuccess) return e;
  tc_latch_selftest_kernel<<<1, 1, 0, stream>>>();
  return cudaGetLastError();
}

// One lane of a fully converged warp.  The TMA / MMA roles run warp-wide (uniform control flow, so descriptors
// live in uniform registers and UTCHMMA / UTMALDG issue without a divergence waterfall) and only the issue
// instructions themselves are predicated on the elected lane.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.b32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];\n" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
// 2-D tile load: coordinates are (inner element index, row index).
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap* map, int32_t c0, int32_t c1,
                                            uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(
          smem_dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

// ---------------------------------------------------------------- TMEM management
__device__ __forceinline__ void tmem_alloc(uint32_t smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_slot), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory");
}
__device__ __forceinline__ void fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory");
}

// ---------------------------------------------------------------- descriptors
// Shared-memory matrix descriptor (64-bit), sm_100 format:
//   [0,14)  start address >> 4        [16,30) leading-dim byte offset >> 4
//   [32,46) stride-dim byte offset >> 4   [46,49) fixed 0b001   [61,64) layout: 2 = SWIZZLE_128B
// K-major SW128 operand (rows of 64 bf16 = 128 B, 8-row groups of 1024 B): LBO unused, SBO = 1024.
// MN-major SW128 operand (memory rows are K, 64 MN elements = 128 B per row): LBO = byte distance
// between 64-element MN groups, SBO = 1024 (distance between 8-row K groups).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

// Instruction descriptor for kind::f16, BF16 x BF16 -> F32.
//   [4,6) D format (1 = F32)  [7,10) A format (1 = BF16)  [10,13) B format (1 = BF16)
//   [15] A major (0 = K)  [16] B major (0 = K, 1 = MN)  [17,23) N >> 3  [24,29) M >> 4
__host__ __device__ constexpr uint32_t make_idesc_bf16(uint32_t m, uint32_t n, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn_major ? 1u : 0u) << 16) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

// ---------------------------------------------------------------- MMA issue / commit
// D[tmem] (+)= A[smem desc] * B[smem desc]
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                       uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                       uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives once every previously issued tcgen05.mma of this thread has completed.
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar)
               : "memory");
}

// ---------------------------------------------------------------- TMEM <-> registers (32 lanes x 32-bit, one row per thread)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// pointer form for slices of a larger register array (indices must be compile-time after unrolling)
__device__ __forceinline__ void tmem_ld32p(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
// warpgroup-wide register re-budgeting (all four warps of the warpgroup execute it)
template <int N>
__device__ __forceinline__ void reg_dealloc() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(N));
}
template <int N>
__device__ __forceinline__ void reg_alloc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(N));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;\n" ::: "memory"); }

// ---------------------------------------------------------------- host: tensor maps
// cuTensorMapEncodeTiled through the runtime's driver entry point (libcuda is not a link dependency).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
    else
      cudaGetLastError();
  }
  return fn;
}
// [rows][cols] bf16 matrix, row stride in elements; box = 128 rows x 64 cols (128 B), SWIZZLE_128B, out-of-bounds
// rows/columns read as zero.
inline bool make_tile_map_bf16(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t row_stride_elems) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace tc
}  // namespace nvllm
