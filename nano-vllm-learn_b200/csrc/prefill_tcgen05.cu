// tcgen05/TMEM varlen prefill kernel (sm_100a).  Placeholder until the tensor-core-tile kernel
// lands: reports "not handled" so the caller falls through to the mma.sync tile kernel.
#include "common.cuh"

namespace nvllm {
struct PrefillParams;
int launch_prefill_tcgen05(const PrefillParams&, int, int, int, cudaStream_t, bool* handled) {
  *handled = false;
  return NVLLM_OK;
}
}  // namespace nvllm
