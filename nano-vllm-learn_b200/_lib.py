"""ctypes binding of the C ABI in include/nvllm_attn.h.  Fails loudly when the library is missing."""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint64, c_void_p
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "lib" / "libnvllm_attn.so"

_P32 = ctypes.POINTER(c_int32)

# name -> (restype, argtypes); mirrors include/nvllm_attn.h one to one
SIGNATURES = {
    "nvllm_abi_version": (c_int, []),
    "nvllm_status_string": (c_char_p, [c_int]),
    "nvllm_last_cuda_error": (c_int, []),
    "nvllm_launch_count": (c_int64, []),
    "nvllm_reset_launch_count": (None, []),
    "nvllm_store_kvcache": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p,
                                    c_int32, c_int32, c_int32, c_int32, c_void_p]),
    "nvllm_paged_decode": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64,
                                   c_void_p, c_int64, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32,
                                   c_float, c_int32, c_void_p, c_size_t, c_void_p]),
    "nvllm_paged_decode_noalloc": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64,
                                           c_void_p, c_int64, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32,
                                           c_float, c_int32, c_void_p]),
    "nvllm_paged_decode_workspace_bytes": (c_size_t, [c_int32, c_int32, c_int32, c_int32]),
    "nvllm_paged_decode_plan": (c_int32, [c_int32, c_int32, c_int32, c_int32, c_int32]),
    "nvllm_paged_decode_set_variant": (c_int, [c_int32]),
    "nvllm_paged_decode_num_variants": (c_int32, []),
    "nvllm_paged_decode_set_tuning": (c_int, [c_int32, c_int32]),
    "nvllm_paged_decode_set_cluster": (c_int, [c_int32]),
    "nvllm_varlen_prefill": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64,
                                     c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_int32, c_void_p,
                                     c_int64, c_int32, c_int32, c_int32, c_int32, c_float, c_void_p]),
    "nvllm_varlen_prefill_set_variant": (c_int, [c_int32]),
    "nvllm_varlen_prefill_last_path": (c_int32, []),
    "nvllm_varlen_prefill_path_count": (c_int64, [c_int32]),
    "nvllm_varlen_prefill_error": (c_int, [ctypes.POINTER(ctypes.c_uint32)]),
    "nvllm_varlen_prefill_poll_error": (c_int, [ctypes.POINTER(ctypes.c_uint32), c_int32]),
    "nvllm_tc_latch_selftest": (c_int, [c_int32, c_void_p]),
    "nvllm_qk_norm_rope_store": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p,
                                         c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int32, c_int32,
                                         c_int32, c_int32, c_void_p]),
    "nvllm_sample_tokens": (c_int, [c_void_p, c_void_p, c_int32, c_int64, c_void_p, c_void_p, c_int64, c_int32, c_int32,
                                    c_uint64, c_uint64, c_void_p, c_size_t, c_void_p]),
    "nvllm_sample_workspace_bytes": (c_size_t, [c_int32]),
    "nvllm_sample_comm_bytes": (c_size_t, [c_int32, c_int32]),
    "nvllm_sample_tokens_sharded": (c_int, [c_void_p, c_void_p, c_void_p, c_int32, c_int64, c_void_p, c_void_p, c_int64, c_int32,
                                            c_int32, c_int32, c_uint64, c_uint64, c_void_p, c_size_t,
                                            ctypes.POINTER(c_void_p), c_size_t, c_int32, c_int32, c_void_p]),
    "nvllm_sample_sharded_poll_error": (c_int, [ctypes.POINTER(ctypes.c_uint32), c_int32]),
    "nvllm_sample_sharded_set_timeout_ms": (c_int, [c_int64]),
    "nvllm_rowparallel_comm_bytes": (c_size_t, [c_int32, c_int32, c_int32, c_int32]),
    "nvllm_rowparallel_comm_layout": (c_int, [c_int32, c_int32, c_int32, c_int32, ctypes.POINTER(c_size_t),
                                              ctypes.POINTER(c_size_t), ctypes.POINTER(c_size_t)]),
    "nvllm_rowparallel_poll_error": (c_int, [ctypes.POINTER(ctypes.c_uint32), c_int32]),
    "nvllm_tp_comm_alloc": (c_int, [c_size_t, ctypes.POINTER(c_void_p), c_void_p]),
    "nvllm_tp_comm_open": (c_int, [c_void_p, ctypes.POINTER(c_void_p)]),
    "nvllm_tp_comm_close": (c_int, [c_void_p]),
    "nvllm_tp_comm_free": (c_int, [c_void_p]),
    "nvllm_tp_comm_reset": (c_int, [c_void_p, c_void_p]),
    "nvllm_tp_comm_enable_peer": (c_int, [c_int32]),
    "nvllm_tp_comm_status": (c_int, [c_void_p, ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_uint32)]),
    "nvllm_rowparallel_linear_allreduce": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                                   c_int32, c_int32, c_int32, ctypes.POINTER(c_void_p), c_size_t,
                                                   c_int32, c_int32, c_void_p]),
    "nvllm_rowparallel_plan": (c_int, [c_int32, c_int32, c_int32, c_int32, _P32, _P32, _P32]),
    "nvllm_rowparallel_set_plan": (c_int, [c_int32, c_int32, c_int32, c_int64]),
    "nvllm_rowparallel_set_tc_min_tokens": (c_int, [c_int32]),
    "nvllm_advance_decode_state": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int32,
                                           c_int32, _P32, _P32, _P32, c_int32, c_int32, c_int32, c_void_p]),
    "nvllm_advance_decode_state_max_patches": (c_int32, []),
    "nvllm_advance_decode_state_poll_error": (c_int, [ctypes.POINTER(ctypes.c_uint32), c_int32]),
    "nvllm_vocab_embedding_comm_bytes": (c_size_t, [c_int32, c_int32, c_int32]),
    "nvllm_vocab_parallel_embedding": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_int32,
                                               c_int32, c_int32, ctypes.POINTER(c_void_p), c_size_t, c_int32, c_int32,
                                               c_void_p]),
    "nvllm_vocab_embedding_poll_error": (c_int, [ctypes.POINTER(ctypes.c_uint32), c_int32]),
    "nvllm_vocab_embedding_set_timeout_ms": (c_int, [c_int64]),
    "nvllm_prepare_decode_slots": (c_int, [_P32, _P32, c_int32, c_int32, _P32]),
    "nvllm_prepare_prefill_slots": (c_int64, [_P32, c_int32, c_int32, c_int32, c_int32, _P32, c_int64]),
}

# include/nvllm_tools.h (lib/libnvllm_tools.so only; tools/*.py call load_tools())
TOOLS_SIGNATURES = {
    "nvllm_varlen_prefill_set_debug_buffer": (None, [c_void_p]),
    "nvllm_varlen_prefill_set_exp2_poly": (None, [c_int]),
    "nvllm_paged_decode_set_trace": (c_int, [c_void_p]),
    "nvllm_debug_stream_read": (c_int, [c_void_p, c_int64, c_int32, c_int32, c_int32, c_void_p, c_void_p]),
}

ABI_VERSION = 2
_LIB = None


class NvllmError(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Load (once) and type the library.  No fallback: a missing .so is an error."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = Path(os.environ.get("NVLLM_ATTN_LIB", LIB_PATH))
    if not path.exists():
        raise NvllmError(
            f"{path} not found: build it with `python nano-vllm-learn_b200/build.py` "
            "(or __graft_entry__.build()); the cuda attention backend has no CPU/eager fallback")
    lib = ctypes.CDLL(str(path))
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    got = lib.nvllm_abi_version()
    if got != ABI_VERSION:
        raise NvllmError(f"ABI mismatch: library reports {got}, binding expects {ABI_VERSION}")
    _LIB = lib
    return lib


def load_tools() -> ctypes.CDLL:
    """Measurement/debug build (build.py --tools): becomes THE library of this process, so ops.* run the traced
    kernels.  Must be called before anything else loaded the product library."""
    global _LIB
    path = _PKG / "lib" / "libnvllm_tools.so"
    if not path.exists():
        raise NvllmError(f"{path} not found: build it with `python nano-vllm-learn_b200/build.py --tools`")
    if _LIB is not None and not hasattr(_LIB, "nvllm_debug_stream_read"):
        raise NvllmError("the product library is already loaded; call load_tools() first")
    os.environ["NVLLM_ATTN_LIB"] = str(path)
    lib = load()
    for name, (restype, argtypes) in TOOLS_SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    return lib


def check(status: int, what: str) -> None:
    """Map a C status to the exception types the reference raises (Errors row)."""
    if status == 0:
        return
    lib = load()
    msg = lib.nvllm_status_string(status).decode()
    if status in (1, 3):      # bad argument / alignment -> the reference asserts on layout
        raise AssertionError(f"{what}: {msg}")
    if status == 2:           # unsupported shape -> ValueError like an unknown backend
        raise ValueError(f"{what}: {msg}")
    if status == 4:
        raise NvllmError(f"{what}: {msg} (cudaError {lib.nvllm_last_cuda_error()})")
    raise NvllmError(f"{what}: {msg}")
