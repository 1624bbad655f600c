"""Torch-tensor front end of the C ABI: pointer/stride extraction only, no arithmetic on the host.

Function names and argument meaning follow the reference operators they replace:
  store_kvcache           nanovllm/layers/attention.py:44-55
  flash_attn_with_kvcache nanovllm/layers/attention.py:99-101  (paged decode)
  flash_attn_varlen_func  nanovllm/layers/attention.py:93-96   (varlen prefill, optional block_table)
Every function requires CUDA tensors and the built library; there is no fallback path.
"""
from __future__ import annotations

from typing import Optional

import os

import torch

from . import _lib

_PLAN_CACHE: dict[tuple[int, int, int, int], int] = {}
# Split-KV workspace, one per device, shared by all layers (they run back to back on one stream).  It is
# zero-filled once (the kernel's arrival counters reset themselves) and only ever grows; superseded buffers
# stay referenced because captured CUDA graphs may still point at them.
_WORKSPACE: dict[int, torch.Tensor] = {}
_RETIRED_WORKSPACES: list[torch.Tensor] = []


def _decode_workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    ws = _WORKSPACE.get(idx)
    if ws is None or ws.numel() < nbytes:
        if ws is not None:
            _RETIRED_WORKSPACES.append(ws)
        ws = torch.zeros(max(nbytes, 4 << 20), dtype=torch.uint8, device=device)
        _WORKSPACE[idx] = ws
    return ws


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("nvllm cuda attention backend needs CUDA tensors (no CPU fallback)")


def launch_count() -> int:
    return int(_lib.load().nvllm_launch_count())


def reset_launch_count() -> None:
    _lib.load().nvllm_reset_launch_count()


def set_prefill_variant(variant: int) -> None:
    """0 = mma.sync tile kernel, 1 = tcgen05/TMEM tile kernel, -1 = library default."""
    _lib.check(_lib.load().nvllm_varlen_prefill_set_variant(variant), "set_prefill_variant")


PREFILL_KERNEL_NAMES = {1: "prefill_tcgen05_kernel", 0: "varlen_prefill_mma_kernel", -1: "none"}


def prefill_last_path() -> int:
    """1 = the tcgen05/TMEM kernel served the last flash_attn_varlen_func call, 0 = the mma.sync kernel, -1 = none yet."""
    return int(_lib.load().nvllm_varlen_prefill_last_path())


def prefill_path_count(path: int) -> int:
    return int(_lib.load().nvllm_varlen_prefill_path_count(path))


def prefill_error() -> int:
    """mbarrier-timeout latch of the tcgen05 prefill kernel on the current device (0 = none); clears it.
    Synchronises the device: not for use under graph capture."""
    import ctypes
    code = ctypes.c_uint32(0)
    _lib.check(_lib.load().nvllm_varlen_prefill_error(ctypes.byref(code)), "varlen_prefill_error")
    return int(code.value)


def prefill_poll_error(clear: bool = False) -> int:
    """The same latch through its mapped host word: one ctypes call and one host memory read, no CUDA call (0 = none).
    clear=True with a pending error also clears the device-side latch (that path synchronises)."""
    import ctypes
    code = ctypes.c_uint32(0)
    _lib.check(_lib.load().nvllm_varlen_prefill_poll_error(ctypes.byref(code), 1 if clear else 0),
               "varlen_prefill_poll_error")
    return int(code.value)


def raise_on_prefill_timeout() -> None:
    """Product-path check: flash_attn_varlen_func calls this before every eager launch, so an mbarrier
    timeout of an EARLIER tcgen05 prefill launch becomes an exception at the next call instead of a sticky latch that
    lets every later wait return early.  The latch is cleared before raising: launches after the exception are clean."""
    code = prefill_poll_error()
    if code:
        prefill_poll_error(clear=True)
        raise _lib.NvllmError(f"tcgen05 prefill: an mbarrier wait timed out (code {code:#x}); the outputs of the prefill "
                              f"launches queued since are undefined.  The latch has been cleared")


def set_decode_variant(variant: int) -> None:
    _lib.check(_lib.load().nvllm_paged_decode_set_variant(variant), "set_decode_variant")


def store_kvcache(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                  slot_mapping: torch.Tensor) -> None:
    """key/value [N, Hkv, D] (arbitrary row stride) -> k_cache/v_cache [num_blocks, block, Hkv, D]."""
    lib = _lib.load()
    _require_cuda(key, value, k_cache, v_cache, slot_mapping)
    n, num_kv_heads, head_dim = key.shape
    row = num_kv_heads * head_dim
    # same layout asserts as the reference (attention.py:51-54)
    assert key.stride(-1) == 1 and value.stride(-1) == 1
    assert key.stride(1) == head_dim and value.stride(1) == head_dim
    assert k_cache.stride(1) == row and v_cache.stride(1) == row
    assert slot_mapping.numel() == n
    assert slot_mapping.dtype == torch.int32 and slot_mapping.is_contiguous()
    assert key.dtype == value.dtype == k_cache.dtype == v_cache.dtype
    assert k_cache.stride(0) == k_cache.size(1) * row and v_cache.stride(0) == v_cache.size(1) * row
    if n == 0:
        return
    st = lib.nvllm_store_kvcache(key.data_ptr(), key.stride(0), value.data_ptr(), value.stride(0),
                                 k_cache.data_ptr(), v_cache.data_ptr(), slot_mapping.data_ptr(),
                                 n, num_kv_heads, head_dim, key.element_size(), _stream())
    _lib.check(st, "store_kvcache")


def qk_norm_rope_store(qkv: torch.Tensor, positions: torch.Tensor, q_weight: torch.Tensor, k_weight: torch.Tensor,
                       eps: float, cos_sin_cache: torch.Tensor, num_q_heads: int, num_kv_heads: int, head_dim: int,
                       k_cache: Optional[torch.Tensor] = None, v_cache: Optional[torch.Tensor] = None,
                       slot_mapping: Optional[torch.Tensor] = None, want_k: bool = True):
    """Fused q/k RMSNorm + RoPE + store_kvcache on the fused projection output qkv [N, (Hq+2Hkv)*D].

    Returns (q [N, Hq*D], k [N, Hkv*D] or None).  Reference sequence: qwen3.py:106-117 (q_norm, k_norm,
    rotary_emb) followed by attention.py:86 (store_kvcache)."""
    lib = _lib.load()
    _require_cuda(qkv, positions, q_weight, k_weight, cos_sin_cache, k_cache, v_cache, slot_mapping)
    n = qkv.shape[0]
    assert qkv.dim() == 2 and qkv.stride(-1) == 1 and qkv.shape[1] == (num_q_heads + 2 * num_kv_heads) * head_dim
    assert qkv.dtype == torch.bfloat16 and q_weight.dtype == torch.bfloat16 and k_weight.dtype == torch.bfloat16
    assert q_weight.numel() == head_dim and k_weight.numel() == head_dim
    assert q_weight.is_contiguous() and k_weight.is_contiguous()
    assert positions.dtype == torch.int64 and positions.is_contiguous() and positions.numel() == n
    cs = cos_sin_cache.view(cos_sin_cache.shape[0], -1)
    assert cs.dtype == torch.float32 and cs.is_contiguous() and cs.shape[1] == head_dim
    q = torch.empty((n, num_q_heads * head_dim), dtype=qkv.dtype, device=qkv.device)
    k = torch.empty((n, num_kv_heads * head_dim), dtype=qkv.dtype, device=qkv.device) if want_k else None
    kc_ptr = vc_ptr = sm_ptr = 0
    if k_cache is not None and k_cache.numel():
        row = num_kv_heads * head_dim
        assert k_cache.stride(1) == row and v_cache.stride(1) == row and k_cache.stride(0) == k_cache.size(1) * row
        assert v_cache.stride() == k_cache.stride() and k_cache.dtype == qkv.dtype
        assert slot_mapping is not None and slot_mapping.dtype == torch.int32 and slot_mapping.is_contiguous()
        assert slot_mapping.numel() == n
        kc_ptr, vc_ptr, sm_ptr = k_cache.data_ptr(), v_cache.data_ptr(), slot_mapping.data_ptr()
    if n == 0:
        return q, k
    st = lib.nvllm_qk_norm_rope_store(qkv.data_ptr(), qkv.stride(0), q_weight.data_ptr(), k_weight.data_ptr(), float(eps),
                                      cs.data_ptr(), positions.data_ptr(), q.data_ptr(), q.stride(0),
                                      k.data_ptr() if k is not None else 0, k.stride(0) if k is not None else 0,
                                      kc_ptr, vc_ptr, sm_ptr, n, num_q_heads, num_kv_heads, head_dim, _stream())
    _lib.check(st, "qk_norm_rope_store")
    return q, k


def plan_num_splits(bs: int, num_q_heads: int, num_kv_heads: int, max_ctx: int) -> int:
    """Split count (>= 1) of the split-KV decode grid; 0 (balanced persistent grid) only under NVLLM_DECODE_PLAN=balanced."""
    # the plan depends on NVLLM_DECODE_PLAN, which the C side reads per call: the cached answer must too
    key = (bs, num_q_heads, num_kv_heads, max_ctx, os.environ.get("NVLLM_DECODE_PLAN"))
    s = _PLAN_CACHE.get(key)
    if s is None:
        s = int(_lib.load().nvllm_paged_decode_plan(bs, num_q_heads, num_kv_heads, max_ctx, 0))
        _PLAN_CACHE[key] = s
    return s


def decode_workspace_bytes(bs: int, num_q_heads: int, head_dim: int, num_splits: int) -> int:
    return int(_lib.load().nvllm_paged_decode_workspace_bytes(bs, num_q_heads, head_dim, num_splits))


def flash_attn_with_kvcache(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                            cache_seqlens: torch.Tensor, block_table: torch.Tensor,
                            softmax_scale: Optional[float] = None, causal: bool = True,
                            num_splits: Optional[int] = None, out: Optional[torch.Tensor] = None,
                            workspace: Optional[torch.Tensor] = None,
                            balanced_ctas: Optional[int] = None) -> torch.Tensor:
    """Paged decode.  q [bs, 1, Hq, D] (or [bs, Hq, D]); returns the same shape.

    `causal` is accepted for signature parity; with one query token the bottom-right causal mask is
    a no-op.  num_splits=None or 0 ("choose automatically", as in flash-attn's signature) lets the library plan from
    (bs, Hkv, table width) only; >= 1 is the split-KV grid with that many splits.  The opt-in balanced persistent grid
    (slower on every measured shape) is selected by `balanced_ctas` alone: 0 = fill the SMs, P > 0 = P CTAs.
    """
    lib = _lib.load()
    _require_cuda(q, k_cache, v_cache, cache_seqlens, block_table)
    squeeze = q.dim() == 4
    if squeeze:
        assert q.shape[1] == 1, "Decode stage should have seq_len=1"
        q3 = q[:, 0]
    else:
        q3 = q
    bs, num_q_heads, head_dim = q3.shape
    _, block_size, num_kv_heads, _ = k_cache.shape
    assert q3.dtype == torch.bfloat16 and k_cache.dtype == torch.bfloat16 and v_cache.dtype == torch.bfloat16
    assert q3.stride(-1) == 1 and q3.stride(1) == head_dim
    assert k_cache.stride(-1) == 1 and k_cache.stride(2) == head_dim
    assert v_cache.stride() == k_cache.stride() and v_cache.shape == k_cache.shape
    assert cache_seqlens.dtype == torch.int32 and cache_seqlens.is_contiguous() and cache_seqlens.numel() == bs
    assert block_table.dtype == torch.int32 and block_table.stride(-1) == 1 and block_table.size(0) == bs
    if softmax_scale is None:
        softmax_scale = head_dim ** -0.5
    if balanced_ctas is not None:
        assert balanced_ctas >= 0 and not num_splits, "balanced_ctas excludes an explicit split count"
        num_splits = -int(balanced_ctas)       # the C ABI's encoding (include/nvllm_attn.h): <= 0 is the balanced grid
    elif not num_splits:
        num_splits = plan_num_splits(bs, num_q_heads, num_kv_heads, block_table.size(1) * block_size)
    elif num_splits < 0:
        raise ValueError("num_splits must be None/0 (automatic) or >= 1; the balanced grid is balanced_ctas=")
    if out is None:
        out = torch.empty((bs, num_q_heads, head_dim), dtype=q.dtype, device=q.device)
    ws_bytes = decode_workspace_bytes(bs, num_q_heads, head_dim, num_splits)
    ws_ptr = 0
    if ws_bytes:
        # a caller-provided workspace must have been zero-filled once (see include/nvllm_attn.h)
        if workspace is None or workspace.numel() * workspace.element_size() < ws_bytes:
            workspace = _decode_workspace(q.device, ws_bytes)
        ws_ptr = workspace.data_ptr()
        ws_bytes = workspace.numel() * workspace.element_size()
    st = lib.nvllm_paged_decode(out.data_ptr(), q3.data_ptr(), q3.stride(0), k_cache.data_ptr(), v_cache.data_ptr(),
                                k_cache.stride(0), k_cache.stride(1), block_table.data_ptr(), block_table.stride(0),
                                cache_seqlens.data_ptr(), bs, num_q_heads, num_kv_heads, head_dim, block_size,
                                float(softmax_scale), num_splits, ws_ptr, ws_bytes, _stream())
    _lib.check(st, "paged_decode")
    return out.unsqueeze(1) if squeeze else out


def flash_attn_varlen_func(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, max_seqlen_q: int,
                           cu_seqlens_q: torch.Tensor, max_seqlen_k: int, cu_seqlens_k: torch.Tensor,
                           softmax_scale: Optional[float] = None, causal: bool = True,
                           block_table: Optional[torch.Tensor] = None,
                           out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Varlen causal prefill.  q [sum Lq, Hq, D]; k/v dense [sum Lk, Hkv, D] or, with block_table,
    the paged caches [num_blocks, block, Hkv, D].  Mask is bottom-right aligned (flash-attn)."""
    lib = _lib.load()
    _require_cuda(q, k, v, cu_seqlens_q, cu_seqlens_k, block_table)
    if not causal:
        raise ValueError("the cuda backend implements the causal prefill the reference calls (causal=True)")
    if cu_seqlens_q is None:
        raise ValueError("cu_seqlens_q must be provided for varlen attention")
    total_q, num_q_heads, head_dim = q.shape
    assert q.dtype == torch.bfloat16 and k.dtype == torch.bfloat16 and v.dtype == torch.bfloat16
    assert q.stride(-1) == 1 and q.stride(1) == head_dim
    assert cu_seqlens_q.dtype == torch.int32 and cu_seqlens_k.dtype == torch.int32
    assert cu_seqlens_q.is_contiguous() and cu_seqlens_k.is_contiguous()
    batch = cu_seqlens_q.numel() - 1
    if softmax_scale is None:
        softmax_scale = head_dim ** -0.5
    if out is None:
        out = torch.empty((total_q, num_q_heads, head_dim), dtype=q.dtype, device=q.device)
    if total_q == 0 or batch <= 0:
        return out
    if not torch.cuda.is_current_stream_capturing():
        raise_on_prefill_timeout()
    if block_table is not None:
        assert k.dim() == 4 and v.shape == k.shape and v.stride() == k.stride()
        assert block_table.dtype == torch.int32 and block_table.stride(-1) == 1 and block_table.size(0) == batch
        _, block_size, num_kv_heads, _ = k.shape
        assert k.stride(-1) == 1 and k.stride(2) == head_dim
        k_row, v_row, blk = k.stride(1), v.stride(1), k.stride(0)
        assert blk == block_size * k_row, "paged cache must be one [num_blocks * block_size, Hkv * D] matrix"
        total_k = k.shape[0] * block_size
        bt_ptr, bt_stride = block_table.data_ptr(), block_table.stride(0)
    else:
        assert k.dim() == 3 and v.dim() == 3
        num_kv_heads = k.shape[1]
        assert k.stride(-1) == 1 and k.stride(1) == head_dim and v.stride(-1) == 1 and v.stride(1) == head_dim
        k_row, v_row, blk = k.stride(0), v.stride(0), 0
        total_k = k.shape[0]
        block_size, bt_ptr, bt_stride = 0, 0, 0
    st = lib.nvllm_varlen_prefill(out.data_ptr(), q.data_ptr(), q.stride(0), k.data_ptr(), k_row, v.data_ptr(), v_row,
                                  blk, cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr(), batch, total_q, total_k,
                                  int(max_seqlen_q), int(max_seqlen_k), bt_ptr, bt_stride, num_q_heads, num_kv_heads, head_dim,
                                  block_size, float(softmax_scale), _stream())
    _lib.check(st, "varlen_prefill")
    return out


# Sampler workspace (arrival counters + chunk partials), one per device, zero-filled once.
_SAMPLE_WS: dict = {}


def sample_tokens(logits: torch.Tensor, temperatures: torch.Tensor, noise: Optional[torch.Tensor] = None,
                  seed: int = 0, offset: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused Sampler.forward (nanovllm/layers/sampler.py:10-18): logits [bs, vocab] bf16/fp32, temperatures [bs]
    fp32 -> int64 token ids [bs].  `noise` (fp32 [bs, vocab], Exp(1) draws) replaces the in-kernel Philox stream."""
    lib = _lib.load()
    _require_cuda(logits, temperatures, noise)
    assert logits.dim() == 2 and logits.stride(1) == 1
    assert logits.dtype in (torch.bfloat16, torch.float32)
    bs, vocab = logits.shape
    assert temperatures.dtype == torch.float32 and temperatures.numel() == bs and temperatures.is_contiguous()
    if noise is not None:
        assert noise.dtype == torch.float32 and noise.shape == logits.shape and noise.stride(1) == 1
    if out is None:
        out = torch.empty(bs, dtype=torch.int64, device=logits.device)
    if bs == 0:
        return out
    idx = logits.device.index if logits.device.index is not None else torch.cuda.current_device()
    ws = _sample_workspace(logits.device, bs, idx)
    st = lib.nvllm_sample_tokens(out.data_ptr(), logits.data_ptr(), logits.element_size(), logits.stride(0),
                                 temperatures.data_ptr(), noise.data_ptr() if noise is not None else None,
                                 noise.stride(0) if noise is not None else 0, bs, vocab,
                                 seed & 0xFFFFFFFFFFFFFFFF, offset & 0xFFFFFFFFFFFFFFFF, ws.data_ptr(), ws.numel(),
                                 _stream())
    _lib.check(st, "sample_tokens")
    return out


def _sample_workspace(device: torch.device, bs: int, key) -> torch.Tensor:
    need = int(_lib.load().nvllm_sample_workspace_bytes(bs))
    ws = _SAMPLE_WS.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(max(need, 1 << 20), dtype=torch.uint8, device=device)
        _SAMPLE_WS[key] = ws
    return ws


def sample_tokens_sharded(logits_shard: torch.Tensor, temperatures: Optional[torch.Tensor], vocab_offset: int, comm,
                          noise: Optional[torch.Tensor] = None, seed: int = 0, offset: int = 0,
                          out: Optional[torch.Tensor] = None, host_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Vocabulary-parallel Sampler: replaces the logits gather of ParallelLMHead.forward
    (nanovllm/layers/embed_head.py:66-70) + Sampler.forward on rank 0 (nanovllm/layers/sampler.py:10-18).
    `logits_shard` [bs, vocab/tp] are this rank's columns [vocab_offset, vocab_offset + vocab/tp) of the full logits;
    `comm` is a TPComm sized with sampler_cuda.comm_bytes().  Every rank gets the int64 token ids [bs], equal to
    sample_tokens() on the concatenated logits.  `host_out` (pinned int64 host tensor) receives a second copy of the
    ids, written by the kernel through the buffer's device mapping: the host reads them after a stream/event
    synchronise with no device-to-host copy (what `.tolist()` does at engine/model_runner.py:311).
    `temperatures` may be None on ranks > 0: rank 0's are used (it hands them over inside the kernel)."""
    lib = _lib.load()
    _require_cuda(logits_shard, temperatures, noise)
    assert logits_shard.dim() == 2 and logits_shard.stride(1) == 1
    assert logits_shard.dtype in (torch.bfloat16, torch.float32)
    bs, vocab = logits_shard.shape
    if temperatures is None:
        assert comm.rank > 0, "rank 0 must supply the temperatures"
    else:
        assert temperatures.dtype == torch.float32 and temperatures.numel() == bs and temperatures.is_contiguous()
    if noise is not None:
        assert noise.dtype == torch.float32 and noise.shape == logits_shard.shape and noise.stride(1) == 1
    if out is None:
        out = torch.empty(bs, dtype=torch.int64, device=logits_shard.device)
    else:
        assert out.dtype == torch.int64 and out.numel() >= bs and out.is_contiguous() and (out.is_cuda or out.is_pinned())
    if host_out is not None:
        assert host_out.dtype == torch.int64 and host_out.numel() >= bs and host_out.is_contiguous() and host_out.is_pinned()
    if bs == 0:
        return out
    idx = logits_shard.device.index if logits_shard.device.index is not None else torch.cuda.current_device()
    # one workspace per (device, communicator): ranks sharing a device (loop-back tests) launch concurrently
    ws = _sample_workspace(logits_shard.device, bs, (idx, comm.local_ptr))
    st = lib.nvllm_sample_tokens_sharded(out.data_ptr(), host_out.data_ptr() if host_out is not None else None,
                                         logits_shard.data_ptr(), logits_shard.element_size(),
                                         logits_shard.stride(0),
                                         temperatures.data_ptr() if temperatures is not None else None,
                                         noise.data_ptr() if noise is not None else None,
                                         noise.stride(0) if noise is not None else 0, bs, vocab, vocab_offset,
                                         seed & 0xFFFFFFFFFFFFFFFF, offset & 0xFFFFFFFFFFFFFFFF, ws.data_ptr(),
                                         ws.numel(), comm.ptr_array, comm.nbytes, comm.rank, comm.world, _stream())
    _lib.check(st, "sample_tokens_sharded")
    return out


def rowparallel_linear_allreduce(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], comm,
                                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused RowParallelLinear.forward (nanovllm/layers/linear.py:185-190): y = all_reduce_r(x_r @ W_r^T) (+ bias),
    GEMM and all-reduce in one launch.  x [n, in/tp] bf16 (row-strided), weight [out, in/tp] bf16, bias bf16 [out] or
    None; `comm` is a rowparallel_cuda.TPComm (peer-mapped buffers of all ranks).  Returns a fresh [n, out] bf16."""
    lib = _lib.load()
    _require_cuda(x, weight, bias)
    assert x.dim() == 2 and weight.dim() == 2 and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
    assert x.stride(1) == 1 and weight.stride(1) == 1
    n, k = x.shape
    h, k2 = weight.shape
    assert k == k2, "x and weight shards disagree on in_features_per_rank"
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.numel() == h and bias.is_contiguous()
    if out is None:
        out = torch.empty((n, h), dtype=torch.bfloat16, device=x.device)
    else:
        assert out.shape == (n, h) and out.dtype == torch.bfloat16 and out.stride(1) == 1
    if n == 0:
        return out
    st = lib.nvllm_rowparallel_linear_allreduce(out.data_ptr(), out.stride(0), x.data_ptr(), x.stride(0),
                                                weight.data_ptr(), weight.stride(0),
                                                bias.data_ptr() if bias is not None else None, n, k, h,
                                                comm.ptr_array, comm.nbytes, comm.rank, comm.world, _stream())
    _lib.check(st, "rowparallel_linear_allreduce")
    return out


EMBEDDING_MAX_TOKENS = 512


def vocab_parallel_embedding(token_ids: torch.Tensor, weight: torch.Tensor, vocab_start: int, comm,
                             out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused VocabParallelEmbedding.forward (nanovllm/layers/embed_head.py:35-43): lookup on this rank's vocabulary
    shard `weight` [vocab/tp, hidden] bf16 (rows [vocab_start, vocab_start + vocab/tp) of the table) + the exchange of
    the rows between ranks in one launch; `comm` is a TPComm sized with nvllm_vocab_embedding_comm_bytes(world,
    EMBEDDING_MAX_TOKENS, hidden).  token_ids int64 [n], n <= 512.  Returns [n, hidden] bf16, identical on every rank."""
    lib = _lib.load()
    _require_cuda(token_ids, weight)
    assert token_ids.dim() == 1 and token_ids.dtype == torch.int64 and token_ids.is_contiguous()
    assert weight.dim() == 2 and weight.dtype == torch.bfloat16 and weight.stride(1) == 1
    n = token_ids.numel()
    vpr, hidden = weight.shape
    if out is None:
        out = torch.empty((n, hidden), dtype=torch.bfloat16, device=weight.device)
    else:
        assert out.shape == (n, hidden) and out.dtype == torch.bfloat16 and out.stride(1) == 1
    if n == 0:
        return out
    st = lib.nvllm_vocab_parallel_embedding(out.data_ptr(), out.stride(0), token_ids.data_ptr(), weight.data_ptr(),
                                            weight.stride(0), vocab_start, vpr, n, hidden, EMBEDDING_MAX_TOKENS,
                                            comm.ptr_array, comm.nbytes, comm.rank, comm.world, _stream())
    _lib.check(st, "vocab_parallel_embedding")
    return out


def rowparallel_plan(n_tokens: int, in_features_per_rank: int, out_features: int, world: int) -> tuple[int, int, int]:
    """(feature_tile, ksplit, mode) the library will use for this shape (needs a CUDA device)."""
    import ctypes
    lib = _lib.load()
    bh, ks, mode = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    _lib.check(lib.nvllm_rowparallel_plan(n_tokens, in_features_per_rank, out_features, world,
                                          ctypes.byref(bh), ctypes.byref(ks), ctypes.byref(mode)), "rowparallel_plan")
    return bh.value, ks.value, mode.value


def rowparallel_set_plan(feature_tile: int = 0, ksplit: int = 0, mode: int = -1, timeout_ms: int = 0) -> None:
    _lib.check(_lib.load().nvllm_rowparallel_set_plan(feature_tile, ksplit, mode, timeout_ms), "rowparallel_set_plan")
