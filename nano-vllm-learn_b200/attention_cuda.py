"""`--attn-backend cuda`: drop-in for the reference attention operator.

Same constructor, mutable attributes and `forward(q, k, v)` contract as
`nanovllm/layers/attention.py:58-103` (class Attention): per-rank head counts,
`k_cache`/`v_cache` start as empty tensors and are later bound by
`ModelRunner.allocate_kv_cache` (model_runner.py:148-157), all other inputs come from the
module-global Context.  The body only orders the three C-ABI calls; store precedes attention on
the same stream, which same-batch prefix-cache hits rely on.
"""
from __future__ import annotations

import torch
from torch import nn

from . import context as _own_context
from . import ops

_get_context = _own_context.get_context


def bind_context(get_context_fn) -> None:
    """Point the backend at the host engine's context getter (nanovllm.utils.context.get_context)."""
    global _get_context
    _get_context = get_context_fn


class Attention(nn.Module):

    def __init__(self, num_heads: int, head_dim: int, scale: float, num_kv_heads: int, block_size: int = 256):
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.scale = scale
        self.num_kv_heads = num_kv_heads
        self.block_size = block_size
        self.k_cache = self.v_cache = torch.tensor([])

    def forward_fused(self, qkv: torch.Tensor, positions: torch.Tensor, q_norm, k_norm, rotary_emb) -> torch.Tensor:
        """q/k RMSNorm + RoPE + store_kvcache in one kernel, then attention.

        Takes the fused projection output (qwen3.py:104) and the reference's own norm / rotary modules
        (their weights, eps and cos_sin_cache are read, their forward is not called); replaces
        qwen3.py:106-117.  Store still precedes attention on the same stream."""
        context = _get_context()
        k_cache, v_cache = self.k_cache, self.v_cache
        have_cache = bool(k_cache.numel() and v_cache.numel() and context.slot_mapping is not None)
        dense_kv = context.is_prefill and context.block_tables is None
        q, k = ops.qk_norm_rope_store(qkv, positions, q_norm.weight, k_norm.weight, q_norm.eps,
                                      rotary_emb.cos_sin_cache, self.num_heads, self.num_kv_heads, self.head_dim,
                                      k_cache if have_cache else None, v_cache if have_cache else None,
                                      context.slot_mapping if have_cache else None, want_k=dense_kv)
        q = q.view(-1, self.num_heads, self.head_dim)
        if context.is_prefill:
            if dense_kv:
                kv_off = (self.num_heads + self.num_kv_heads) * self.head_dim
                k = k.view(-1, self.num_kv_heads, self.head_dim)
                v = qkv[:, kv_off:].view(-1, self.num_kv_heads, self.head_dim)
            else:
                k, v = k_cache, v_cache
            o = ops.flash_attn_varlen_func(q, k, v,
                                           max_seqlen_q=context.max_seqlen_q, cu_seqlens_q=context.cu_seqlens_q,
                                           max_seqlen_k=context.max_seqlen_k, cu_seqlens_k=context.cu_seqlens_k,
                                           softmax_scale=self.scale, causal=True, block_table=context.block_tables)
        else:
            o = ops.flash_attn_with_kvcache(q.unsqueeze(1), k_cache, v_cache,
                                            cache_seqlens=context.context_lens, block_table=context.block_tables,
                                            softmax_scale=self.scale, causal=True)
        return o.view(-1, self.num_heads * self.head_dim)

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
        q = q.view(-1, self.num_heads, self.head_dim)
        k = k.view(-1, self.num_kv_heads, self.head_dim)
        v = v.view(-1, self.num_kv_heads, self.head_dim)

        context = _get_context()
        k_cache, v_cache = self.k_cache, self.v_cache
        if k_cache.numel() and v_cache.numel() and context.slot_mapping is not None:
            ops.store_kvcache(k, v, k_cache, v_cache, context.slot_mapping)

        if context.is_prefill:
            if context.block_tables is not None:  # prefix cache: keys come from the paged cache
                k, v = k_cache, v_cache
            o = ops.flash_attn_varlen_func(q, k, v,
                                           max_seqlen_q=context.max_seqlen_q, cu_seqlens_q=context.cu_seqlens_q,
                                           max_seqlen_k=context.max_seqlen_k, cu_seqlens_k=context.cu_seqlens_k,
                                           softmax_scale=self.scale, causal=True, block_table=context.block_tables)
        else:
            o = ops.flash_attn_with_kvcache(q.unsqueeze(1), k_cache, v_cache,
                                            cache_seqlens=context.context_lens, block_table=context.block_tables,
                                            softmax_scale=self.scale, causal=True)
        return o.view(-1, self.num_heads * self.head_dim)
