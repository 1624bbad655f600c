"""TEST INFRASTRUCTURE ONLY — never imported by the product path (nano-vllm-learn_b200/).

CPU oracle for the attention hot path of ztxdcyy/nano-vllm-learn.  Plain torch on CPU tensors,
fp32 arithmetic, no CUDA.  Every function cites the reference lines it restates.

Parity status: PINNED.  `tests/golden/make_golden.py` runs the reference's own CPU-runnable
functions (`nanovllm/layers/attention_sdpa.py::flash_attn_with_kvcache`, `::flash_attn_varlen_func`,
`::_gather_cache_tokens`) imported unmodified from /root/reference and freezes their outputs under
tests/golden/; tests/test_oracle_golden.py checks this oracle against those files (and against the
live reference in baseline/_ref when it is importable).  The prefix-cache prefill case has no
correct reference implementation that runs on CPU (the reference sdpa path is wrong there,
SURVEY §8c; flash-attn needs a GPU): it follows flash-attn's documented bottom-right causal
alignment, is pinned on CPU only through the reference's `_gather_cache_tokens` + the no-prefix
limit, and is cross-checked against flash-attn itself in the `-m gpu` tests.
"""
from __future__ import annotations

from typing import Optional

import torch


def store_kvcache_ref(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                      slot_mapping: torch.Tensor, skip_negative: bool = True) -> None:
    """In-place slot scatter.  Restates store_kvcache_kernel, nanovllm/layers/attention.py:30-41:
    row i of key/value (D = Hkv*head_dim contiguous elements) goes to flat cache row slot_mapping[i].
    Rows are applied in order, so a duplicated slot keeps the LAST writer's row only when all
    duplicates carry identical data (the only case the reference produces: zero-padded graph rows,
    model_runner.py:291-293).  skip_negative follows attention_triton.py:30-31."""
    n, hkv, d = key.shape
    row = hkv * d
    kc = k_cache.view(-1, row)
    vc = v_cache.view(-1, row)
    kk = key.reshape(n, row)
    vv = value.reshape(n, row)
    for i, s in enumerate(slot_mapping.tolist()):
        if s < 0:
            if skip_negative:
                continue
            raise IndexError("negative slot with skip_negative=False")
        kc[s] = kk[i]
        vc[s] = vv[i]


def gather_cache_tokens(cache: torch.Tensor, block_ids, total_tokens: int) -> torch.Tensor:
    """First `total_tokens` tokens of a sequence from the paged cache [num_blocks, block, Hkv, D].
    Restates _gather_cache_tokens, nanovllm/layers/attention_sdpa.py:185-205."""
    block_size = cache.shape[1]
    out = []
    left = total_tokens
    for bid in (block_ids.tolist() if isinstance(block_ids, torch.Tensor) else list(block_ids)):
        if bid < 0 or left <= 0:
            break
        take = min(block_size, left)
        out.append(cache[bid, :take])
        left -= take
    if left > 0:
        raise RuntimeError(f"block table too short: {left} of {total_tokens} tokens missing")
    if not out:
        return cache.new_zeros((0, cache.shape[2], cache.shape[3]))
    return torch.cat(out, dim=0)


def _attend(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, scale: float, mask: Optional[torch.Tensor]):
    """q [Lq, Hq, D], k/v [Lk, Hkv, D] -> [Lq, Hq, D] fp32; GQA by head // G (SDPA enable_gqa=True,
    attention_sdpa.py:104-112); softmax in fp32."""
    lq, hq, d = q.shape
    lk, hkv, _ = k.shape
    g = hq // hkv
    qf = q.float().view(lq, hkv, g, d)
    kf = k.float()
    vf = v.float()
    s = torch.einsum("qhgd,khd->hgqk", qf, kf) * scale
    if mask is not None:
        s = s.masked_fill(~mask.view(1, 1, lq, lk), float("-inf"))
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)  # a row with no visible key yields zeros (flash-attn behaviour)
    o = torch.einsum("hgqk,khd->qhgd", p, vf)
    return o.reshape(lq, hq, d)


def paged_decode_ref(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, context_lens: torch.Tensor,
                     block_tables: torch.Tensor, scale: float) -> torch.Tensor:
    """q [bs, Hq, D] -> [bs, Hq, D] in q.dtype.  Restates the decode branch,
    nanovllm/layers/attention.py:97-101 with the semantics of attention_sdpa.py:122-182: sequence b
    attends tokens 0..context_lens[b]-1 found through block_tables[b]; ctx == 0 -> zeros."""
    bs = q.shape[0]
    out = torch.zeros(q.shape, dtype=torch.float32)
    for b in range(bs):
        ctx = int(context_lens[b])
        if ctx <= 0:
            continue
        k = gather_cache_tokens(k_cache, block_tables[b], ctx)
        v = gather_cache_tokens(v_cache, block_tables[b], ctx)
        out[b] = _attend(q[b:b + 1], k, v, scale, None)[0]
    return out.to(q.dtype)


def varlen_prefill_ref(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: torch.Tensor,
                       cu_seqlens_k: torch.Tensor, scale: float,
                       block_tables: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q [sum Lq, Hq, D] -> same shape.  Restates the prefill branch, nanovllm/layers/attention.py:89-96.
    Without block_tables follows attention_sdpa.py:65-119 (per-sequence causal SDPA; Lq == Lk so
    top-left and bottom-right masks coincide).  With block_tables, keys are gathered from the paged
    cache and query j sees keys 0..(Lk-Lq)+j — flash-attn's bottom-right alignment
    (flash_attn_interface.py docstring of flash_attn_varlen_func, `causal`)."""
    out = torch.zeros(q.shape, dtype=torch.float32)
    batch = cu_seqlens_q.numel() - 1
    for i in range(batch):
        q0, q1 = int(cu_seqlens_q[i]), int(cu_seqlens_q[i + 1])
        k0, k1 = int(cu_seqlens_k[i]), int(cu_seqlens_k[i + 1])
        lq, lk = q1 - q0, k1 - k0
        if lq == 0:
            continue
        if block_tables is not None:
            ki = gather_cache_tokens(k, block_tables[i], lk)
            vi = gather_cache_tokens(v, block_tables[i], lk)
        else:
            ki, vi = k[k0:k1], v[k0:k1]
        mask = torch.ones(lq, lk, dtype=torch.bool).tril(diagonal=lk - lq)
        out[q0:q1] = _attend(q[q0:q1], ki, vi, scale, mask)
    return out.to(q.dtype)


def qk_norm_rope_ref(qkv: torch.Tensor, positions: torch.Tensor, q_weight: torch.Tensor, k_weight: torch.Tensor,
                     eps: float, cos_sin_cache: torch.Tensor, num_q_heads: int, num_kv_heads: int, head_dim: int = 128):
    """(q, k, v) after per-head RMSNorm and NeoX RoPE, each [N, heads, D] in qkv.dtype; v is a plain slice.
    Restates RMSNorm.rms_forward (nanovllm/layers/layernorm.py:17-27: fp32 normalise, round to bf16,
    multiply by the weight in bf16) applied per head as in qwen3.py:108-114, then
    apply_rotary_emb / RotaryEmbedding.forward (nanovllm/layers/rotary_embedding.py:6-55: fp32,
    y1 = x1*cos - x2*sin, y2 = x2*cos + x1*sin on the two halves, round to the input dtype)."""
    n = qkv.shape[0]
    qs, ks = num_q_heads * head_dim, num_kv_heads * head_dim
    q, k, v = qkv.split([qs, ks, ks], dim=-1)

    def rms(x, w):
        xf = x.reshape(n, -1, head_dim).to(torch.float32)
        var = xf.pow(2).mean(dim=-1, keepdim=True)
        xf = xf * torch.rsqrt(var + eps)
        return xf.to(qkv.dtype) * w

    def rope(x):
        cs = cos_sin_cache.view(cos_sin_cache.shape[0], -1)[positions]
        cos, sin = cs.chunk(2, dim=-1)
        cos, sin = cos.unsqueeze(-2), sin.unsqueeze(-2)
        x1, x2 = torch.chunk(x.to(torch.float32), 2, dim=-1)
        return torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin), dim=-1).to(x.dtype)

    return rope(rms(q, q_weight)), rope(rms(k, k_weight)), v.reshape(n, num_kv_heads, head_dim)


def decode_algorithmic_bytes(context_lens, num_q_heads: int, num_kv_heads: int, head_dim: int = 128,
                             block_size: int = 256, elem: int = 2) -> int:
    """K+V rows read once + q in + o out + context_lens + live table entries."""
    lens = [int(x) for x in context_lens]
    bs = len(lens)
    kv = sum(lens) * num_kv_heads * head_dim * 2 * elem
    qo = 2 * bs * num_q_heads * head_dim * elem
    idx = 4 * bs + 4 * sum((c + block_size - 1) // block_size for c in lens)
    return kv + qo + idx
