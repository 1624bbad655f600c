"""Per-step attention metadata record: the drop-in boundary of the hot path.

Mirror of the reference contract `nanovllm/utils/context.py:5-27` (same field names, defaults and
module-global get/set/reset protocol) so that `Attention.forward(q, k, v)` reads exactly what the
reference `ModelRunner.prepare_prefill/prepare_decode` publish.  When this package is plugged into
nano-vllm the reference's own module is used instead (see `attention_cuda.bind_context`).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch


@dataclass
class Context:
    # prefill: varlen causal attention over the new tokens; decode: one query token per sequence
    is_prefill: bool = False
    # int32 [B+1] exclusive prefix sums of new-token counts / key counts (prefill only)
    cu_seqlens_q: Optional[torch.Tensor] = None
    cu_seqlens_k: Optional[torch.Tensor] = None
    max_seqlen_q: int = 0
    max_seqlen_k: int = 0
    # int32 [N] flat cache slot (block * block_size + offset) of every token stored this step
    slot_mapping: Optional[torch.Tensor] = None
    # int32 [B] keys per sequence, including the token stored this step (decode only)
    context_lens: Optional[torch.Tensor] = None
    # int32 [B, W] page table; None in prefill without a prefix-cache hit
    block_tables: Optional[torch.Tensor] = None


# One record per process: each tensor-parallel rank is its own spawned process (llm_engine.py:29-36) and runs
# a single Python thread, so a module-level slot is all the engine needs.
_ACTIVE = Context()


def get_context() -> Context:
    return _ACTIVE


def set_context(is_prefill, cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=0, max_seqlen_k=0,
                slot_mapping=None, context_lens=None, block_tables=None) -> None:
    global _ACTIVE
    _ACTIVE = Context(is_prefill, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                      slot_mapping, context_lens, block_tables)


def reset_context() -> None:
    """Back to the all-defaults record, like the reference (`utils/context.py:25-27`)."""
    global _ACTIVE
    _ACTIVE = Context()
