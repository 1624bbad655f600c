"""TEST INFRASTRUCTURE ONLY — never imported by the product path.

Oracle for the runner glue index math.  Two layers:

1. `*_ref` functions: plain-Python restatements of `ModelRunner.prepare_block_tables /
   prepare_prefill / prepare_decode` (nanovllm/engine/model_runner.py:160-269) returning CPU tensors
   with the reference's dtypes.
2. `ReferenceGlue`: runs the reference's UNMODIFIED methods on CPU by binding them to a stub `self`
   and neutralising the only CUDA-only calls (`pin_memory=True`, `.cuda()`), as probed in SURVEY §8c.
   Used by tests/golden/make_golden.py (against /root/reference) and by the live check in
   tests/test_runner_glue.py when baseline/_ref is importable.  Parity status: PINNED by (2).
"""
from __future__ import annotations

import importlib
import sys
import types
from typing import Optional

import torch


# ------------------------------------------------------------------ restatement
def prepare_block_tables_ref(seqs) -> torch.Tensor:
    """model_runner.py:160-169: right-pad with -1 to the longest table, int32."""
    width = max(len(s.block_table) for s in seqs)
    rows = [list(s.block_table) + [-1] * (width - len(s.block_table)) for s in seqs]
    return torch.tensor(rows, dtype=torch.int32)


def prepare_prefill_ref(seqs, block_size: int = 256) -> dict:
    """model_runner.py:171-242."""
    input_ids, positions, slots = [], [], []
    cu_q, cu_k = [0], [0]
    max_q = max_k = 0
    for s in seqs:
        n = len(s)
        cached = s.num_cached_tokens
        input_ids.extend(s[cached:])
        positions.extend(range(cached, n))
        cu_q.append(cu_q[-1] + n - cached)
        cu_k.append(cu_k[-1] + n)
        max_q = max(max_q, n - cached)
        max_k = max(max_k, n)
        if not s.block_table:
            continue
        num_blocks = (n + block_size - 1) // block_size
        last = n - (num_blocks - 1) * block_size
        for i in range(cached // block_size, num_blocks):
            start = s.block_table[i] * block_size
            slots.extend(range(start, start + (block_size if i != num_blocks - 1 else last)))
    tables = prepare_block_tables_ref(seqs) if cu_k[-1] > cu_q[-1] else None
    return dict(
        input_ids=torch.tensor(input_ids, dtype=torch.int64),
        positions=torch.tensor(positions, dtype=torch.int64),
        cu_seqlens_q=torch.tensor(cu_q, dtype=torch.int32),
        cu_seqlens_k=torch.tensor(cu_k, dtype=torch.int32),
        max_seqlen_q=max_q, max_seqlen_k=max_k,
        slot_mapping=torch.tensor(slots, dtype=torch.int32),
        block_tables=tables,
    )


def prepare_decode_ref(seqs, block_size: int = 256) -> dict:
    """model_runner.py:244-269 — note positions = len(seq) (sic)."""
    ids, pos, slots, lens = [], [], [], []
    for s in seqs:
        n = len(s)
        ids.append(s.last_token)
        pos.append(n)
        lens.append(n)
        num_blocks = (n + block_size - 1) // block_size
        last = n - (num_blocks - 1) * block_size
        slots.append(s.block_table[-1] * block_size + last - 1)
    return dict(
        input_ids=torch.tensor(ids, dtype=torch.int64),
        positions=torch.tensor(pos, dtype=torch.int64),
        slot_mapping=torch.tensor(slots, dtype=torch.int32),
        context_lens=torch.tensor(lens, dtype=torch.int32),
        block_tables=prepare_block_tables_ref(seqs),
    )


# ------------------------------------------------------------------ the real thing, on CPU
class _TorchNoPin:
    """Proxy for the `torch` name inside model_runner: tensor(..., pin_memory=True) -> tensor(...)."""

    def __init__(self, real):
        self._real = real

    def __getattr__(self, name):
        return getattr(self._real, name)

    def tensor(self, data, *args, **kwargs):
        kwargs.pop("pin_memory", None)
        return self._real.tensor(data, *args, **kwargs)


class ReferenceGlue:
    """Drives the reference's own prepare_* on CPU.  `ref_root` is a directory holding `nanovllm/`."""

    def __init__(self, ref_root: str, block_size: int = 256):
        self.ref_root = ref_root
        if ref_root not in sys.path:
            sys.path.insert(0, ref_root)
        self.mr = importlib.import_module("nanovllm.engine.model_runner")
        self.ctx = importlib.import_module("nanovllm.utils.context")
        self.sequence = importlib.import_module("nanovllm.engine.sequence")
        self.block_manager = importlib.import_module("nanovllm.engine.block_manager")
        self.sampling = importlib.import_module("nanovllm.sampling_params")
        self.stub = types.SimpleNamespace(block_size=block_size)
        self.stub.prepare_block_tables = types.MethodType(self.mr.ModelRunner.prepare_block_tables, self.stub)

    def _neutralised(self):
        mr = self.mr
        outer = self

        class _Guard:
            def __enter__(self_inner):
                self_inner.saved_torch = mr.torch
                self_inner.saved_cuda = torch.Tensor.cuda
                mr.torch = _TorchNoPin(mr.torch)
                torch.Tensor.cuda = lambda t, *a, **k: t
                return outer

            def __exit__(self_inner, *exc):
                mr.torch = self_inner.saved_torch
                torch.Tensor.cuda = self_inner.saved_cuda
                return False

        return _Guard()

    def prepare_prefill(self, seqs) -> dict:
        with self._neutralised():
            ids, pos = self.mr.ModelRunner.prepare_prefill(self.stub, seqs)
            c = self.ctx.get_context()
            out = dict(input_ids=ids, positions=pos, cu_seqlens_q=c.cu_seqlens_q, cu_seqlens_k=c.cu_seqlens_k,
                       max_seqlen_q=c.max_seqlen_q, max_seqlen_k=c.max_seqlen_k, slot_mapping=c.slot_mapping,
                       block_tables=c.block_tables)
            self.ctx.reset_context()
        return out

    def prepare_decode(self, seqs) -> dict:
        with self._neutralised():
            ids, pos = self.mr.ModelRunner.prepare_decode(self.stub, seqs)
            c = self.ctx.get_context()
            out = dict(input_ids=ids, positions=pos, slot_mapping=c.slot_mapping, context_lens=c.context_lens,
                       block_tables=c.block_tables)
            self.ctx.reset_context()
        return out


def find_reference_root() -> Optional[str]:
    """/root/reference here; baseline/_ref on the GPU box (the installed, unmodified copy)."""
    import os
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for cand in ("/root/reference", os.path.join(here, "baseline", "_ref")):
        if os.path.isdir(os.path.join(cand, "nanovllm")):
            return cand
    return None


# ------------------------------------------------------------------ device-side decode state
def advance_decode_state_ref(state: dict, sampled, patches, pad_cols, block_size: int = 256) -> dict:
    """What csrc/decode_state.cu does to the published decode tensors between two steps of an UNCHANGED batch, restated
    on CPU: `state` is the dict prepare_decode_ref returned for the previous step (input_ids, positions, slot_mapping,
    context_lens, block_tables), `sampled` the token ids sampled in that step, `patches` = [(row, column, block id)] for
    the blocks BlockManager.may_append opened (block_manager.py:134-159), `pad_cols` = (begin, end) table columns that
    join the live width this step (they carry the reference's -1 padding, model_runner.py:163-166)."""
    ids = torch.as_tensor(sampled, dtype=torch.int64).clone()
    lens = state["context_lens"].clone() + 1
    tab = state["block_tables"].clone()
    if pad_cols[1] > tab.shape[1]:
        tab = torch.cat([tab, torch.zeros(tab.shape[0], pad_cols[1] - tab.shape[1], dtype=tab.dtype)], dim=1)
    tab[:, pad_cols[0]:pad_cols[1]] = -1
    for r, c, v in patches:
        tab[r, c] = v
    last = lens - 1
    blk = tab[torch.arange(tab.shape[0]), (last // block_size).long()]
    slots = (blk * block_size + last % block_size).to(torch.int32)
    return dict(input_ids=ids, positions=lens.to(torch.int64), slot_mapping=slots, context_lens=lens.to(torch.int32),
                block_tables=tab)
