"""The `Context` record against the reference's own module (`nanovllm/utils/context.py:5-27`):
same fields and defaults, and the same set/get/reset protocol."""
import dataclasses
import importlib
import os
import sys

import pytest
import torch

from nano_vllm_learn_b200 import Context, get_context, reset_context, set_context

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference_context_module():
    for base in ("/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        path = os.path.join(base, "nanovllm", "utils", "context.py")
        if os.path.exists(path):
            spec = importlib.util.spec_from_file_location("_ref_context", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    return None


def _args():
    return dict(cu_seqlens_q=torch.tensor([0, 3], dtype=torch.int32), cu_seqlens_k=torch.tensor([0, 9], dtype=torch.int32),
                max_seqlen_q=3, max_seqlen_k=9, slot_mapping=torch.arange(3, dtype=torch.int32),
                context_lens=torch.tensor([9], dtype=torch.int32), block_tables=torch.tensor([[4, -1]], dtype=torch.int32))


def test_reset_context_restores_defaults():
    set_context(True, **_args())
    got = get_context()
    assert got.is_prefill and got.max_seqlen_q == 3 and got.block_tables is not None
    reset_context()
    assert get_context() == Context()
    assert get_context().is_prefill is False and get_context().slot_mapping is None


def test_set_context_replaces_the_whole_record():
    set_context(True, **_args())
    set_context(False, slot_mapping=torch.zeros(1, dtype=torch.int32))
    ctx = get_context()
    assert ctx.is_prefill is False and ctx.cu_seqlens_q is None and ctx.max_seqlen_k == 0 and ctx.block_tables is None
    reset_context()


def test_same_schema_and_protocol_as_reference():
    ref = _reference_context_module()
    if ref is None:
        pytest.skip("reference not available")
    ours = [(f.name, f.default) for f in dataclasses.fields(Context)]
    theirs = [(f.name, f.default) for f in dataclasses.fields(ref.Context)]
    assert ours == theirs
    args = _args()
    for step in ("set_prefill", "reset", "set_decode", "reset"):
        if step == "set_prefill":
            set_context(True, **args), ref.set_context(True, **args)
        elif step == "set_decode":
            set_context(False, slot_mapping=args["slot_mapping"], context_lens=args["context_lens"],
                        block_tables=args["block_tables"])
            ref.set_context(False, slot_mapping=args["slot_mapping"], context_lens=args["context_lens"],
                            block_tables=args["block_tables"])
        else:
            reset_context(), ref.reset_context()
        a, b = get_context(), ref.get_context()
        for f in dataclasses.fields(Context):
            va, vb = getattr(a, f.name), getattr(b, f.name)
            assert (va is vb) or va == vb, (step, f.name)
