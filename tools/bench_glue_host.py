"""Host time per decode step of the runner glue: the staged path (host lists -> pinned ring -> one H2D copy) beside the
device-side path (nvllm_advance_decode_state), and the reference's own prepare_decode
(nanovllm/engine/model_runner.py:244-269) where baseline/_ref is importable.  perf_counter around the call only; the
GPU work is asynchronous and drained once at the end.

    python tools/bench_glue_host.py [--bs 64] [--steps 2000]
"""
from __future__ import annotations

import argparse
import itertools
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
BS = 256


class Seq:
    _ids = itertools.count()

    def __init__(self, n, table):
        self.seq_id = next(Seq._ids)
        self.num_tokens = n
        self.block_table = list(table)
        self.last_token = 1
        self.num_cached_tokens = 0
        self.temperature = 0.0

    def __len__(self):
        return self.num_tokens

    @property
    def last_block_num_tokens(self):
        return self.num_tokens - (len(self.block_table) - 1) * BS


def make(bs, ctx):
    free = itertools.count(10)
    return [Seq(ctx + 3 * i, [next(free) for _ in range((ctx + 3 * i + BS - 1) // BS)]) for i in range(bs)], free


def advance(seqs, free, step):
    for i, s in enumerate(seqs):
        s.last_token = (step + i) % 10000
        s.num_tokens += 1
        if s.num_tokens % BS == 1:
            s.block_table.append(next(free))


def run(bs, steps, ctx, device_state, graphs):
    from nano_vllm_learn_b200.runner_glue import RunnerGlue
    os.environ["NVLLM_DEVICE_DECODE_STATE"] = "1" if device_state else "0"
    g = RunnerGlue(BS, 512, 4096, 16384, "cuda", lambda *a, **k: None)
    if graphs:
        g.graph_bs = [1, 2, 4, 8] + list(range(16, 513, 16))
    seqs, free = make(bs, ctx)
    toks = torch.zeros(bs, dtype=torch.int64, device="cuda")
    g.prepare_decode(seqs, toks)
    spent = 0.0
    for step in range(steps):
        advance(seqs, free, step)
        t0 = time.perf_counter()
        g.prepare_decode(seqs, toks)
        spent += time.perf_counter() - t0
    torch.cuda.synchronize()
    return round(spent / steps * 1e6, 2), g.fast_steps, g.staged_steps


def run_reference(bs, steps, ctx):
    """The reference's UNMODIFIED ModelRunner.prepare_decode (nanovllm/engine/model_runner.py:244-269), imported from
    baseline/_ref (or /root/reference) and called on a stub that carries only block_size -- on the GPU box its pinned
    tensors and .cuda() copies are the real thing.  None when the reference is not importable."""
    import importlib
    import types
    root = next((r for r in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference")
                 if os.path.isdir(os.path.join(r, "nanovllm"))), None)
    if root is None:
        return None
    try:
        sys.path.insert(0, root)
        mr = importlib.import_module("nanovllm.engine.model_runner")
        refctx = importlib.import_module("nanovllm.utils.context")
    except Exception:
        return None
    stub = types.SimpleNamespace(block_size=BS)
    stub.prepare_block_tables = types.MethodType(mr.ModelRunner.prepare_block_tables, stub)
    seqs, free = make(bs, ctx)
    spent = 0.0
    n = min(steps, 300)
    for step in range(n):
        advance(seqs, free, step)
        t0 = time.perf_counter()
        mr.ModelRunner.prepare_decode(stub, seqs)
        spent += time.perf_counter() - t0
        refctx.reset_context()
    torch.cuda.synchronize()
    return round(spent / n * 1e6, 2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bs", type=int, nargs="+", default=[16, 64, 256])
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--ctx", type=int, default=600)
    a = ap.parse_args()
    for bs in a.bs:
        row = {"bs": bs, "ctx_start": a.ctx, "steps": a.steps}
        for graphs in (True, False):
            tag = "graph_layout" if graphs else "eager_layout"
            us, fast, staged = run(bs, a.steps, a.ctx, True, graphs)
            row[f"{tag}_device_state_us"] = us
            row[f"{tag}_device_state_fast_steps"] = fast
            us, _, _ = run(bs, a.steps, a.ctx, False, graphs)
            row[f"{tag}_staged_us"] = us
        row["reference_prepare_decode_us"] = run_reference(bs, a.steps, a.ctx)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
