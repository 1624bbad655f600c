"""TEST INFRASTRUCTURE ONLY — never imported by the product path.

Oracle for the fused VocabParallelEmbedding (third collective of a tensor-parallel step).
  * `vocab_embedding_ref`: CPU restatement of `VocabParallelEmbedding.forward` (nanovllm/layers/embed_head.py:35-43)
    for ALL ranks at once: masked lookup on every rank's shard, foreign rows zeroed, then the all-reduce as a sum.
    Exactly one rank contributes a non-zero row per token, so the sum is that row (up to the sign of zeros).
  * `run_reference_world`: the reference's UNMODIFIED class under torch.distributed (gloo, CPU, one process per
    rank) -- freezes tests/golden/vocab_embedding.npz (tests/golden/make_golden_embedding.py) and re-checks live.
Parity status: PINNED (golden outputs of the reference class itself, reproduced by `vocab_embedding_ref`).
"""
from __future__ import annotations

import os

import torch
import torch.nn.functional as F


def vocab_embedding_ref(x: torch.Tensor, weight: torch.Tensor, world: int) -> torch.Tensor:
    per = weight.shape[0] // world
    y = None
    for r in range(world):
        start, end = r * per, (r + 1) * per                      # embed_head.py:21-23
        mask = (x >= start) & (x < end)                          # :37
        xl = mask * (x - start)                                  # :38
        part = F.embedding(xl, weight[start:end])                # :39
        part = mask.unsqueeze(1) * part                          # :41
        y = part if y is None else y + part                      # :42 dist.all_reduce (SUM)
    return y


def _ref_root():
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for cand in ("/root/reference", os.path.join(here, "baseline", "_ref")):
        if os.path.isdir(os.path.join(cand, "nanovllm")):
            return cand
    return None


def _worker(rank, world, port, x, weight, out_q):
    import importlib.util
    import sys
    import types
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        root = _ref_root()
        for name in ("nanovllm", "nanovllm.utils"):          # package shells: the real package pulls flash-attn/triton
            if name not in sys.modules:
                m = types.ModuleType(name)
                m.__path__ = []
                sys.modules[name] = m

        def load(mod_name, rel):
            spec = importlib.util.spec_from_file_location(mod_name, os.path.join(root, *rel))
            mod = importlib.util.module_from_spec(spec)
            sys.modules[mod_name] = mod
            spec.loader.exec_module(mod)
            return mod
        load("nanovllm.utils.context", ("nanovllm", "utils", "context.py"))
        eh = load("ref_embed_head", ("nanovllm", "layers", "embed_head.py"))
        emb = eh.VocabParallelEmbedding(weight.shape[0], weight.shape[1]).to(weight.dtype)
        emb.weight.weight_loader(emb.weight, weight)             # the reference's own sharding
        with torch.no_grad():
            y = emb(x)
        out_q.put((rank, y.view(torch.int16).numpy()))
    finally:
        dist.destroy_process_group()


def run_reference_world(x: torch.Tensor, weight: torch.Tensor, world: int, port: int = 29801):
    import torch.multiprocessing as mp
    if _ref_root() is None:
        raise FileNotFoundError("reference tree not found (/root/reference or baseline/_ref)")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, x, weight, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    return [torch.from_numpy(got[r].copy()).view(torch.bfloat16) for r in range(world)]
