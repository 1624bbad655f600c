"""CPU, world_size 2 under gloo: the host-side gating of the two embed_head.py patches (INTEGRATION.md edits A8 / A9).

  * vocabulary-parallel sampling OFF (or another backend): the patched `ParallelLMHead.forward` IS the reference's
    (dist.gather to rank 0 + torch.cat, nanovllm/layers/embed_head.py:61-71);  ON: every rank keeps its own logits shard
    -- exactly the reference's pre-gather tensor -- and no collective runs.
  * the patched `VocabParallelEmbedding.forward` hands CPU tensors, tp_size 1, > 512 tokens and NVLLM_FUSED_EMBEDDING=0
    to the reference forward (masked lookup + dist.all_reduce, embed_head.py:35-43); the CUDA op is never reached here.
  * both ranks size their exchange buffers identically (pure host arithmetic of the C ABI).
The reference classes are loaded from their source file with package shells, as the oracles do."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from oracle import sampler_oracle as so  # noqa: E402


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import importlib.util
    import types
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import cases
        root = so._ref_root()
        for name in ("nanovllm", "nanovllm.utils"):
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m

        def load(mod_name, rel):
            spec = importlib.util.spec_from_file_location(mod_name, os.path.join(root, *rel))
            mod = importlib.util.module_from_spec(spec)
            sys.modules[mod_name] = mod
            spec.loader.exec_module(mod)
            return mod
        ctx = load("nanovllm.utils.context", ("nanovllm", "utils", "context.py"))
        eh = load("ref_embed_head", ("nanovllm", "layers", "embed_head.py"))
        from nano_vllm_learn_b200 import embedding_cuda, sampler_cuda
        ref_head_forward = eh.ParallelLMHead.forward
        ref_emb_forward = eh.VocabParallelEmbedding.forward
        sampler_cuda.install_lm_head(eh.ParallelLMHead)           # edit A8
        embedding_cuda.install(eh.VocabParallelEmbedding)         # edit A9
        ok = eh.ParallelLMHead.forward is not ref_head_forward and eh.VocabParallelEmbedding.forward is not ref_emb_forward

        c = cases.make_lm_head_case("vp_decode_tp2")
        head = eh.ParallelLMHead(c["weight"].shape[0], c["weight"].shape[1]).to(torch.bfloat16)
        head.weight.weight_loader(head.weight, c["weight"])
        shards = so.lm_head_shards_ref(c["x"], c["weight"], world)
        ctx.set_context(False)
        # another backend / switched off: the reference's gather, logits on rank 0 only
        for env, backend in (("1", False), ("0", True)):
            os.environ["NVLLM_VOCAB_PARALLEL_SAMPLER"] = env
            sampler_cuda.configure(backend)
            ok = ok and not sampler_cuda.vocab_parallel_enabled()
            with torch.no_grad():
                y = head(c["x"])
            ok = ok and (torch.equal(y, so.lm_head_gather_ref(shards)) if rank == 0 else y is None)
        # on: every rank keeps its shard, bit-identical to the reference's pre-gather logits; the fused sampler is needed
        os.environ.pop("NVLLM_VOCAB_PARALLEL_SAMPLER")
        sampler_cuda.configure(True)
        ok = ok and sampler_cuda.vocab_parallel_enabled()
        with torch.no_grad():
            y = head(c["x"])
        ok = ok and y is not None and torch.equal(y, shards[rank])
        os.environ["NVLLM_FUSED_SAMPLER"] = "0"
        sampler_cuda.configure(True)
        ok = ok and not sampler_cuda.vocab_parallel_enabled()
        os.environ.pop("NVLLM_FUSED_SAMPLER")
        # prefill context: last token of every sequence, as the reference selects them
        c2 = cases.make_lm_head_case("vp_prefill_tp2")
        head2 = eh.ParallelLMHead(c2["weight"].shape[0], c2["weight"].shape[1]).to(torch.bfloat16)
        head2.weight.weight_loader(head2.weight, c2["weight"])
        sampler_cuda.configure(True)
        ctx.set_context(True, cu_seqlens_q=c2["cu_seqlens_q"])
        with torch.no_grad():
            y2 = head2(c2["x"])
        last = (c2["cu_seqlens_q"][1:] - 1).long()
        ok = ok and torch.equal(y2, so.lm_head_shards_ref(c2["x"], c2["weight"], world, last)[rank])
        ctx.reset_context()

        # embedding: CPU tensors always take the reference path, whatever the switch says
        e = cases.make_embedding_case("emb_tp2")
        emb = eh.VocabParallelEmbedding(e["weight"].shape[0], e["weight"].shape[1]).to(torch.bfloat16)
        emb.weight.weight_loader(emb.weight, e["weight"])
        for env in ("1", "0"):
            os.environ["NVLLM_FUSED_EMBEDDING"] = env
            embedding_cuda.configure(True)
            ok = ok and embedding_cuda.enabled() == (env == "1")
            with torch.no_grad():
                ye = emb(e["x"])
            ok = ok and torch.equal(ye, e["weight"][e["x"]])
        embedding_cuda.configure(False)
        ok = ok and not embedding_cuda.enabled()
        sizes = [None] * world
        dist.all_gather_object(sizes, (sampler_cuda.comm_bytes(world), embedding_cuda.comm_bytes(world, 1024)))
        ok = ok and len(set(sizes)) == 1 and min(sizes[0]) > 4096
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_lm_head_and_embedding_patches_world2_gloo(lib):
    if so._ref_root() is None:
        pytest.skip("reference tree not on this machine")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29850 + os.getpid() % 100
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res == {0: True, 1: True}
