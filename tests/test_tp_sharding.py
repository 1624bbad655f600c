"""CPU, gloo, world_size 2: the multi-GPU path shards attention by KV head with NO data-path
collective.  Each rank runs its head shard; gathering the shards along the head axis
must reproduce the unsharded result bit for bit, and the runner glue must publish identical
metadata on every rank.
Scope: this is a property of the ALGORITHM (heads are independent), shown on the CPU oracle because no CUDA kernel can run
here; the product kernels at head-sharded shapes and the engine at tensor_parallel_size 2 are covered on the box
(tests/test_gpu_parity.py, tests/test_gpu_engine_tp.py, bench.py tp_strong); the host logic of the N > 1 collectives has
its own gloo world-2 tests (tests/test_rowparallel_oracle.py, tests/test_tp_head_patches_gloo.py)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cases
from oracle import paged_attention_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q_out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from head_sharding import shard_heads
        c = cases.make_decode_case("c1_bs1_ctx512")
        hq, hkv = c["hq"], c["hkv"]
        q_lo, q_hi, kv_lo, kv_hi = shard_heads(rank, world, hq, hkv)
        out = orc.paged_decode_ref(c["q"][:, q_lo:q_hi], c["k_cache"][:, :, kv_lo:kv_hi].contiguous(),
                                   c["v_cache"][:, :, kv_lo:kv_hi].contiguous(), c["context_lens"], c["block_tables"],
                                   c["scale"])
        parts = [torch.zeros_like(out) for _ in range(world)]
        dist.all_gather(parts, out.contiguous())
        full = torch.cat(parts, dim=1)
        want = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], c["scale"])
        ok = torch.equal(full.view(torch.int16), want.view(torch.int16))
        # identical metadata on every rank (every rank runs the same prepare_* on the same seqs)
        meta = torch.cat([c["context_lens"].flatten(), c["block_tables"].flatten()]).to(torch.int64)
        gathered = [torch.zeros_like(meta) for _ in range(world)]
        dist.all_gather(gathered, meta)
        ok = ok and all(torch.equal(gathered[0], g) for g in gathered)
        if rank == 0:
            q_out.put(ok)
    finally:
        dist.destroy_process_group()


def test_kv_head_sharding_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_shard_heads_partition():
    from head_sharding import shard_heads
    for hq, hkv in ((16, 8), (32, 8)):
        for world in (1, 2, 4, 8):
            seen_q, seen_kv = [], []
            for r in range(world):
                q_lo, q_hi, kv_lo, kv_hi = shard_heads(r, world, hq, hkv)
                assert (q_hi - q_lo) == hq // world and (kv_hi - kv_lo) == hkv // world
                assert q_lo // (hq // hkv) == kv_lo          # q heads of a rank use only that rank's kv heads
                seen_q += list(range(q_lo, q_hi))
                seen_kv += list(range(kv_lo, kv_hi))
            assert seen_q == list(range(hq)) and seen_kv == list(range(hkv))
    with pytest.raises(AssertionError):
        shard_heads(0, 3, 16, 8)
