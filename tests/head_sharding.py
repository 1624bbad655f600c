"""TEST HELPER (not part of the product package): KV-head sharding of the attention path under tensor parallelism.

Mirrors `Qwen3Attention.__init__` (nanovllm/models/qwen3.py:30-39): with tensor_parallel_size `world`
each rank owns `Hq / world` query heads and `Hkv / world` kv heads — contiguous ranges, because
`QKVParallelLinear` slices the fused projection by rank (nanovllm/layers/linear.py).  Attention needs
no collective: heads are independent and every rank sees identical slot_mapping / block_tables /
context_lens.  The product never shards anything itself: the reference's model code hands each rank its heads, and the
kernels are checked at the sharded shapes on the box (tests/test_gpu_parity.py decode cases g1/g8, bench.py tp_strong,
tests/test_gpu_engine_tp.py at tensor_parallel_size 2).
"""
from __future__ import annotations


def shard_heads(rank: int, world: int, num_q_heads: int, num_kv_heads: int) -> tuple[int, int, int, int]:
    """(q_lo, q_hi, kv_lo, kv_hi) head ranges of `rank`."""
    assert 0 <= rank < world
    assert num_q_heads % world == 0 and num_kv_heads % world == 0, "heads must divide the tensor-parallel size"
    hq, hkv = num_q_heads // world, num_kv_heads // world
    return rank * hq, (rank + 1) * hq, rank * hkv, (rank + 1) * hkv
