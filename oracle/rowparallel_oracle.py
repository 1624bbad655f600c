"""TEST INFRASTRUCTURE ONLY — never imported by the product path.

Oracle for the fused RowParallelLinear: GEMM + tensor-parallel all-reduce.
  * `rowparallel_ref`: CPU restatement of `RowParallelLinear.forward` (nanovllm/layers/linear.py:185-190) for ALL
    ranks at once: every rank's `F.linear(x_r, W_r, bias if r == 0 else None)` is a bf16 tensor, and the all-reduce
    adds those bf16 tensors.  For world == 2 the sum is a single bf16 addition and therefore unambiguous; for
    world > 2 the order (and NCCL's intermediate precision) is implementation-defined, so the rank-order bf16
    accumulation returned here is one representative and tests bound it rather than match it bit for bit.
  * `rowparallel_exact`: the same contraction in float64 without intermediate rounding — what the CUDA kernel
    approximates with fp32 accumulation and ONE final rounding to bf16.
  * `run_reference_world`: the reference's UNMODIFIED `RowParallelLinear` under torch.distributed (gloo, CPU, one
    process per rank) — used by tests/golden/make_golden_rowparallel.py to freeze tests/golden/rowparallel.npz and,
    where /root/reference (or baseline/_ref) is present, by the tests to re-check the restatement live.
  * `shard_inputs`: how the reference shards x (the o_proj input is the local heads' attention output, i.e. the
    rank-th chunk of the feature axis) and W (`weight_loader` narrows dim 1, linear.py:177-183).
Parity status: PINNED for world == 2 (golden outputs of the reference class itself, reproduced bit for bit by
`rowparallel_ref`); for world > 2 pinned up to the summation order of the collective.
"""
from __future__ import annotations

import os
import sys

import torch
import torch.nn.functional as F


def shard_inputs(x: torch.Tensor, weight: torch.Tensor, world: int):
    """x [n, in] -> world chunks of the feature axis; weight [out, in] -> world chunks of dim 1 (linear.py:177-183)."""
    xs = [c.contiguous() for c in x.chunk(world, dim=1)]
    ws = [c.contiguous() for c in weight.chunk(world, dim=1)]
    return xs, ws


def rowparallel_ref(xs, ws, bias=None) -> torch.Tensor:
    """linear.py:185-190 on every rank, then the all-reduce as a rank-order bf16 accumulation."""
    y = None
    for r, (x, w) in enumerate(zip(xs, ws)):
        part = F.linear(x, w, bias if r == 0 else None)       # linear.py:186 — rank 0 alone adds the bias
        y = part.clone() if y is None else y.add_(part)       # linear.py:188-189 — dist.all_reduce(y), SUM
    return y


def rowparallel_exact(xs, ws, bias=None) -> torch.Tensor:
    y = sum(x.double() @ w.double().T for x, w in zip(xs, ws))
    if bias is not None:
        y = y + bias.double()
    return y


def _ref_root() -> str | None:
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for cand in ("/root/reference", os.path.join(here, "baseline", "_ref")):
        if os.path.isdir(os.path.join(cand, "nanovllm")):
            return cand
    return None


def _world_worker(rank: int, world: int, port: int, x, weight, bias, out_q):
    import importlib.util
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # load linear.py from its source file: importing the nanovllm package would pull flash-attn/triton
        spec = importlib.util.spec_from_file_location("ref_linear", os.path.join(_ref_root(), "nanovllm", "layers", "linear.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        n, in_features = x.shape
        m = mod.RowParallelLinear(in_features, weight.shape[0], bias=bias is not None).to(torch.bfloat16)
        m.weight.weight_loader(m.weight, weight)               # the reference's own sharding of W
        if bias is not None:
            m.bias.data.copy_(bias)
        xs = x.chunk(world, dim=1)[rank].contiguous()
        with torch.no_grad():
            y = m(xs)
        out_q.put((rank, y.view(torch.int16).numpy()))
    finally:
        dist.destroy_process_group()


def run_reference_world(x: torch.Tensor, weight: torch.Tensor, bias, world: int, port: int = 29731):
    """Outputs of the reference RowParallelLinear on every rank (list of bf16 tensors); needs the reference tree."""
    import torch.multiprocessing as mp
    if _ref_root() is None:
        raise FileNotFoundError("reference tree not found (/root/reference or baseline/_ref)")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_world_worker, args=(r, world, port, x, weight, bias, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    return [torch.from_numpy(got[r].copy()).view(torch.bfloat16) for r in range(world)]


if __name__ == "__main__":
    sys.exit(0)
