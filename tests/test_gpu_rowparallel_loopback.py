"""GPU, ONE device: the EXCHANGE of the fused RowParallelLinear GEMM + all-reduce (csrc/rowparallel_allreduce.cu,
modes 0 = two-shot, 1 = flagged one-shot, 3 = flag-in-data) with every tensor-parallel rank running on the same GPU --
one exchange buffer and one stream per rank, the kernels of all ranks co-resident and waiting for each other exactly as
they do across NVLink.  VERDICT r2 (missing #2): the multi-GPU exchange tests are skipped on a 1-GPU box, so the
driver's test run never executed the exchange; the sampler and the embedding already had such a loop-back test.

Plans are forced small (64-feature tiles, no k split) so that world x tiles CTAs fit the SMs at once: a rank whose CTAs
could not become resident while its peers spin would only time out (5 s bound, then the test fails; nothing hangs).

Bar as in tests/test_gpu_rowparallel_worlds.py: every rank's y bit-identical to rank 0's, and within one bf16 rounding
of the exact contraction.  reference: nanovllm/layers/linear.py:185-190."""
import pytest
import torch

from oracle import rowparallel_oracle as rpo

pytestmark = pytest.mark.gpu
ULP = 2.0 ** -8
TILE = 64
_STATE = {}


def _world(lib, world):
    if world not in _STATE:
        from nano_vllm_learn_b200 import ops
        from nano_vllm_learn_b200.rowparallel_cuda import TPComm
        ops.rowparallel_set_plan(TILE, 1, -1, timeout_ms=5000)
        need = max(int(lib.nvllm_rowparallel_comm_bytes(world, 128, k // world, h)) for k, h in [(2048, 1024), (1024, 256)])
        ops.rowparallel_set_plan(0, 0, -1)
        comms = TPComm.single_process([0] * world, need)
        streams = [torch.cuda.Stream(device="cuda:0") for _ in range(world)]
        _STATE[world] = (comms, streams)
    return _STATE[world]


def _case(n, k, h, world, seed, bias):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, k, generator=g).to(torch.bfloat16)
    w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
    b = torch.randn(h, generator=g).to(torch.bfloat16) if bias else None
    xs, ws = rpo.shard_inputs(x, w, world)
    return xs, ws, b, [t.cuda() for t in xs], [t.cuda() for t in ws], (b.cuda() if bias else None)


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("mode", [0, 1, 3, -1])      # -1: the planner's own exchange for this size
@pytest.mark.parametrize("n,k,h,bias", [(1, 2048, 1024, False), (16, 1024, 256, True), (64, 2048, 1024, True)])
def test_exchange_loopback_vs_oracle(lib, world, mode, n, k, h, bias):
    from nano_vllm_learn_b200 import ops, rowparallel_cuda
    assert world * (h // TILE) <= 148, "all ranks' CTAs must be resident together"
    comms, streams = _world(lib, world)
    xs, ws, b, xd, wd, bd = _case(n, k, h, world, seed=world * 100 + n, bias=bias)
    outs = [torch.empty(n, h, dtype=torch.bfloat16, device="cuda") for _ in range(world)]
    torch.cuda.synchronize()
    rowparallel_cuda.poll_error(clear=True)
    ops.rowparallel_set_plan(TILE, 1, mode, timeout_ms=5000)
    try:
        for it in range(3):      # both epoch parities of the inboxes, and a changing launch order
            order = list(range(world)) if it % 2 == 0 else list(reversed(range(world)))
            for r in order:
                with torch.cuda.stream(streams[r]):
                    ops.rowparallel_linear_allreduce(xd[r], wd[r], bd, comms[r], out=outs[r])
        torch.cuda.synchronize()
    finally:
        ops.rowparallel_set_plan(0, 0, -1)
    what = f"world={world} mode={mode} n={n} k={k} h={h}"
    assert rowparallel_cuda.poll_error() == 0, what
    for comm in comms:
        assert comm.status()[1] == 0, f"{what}: rank {comm.rank} timed out"
    y0 = outs[0].cpu()
    for r in range(1, world):
        assert torch.equal(y0.view(torch.int16), outs[r].cpu().view(torch.int16)), f"{what}: rank {r} differs from rank 0"
    exact = rpo.rowparallel_exact(xs, ws, b)
    err = (y0.double() - exact).abs().max().item()
    assert err <= ULP * exact.abs().max().item(), what
