"""GPU, world sizes 2 / 4 / 8 (each skipped unless the box has that many devices with peer access): the fused
RowParallelLinear GEMM + all-reduce (csrc/rowparallel_allreduce.cu) against the oracle, one process driving every
device over direct peer access, plus the one-process-per-GPU CUDA-IPC layout of the engine.

Same bar as tests/test_gpu_rowparallel.py: every rank's y BIT-IDENTICAL to every other rank's, and within one bf16
rounding of the exact contraction (|y - exact| <= 2^-8 max|exact|).  reference: nanovllm/layers/linear.py:185-190."""
import os
import subprocess
import sys

import pytest
import torch

from oracle import rowparallel_oracle as rpo

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ULP = 2.0 ** -8
SHAPES = [(2048, 1024), (3072, 1024), (4096, 4096), (12288, 4096), (1024, 256)]   # (k, h): 0.6B and 8B o_proj / down_proj
_COMMS = {}


def _comms(lib, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs with peer access")
    if world not in _COMMS:
        from nano_vllm_learn_b200 import ops
        from nano_vllm_learn_b200.rowparallel_cuda import TPComm
        ops.rowparallel_set_plan(timeout_ms=5000)
        need = 0
        for ks in (0, 1, 2, 4, 8):
            ops.rowparallel_set_plan(0, ks, -1)
            for k, h in SHAPES:
                if k % world == 0:
                    need = max(need, int(lib.nvllm_rowparallel_comm_bytes(world, 512, k // world, h)))
        ops.rowparallel_set_plan(0, 0, -1)
        _COMMS[world] = TPComm.single_process(list(range(world)), need)
    return _COMMS[world]


def _launch(comms, xd, wd, bd, outs):
    from nano_vllm_learn_b200 import ops
    for r, comm in enumerate(comms):      # no synchronisation in between: the kernels wait for each other
        with torch.cuda.device(r):
            ops.rowparallel_linear_allreduce(xd[r], wd[r], bd[r] if bd else None, comm, out=outs[r])


def _sync(world):
    for r in range(world):
        torch.cuda.synchronize(r)


def _check(ys, xs, ws, bias, comms, what):
    for comm in comms:
        _, err = comm.status()
        assert err == 0, f"{what}: rank {comm.rank} timed out waiting for tile {err - 1}"
    y0 = ys[0].cpu()
    for r in range(1, len(ys)):
        assert torch.equal(y0.view(torch.int16), ys[r].cpu().view(torch.int16)), f"{what}: rank {r} differs from rank 0"
    exact = rpo.rowparallel_exact(xs, ws, bias)
    scale = exact.abs().max().item()
    err = (y0.double() - exact).abs().max().item()
    assert err <= ULP * scale, f"{what}: max|err| {err:.3e} vs bound {ULP * scale:.3e}"


def _case(n, k, h, world, seed, bias=False, amp=1.0):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(n, k, generator=g) * amp).to(torch.bfloat16)
    w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
    b = torch.randn(h, generator=g).to(torch.bfloat16) if bias else None
    xs, ws = rpo.shard_inputs(x, w, world)
    xd = [t.to(f"cuda:{r}") for r, t in enumerate(xs)]
    wd = [t.to(f"cuda:{r}") for r, t in enumerate(ws)]
    bd = [b.to(f"cuda:{r}") for r in range(world)] if bias else None
    return xs, ws, b, xd, wd, bd


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode", [-1, 0, 1, 3])      # -1: the planner's own choice
@pytest.mark.parametrize("n,k,h,bias", [(1, 2048, 1024, False), (16, 3072, 1024, False), (64, 4096, 4096, True),
                                        (130, 12288, 4096, False), (300, 1024, 256, True)])
def test_world_vs_oracle(lib, world, mode, n, k, h, bias):
    from nano_vllm_learn_b200 import ops
    comms = _comms(lib, world)
    xs, ws, b, xd, wd, bd = _case(n, k, h, world, seed=world * 1000 + n, bias=bias)
    outs = [torch.empty(n, h, dtype=torch.bfloat16, device=f"cuda:{r}") for r in range(world)]
    ops.rowparallel_set_plan(0, 0, mode)
    try:
        for _ in range(2):      # twice: both epoch parities of the inboxes
            _launch(comms, xd, wd, bd, outs)
        _sync(world)
    finally:
        ops.rowparallel_set_plan(0, 0, -1)
    _check(outs, xs, ws, b, comms, f"world={world} mode={mode} n={n} k={k} h={h}")


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ll_exchange_stress(lib, world):
    """Mode 3 (8-byte {payload, epoch} words, no fence, no flag): 240 back-to-back calls with NO host synchronisation,
    cycling over inputs whose partial sums differ in sign and by orders of magnitude, so a word whose epoch matched but
    whose payload was stale (a torn 8-byte half) would be off by far more than the bound in some call."""
    from nano_vllm_learn_b200 import ops
    comms = _comms(lib, world)
    k, h = 2048, 1024
    sets = []
    for i, (n, amp) in enumerate([(1, 1.0), (16, -64.0), (64, 1e-3), (7, 4096.0), (33, -1.0), (128, 0.25)]):
        xs, ws, _, xd, wd, _ = _case(n, k, h, world, seed=900 + i, amp=abs(amp))
        if amp < 0:
            xd = [-t for t in xd]
            xs = [-t for t in xs]
        sets.append((xs, ws, xd, wd, n))
    calls = 240
    outs = [[torch.empty(sets[c % len(sets)][4], h, dtype=torch.bfloat16, device=f"cuda:{r}") for r in range(world)]
            for c in range(calls)]
    _sync(world)
    e_before = comms[0].status()[0]
    ops.rowparallel_set_plan(0, 0, 3)
    try:
        for c in range(calls):
            xs, ws, xd, wd, n = sets[c % len(sets)]
            _launch(comms, xd, wd, None, outs[c])
        _sync(world)
    finally:
        ops.rowparallel_set_plan(0, 0, -1)
    assert comms[0].status()[0] - e_before == calls
    for c in range(calls):
        xs, ws, _, _, _ = sets[c % len(sets)]
        _check(outs[c], xs, ws, None, comms, f"world={world} call {c}")


@pytest.mark.parametrize("world", [4, 8])
def test_one_process_per_gpu_over_cuda_ipc(lib, world):
    """the engine's process layout at 4 and 8 ranks (tests/test_gpu_rowparallel.py covers 2)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([ROOT, os.path.join(ROOT, "tests"), os.environ.get("PYTHONPATH", "")]))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(29700 + os.getpid() % 90 + world), os.path.join(ROOT, "tools", "bench_rowparallel.py"),
           "--check-only"]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    sys.stdout.write(p.stdout[-4000:])
    sys.stderr.write(p.stderr[-4000:])
    assert p.returncode == 0
    assert "ROWPARALLEL_IPC_OK" in p.stdout
