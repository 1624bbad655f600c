"""GPU (needs >= 2 devices; skipped on a 1-GPU box): fused RowParallelLinear GEMM + all-reduce (SURVEY §8f rank 2,
csrc/rowparallel_allreduce.cu) against the oracle.

Bar (floating point): every rank's y is BIT-IDENTICAL to every other rank's (tensor parallelism relies on it), and
within one bf16 rounding of the exact contraction: |y - exact| <= 2^-8 * max|exact| (fp32 accumulate, single
rounding) — tighter than the reference's own result, which rounds every rank's partial to bf16 first.  Also checked
against the reference-style result `rowparallel_ref` (<= 3 bf16 ulp of the scale) and the frozen reference fixture.
Two transports: one process driving both GPUs over direct peer access, and one process per GPU over CUDA IPC with the
reference's NCCL all-reduce timed beside it."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import cases
from oracle import rowparallel_oracle as rpo

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ULP = 2.0 ** -8
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "rowparallel.npz"))


def _need_two():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs with peer access")


@pytest.fixture(scope="module")
def comms(lib):
    _need_two()
    from nano_vllm_learn_b200 import ops
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    ops.rowparallel_set_plan(timeout_ms=5000)
    need = 0
    shapes = [(c[2] // c[4], c[3]) for c in cases.ROWPARALLEL_CASES.values() if c[4] == 2] + [(2048, 4096), (1024, 1024)]
    for ks in (0, 1, 2, 4, 8):
        ops.rowparallel_set_plan(0, ks, -1)
        for k, h in shapes:
            need = max(need, int(lib.nvllm_rowparallel_comm_bytes(2, 512, k, h)))
    ops.rowparallel_set_plan(0, 0, -1)
    return TPComm.single_process([0, 1], need)


def _run_world(comms, xs, ws, bias, outs=None):
    """Launch rank 0 and rank 1 back to back WITHOUT synchronising in between (the kernels wait for each other)."""
    from nano_vllm_learn_b200 import ops
    if outs is None:   # allocate before the first launch: no cudaMalloc while a peer's kernel is spinning
        outs = [torch.empty((xs[r].shape[0], ws[r].shape[0]), dtype=torch.bfloat16, device=f"cuda:{r}")
                for r in range(len(comms))]
    ys = []
    for r, comm in enumerate(comms):
        with torch.cuda.device(r):
            ys.append(ops.rowparallel_linear_allreduce(xs[r], ws[r], bias[r] if bias else None, comm, out=outs[r]))
    for r in range(len(comms)):
        torch.cuda.synchronize(r)
    return ys


def _to_dev(c):
    xs, ws = rpo.shard_inputs(c["x"], c["weight"], c["world"])
    xd = [x.to(f"cuda:{r}") for r, x in enumerate(xs)]
    wd = [w.to(f"cuda:{r}") for r, w in enumerate(ws)]
    bd = [c["bias"].to(f"cuda:{r}") for r in range(c["world"])] if c["bias"] is not None else None
    return xs, ws, xd, wd, bd


def _check(ys, xs, ws, bias, comms, what):
    for comm in comms:
        _, err = comm.status()
        assert err == 0, f"{what}: rank {comm.rank} timed out waiting for tile {err - 1}"
    y0 = ys[0].cpu()
    for r in range(1, len(ys)):
        assert torch.equal(y0.view(torch.int16), ys[r].cpu().view(torch.int16)), f"{what}: ranks disagree"
    exact = rpo.rowparallel_exact(xs, ws, bias)
    scale = exact.abs().max().item()
    err = (y0.double() - exact).abs().max().item()
    assert err <= ULP * scale, f"{what}: max|err| {err:.3e} vs bound {ULP * scale:.3e}"
    ref = rpo.rowparallel_ref(xs, ws, bias)
    assert (y0.float() - ref.float()).abs().max().item() <= 3 * ULP * scale, what
    return y0


@pytest.mark.parametrize("name", [n for n, c in cases.ROWPARALLEL_CASES.items() if c[4] == 2])
@pytest.mark.parametrize("mode", [0, 1, 3])
def test_cases_vs_oracle_and_fixture(comms, name, mode):
    from nano_vllm_learn_b200 import ops
    c = cases.make_rowparallel_case(name)
    xs, ws, xd, wd, bd = _to_dev(c)
    for bh, ks in ((0, 0), (32, 1), (64, 2), (32, 4)):
        if ks and (c["k"] // 2) % (ks * 64):
            continue
        ops.rowparallel_set_plan(bh, ks, mode)
        try:
            ys = _run_world(comms, xd, wd, bd)
        finally:
            ops.rowparallel_set_plan(0, 0, -1)
        y = _check(ys, xs, ws, c["bias"], comms, f"{name} mode={mode} bh={bh} ks={ks}")
        want = cases.from_bits(GOLD[f"{name}_y"])
        scale = want.float().abs().max().item()
        assert (y.float() - want.float()).abs().max().item() <= 3 * ULP * scale


@pytest.mark.parametrize("n", [1, 15, 16, 17, 64, 128, 129, 300, 512])
def test_token_counts_and_strided_rows(comms, n):
    """ragged token counts across the 16-row MMA tiles and 128-token CTA blocks; x is a strided view (the o_proj input is
    contiguous, the general contract is row-strided), y is written into a wider buffer."""
    g = torch.Generator().manual_seed(100 + n)
    k, h = 2048, 1024
    x = torch.randn(n, k, generator=g).to(torch.bfloat16)
    w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
    xs, ws = rpo.shard_inputs(x, w, 2)
    wide = [torch.randn(n, k // 2 + 64, generator=g).to(torch.bfloat16).to(f"cuda:{r}") for r in range(2)]
    for r in range(2):
        wide[r][:, 8:8 + k // 2] = xs[r].to(f"cuda:{r}")
    xd = [wide[r][:, 8:8 + k // 2] for r in range(2)]
    wd = [ws[r].to(f"cuda:{r}") for r in range(2)]
    outs_full = [torch.full((n, h + 16), 7.0, dtype=torch.bfloat16, device=f"cuda:{r}") for r in range(2)]
    outs = [o[:, 8:8 + h] for o in outs_full]
    ys = _run_world(comms, xd, wd, None, outs)
    _check(ys, xs, ws, None, comms, f"n={n}")
    for o in outs_full:   # nothing outside the [n, h] window was touched
        assert (o[:, :8] == 7).all() and (o[:, 8 + h:] == 7).all()


@pytest.mark.parametrize("mode", [0, 1, 3])
@pytest.mark.parametrize("n", [33, 64, 128, 200, 512])
def test_tcgen05_tile_through_every_exchange(comms, n, mode):
    """The swap-AB tcgen05 tile (feature tile 128) feeds the same staged-tile exchanges as the mma.sync tile: one-shot
    flagged (0), two-shot (1) and flag-in-data (3), ragged token counts over 1-4 token blocks, with a bias."""
    from nano_vllm_learn_b200 import ops
    g = torch.Generator().manual_seed(700 + n)
    k, h = 2048, 1024
    x = torch.randn(n, k, generator=g).to(torch.bfloat16)
    w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
    b = torch.randn(h, generator=g).to(torch.bfloat16)
    xs, ws = rpo.shard_inputs(x, w, 2)
    xd = [xs[r].to(f"cuda:{r}") for r in range(2)]
    wd = [ws[r].to(f"cuda:{r}") for r in range(2)]
    bd = [b.to(f"cuda:{r}") for r in range(2)]
    ops.rowparallel_set_plan(128, 0, mode)
    try:
        assert ops.rowparallel_plan(n, k // 2, h, 2)[0] == 128
        ys = _run_world(comms, xd, wd, bd)
    finally:
        ops.rowparallel_set_plan(0, 0, -1)
    _check(ys, xs, ws, b, comms, f"tcgen05 n={n} mode={mode}")


def test_back_to_back_calls_alternate_buffers(comms):
    """40 calls with changing inputs, shapes and modes and no host synchronisation: the device-side epoch, the
    double-buffered inboxes and stale flags from other tile geometries."""
    from nano_vllm_learn_b200 import ops
    g = torch.Generator().manual_seed(7)
    plans = []
    for i in range(40):
        n = [16, 3, 64, 130][i % 4]
        k, h = [(2048, 1024), (1024, 1024), (3072, 1024)][i % 3]
        x = torch.randn(n, k, generator=g).to(torch.bfloat16)
        w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
        xs, ws = rpo.shard_inputs(x, w, 2)
        plans.append((xs, ws, [t.to(f"cuda:{r}") for r, t in enumerate(xs)], [t.to(f"cuda:{r}") for r, t in enumerate(ws)],
                      [0, 1, 3][i % 3]))
    results = []
    for xs, ws, xd, wd, mode in plans:
        ops.rowparallel_set_plan(0, 0, mode)
        ys = []
        for r, comm in enumerate(comms):
            with torch.cuda.device(r):
                ys.append(ops.rowparallel_linear_allreduce(xd[r], wd[r], None, comm))
        results.append(ys)
    ops.rowparallel_set_plan(0, 0, -1)
    for r in range(2):
        torch.cuda.synchronize(r)
    e0 = comms[0].status()[0]
    assert e0 == comms[1].status()[0] and e0 >= 40
    for i, ((xs, ws, _, _, mode), ys) in enumerate(zip(plans, results)):
        _check(ys, xs, ws, None, comms, f"call {i} mode={mode}")


def test_cuda_graph_replay(comms):
    """the decode path runs under CUDA graphs (model_runner.py:316-370): capture 6 chained calls per rank, replay 5x."""
    from nano_vllm_learn_b200 import ops
    g = torch.Generator().manual_seed(11)
    n, k, h = 32, 2048, 1024
    w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
    x = torch.randn(n, k, generator=g).to(torch.bfloat16)
    xs, ws = rpo.shard_inputs(x, w, 2)
    xd = [t.to(f"cuda:{r}") for r, t in enumerate(xs)]
    wd = [t.to(f"cuda:{r}") for r, t in enumerate(ws)]
    outs = [[torch.zeros(n, h, dtype=torch.bfloat16, device=f"cuda:{r}") for _ in range(6)] for r in range(2)]
    _run_world(comms, xd, wd, None)                       # warm: attributes set outside capture
    graphs = []
    for r in range(2):
        with torch.cuda.device(r):
            # torch.cuda.graph's default capture stream is ONE process-wide stream on whichever device captured first;
            # entering it switches the current device, so rank 1's kernels would be recorded on cuda:0.  One capture
            # stream per device (the engine has one process per GPU and never sees this).
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=torch.cuda.Stream(device=r)):
                assert torch.cuda.current_device() == r
                for j in range(6):
                    ops.rowparallel_linear_allreduce(xd[r], wd[r], None, comms[r], out=outs[r][j])
            graphs.append(gr)
    for rep in range(5):
        xnew = torch.randn(n, k, generator=g).to(torch.bfloat16)
        xs, _ = rpo.shard_inputs(xnew, w, 2)
        for r in range(2):
            xd[r].copy_(xs[r])
        for r in range(2):
            torch.cuda.synchronize(r)
        for r in range(2):
            with torch.cuda.device(r):
                graphs[r].replay()
        for r in range(2):
            torch.cuda.synchronize(r)
        for j in range(6):
            _check([outs[0][j], outs[1][j]], xs, ws, None, comms, f"replay {rep} call {j}")


def test_one_process_per_gpu_over_cuda_ipc(lib):
    """the reference's process layout: torchrun, NCCL process group, CUDA IPC handles through all_gather_object; the
    worker compares against F.linear + dist.all_reduce (the reference's RowParallelLinear.forward) on the same box."""
    _need_two()
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([ROOT, os.path.join(ROOT, "tests"), os.environ.get("PYTHONPATH", "")]))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(29900 + os.getpid() % 90), os.path.join(ROOT, "tools", "bench_rowparallel.py"), "--check-only"]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    sys.stdout.write(p.stdout[-4000:])
    sys.stderr.write(p.stderr[-4000:])
    assert p.returncode == 0
    assert "ROWPARALLEL_IPC_OK" in p.stdout
