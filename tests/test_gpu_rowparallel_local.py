"""GPU, ONE device: the GEMM half of the fused RowParallelLinear kernel (csrc/rowparallel_allreduce.cu) through its
exchange mode 2 — the rank's own fp32 partial tile, k splits summed inside the thread-block cluster through distributed
shared memory, rounded to bf16 once and written to y with no cross-rank traffic — against `x_r @ W_r^T` in fp64.

The cross-rank exchange needs >= 2 GPUs (tests/test_gpu_rowparallel.py, skipped on a 1-GPU box); this file keeps the
tensor-core tiles, the cp.async ring, the cluster reduction and the epilogue of EVERY instantiation the planner can pick
(feature tile 32|64 x staged row 128|256 B x 1|2|4|8 token tiles) under test on any box.
reference: nanovllm/layers/linear.py:185-190 (`F.linear(x, self.weight, bias)` before the all-reduce).
Bar: |y - exact| <= 2^-8 * max|exact| (fp32 accumulate, a single bf16 rounding)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
ULP = 2.0 ** -8
DEV = "cuda:0"


@pytest.fixture(scope="module")
def local_comm(lib):
    """A 'world of 2' whose two inboxes are the same local buffer: mode 2 never touches them."""
    from nano_vllm_learn_b200 import _lib, ops
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    torch.cuda.set_device(0)
    need = 0
    for ks in (0, 1, 2, 4, 8):
        ops.rowparallel_set_plan(0, ks, 2)
        for shape in [(1024, 256), (960, 256), (2048, 1024), (512, 128)]:
            try:
                need = max(need, TPComm.required_bytes(2, [shape]))
            except ValueError:      # this forced k split does not divide the shape
                assert ks > 1
    ops.rowparallel_set_plan(0, 0, -1)
    p = ctypes.c_void_p()
    _lib.check(lib.nvllm_tp_comm_alloc(need, ctypes.byref(p), None), "tp_comm_alloc")
    comm = TPComm(0, 2, need, p.value, [p.value, p.value], "local")
    yield comm
    ops.rowparallel_set_plan(0, 0, -1)
    torch.cuda.synchronize()
    lib.nvllm_tp_comm_free(p)


def _case(n, k, h, bias, seed, x_pad=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, k + x_pad, generator=g).to(torch.bfloat16)[:, :k]      # row-strided when x_pad > 0
    w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
    b = torch.randn(h, generator=g).to(torch.bfloat16) if bias else None
    exact = x.double() @ w.double().t()
    if b is not None:
        exact = exact + b.double()
    return x, w, b, exact


def _check(ops, comm, n, k, h, bias, seed, x_pad=0):
    x, w, b, exact = _case(n, k, h, bias, seed, x_pad)
    xd = x.to(DEV) if x_pad == 0 else torch.randn(n, k + x_pad).to(torch.bfloat16).to(DEV)
    if x_pad:
        xd[:, :k] = x.to(DEV)
        xd = xd[:, :k]
    y = ops.rowparallel_linear_allreduce(xd, w.to(DEV), b.to(DEV) if b is not None else None, comm)
    torch.cuda.synchronize()
    err = (y.double().cpu() - exact).abs().max().item()
    assert err <= ULP * exact.abs().max().item() + 1e-6, (n, k, h, err)
    return y


@pytest.mark.parametrize("tile", [32, 64])
@pytest.mark.parametrize("n", [1, 16, 17, 32, 33, 64, 100, 128])          # token tiles 1 / 2 / 4 / 8, ragged and full
@pytest.mark.parametrize("k", [1024, 960])                                  # 256-byte staged rows where they fit / 128-byte
def test_gemm_half_every_instantiation(local_comm, tile, n, k):
    from nano_vllm_learn_b200 import ops
    ops.rowparallel_set_plan(tile, 0, 2)
    assert ops.rowparallel_plan(n, k, 256, 2)[0] == tile
    _check(ops, local_comm, n, k, 256, bias=False, seed=n * 7 + tile)


@pytest.mark.parametrize("ksplit", [1, 2, 4, 8])
def test_gemm_half_cluster_ksplit_and_bias(local_comm, ksplit):
    """k splits of a tile = one thread-block cluster summed through DSMEM; bias added once after the sum."""
    from nano_vllm_learn_b200 import ops
    ops.rowparallel_set_plan(0, ksplit, 2)
    _check(ops, local_comm, 33, 2048, 1024, bias=True, seed=90 + ksplit)
    _check(ops, local_comm, 5, 512, 128, bias=True, seed=190 + ksplit)


def test_gemm_half_strided_rows_multiple_token_blocks_and_determinism(local_comm):
    from nano_vllm_learn_b200 import ops
    ops.rowparallel_set_plan(0, 0, 2)
    _check(ops, local_comm, 200, 1024, 256, bias=False, seed=5, x_pad=64)    # 2 token blocks, x row stride > k
    a = _check(ops, local_comm, 48, 2048, 1024, bias=True, seed=6)
    b = _check(ops, local_comm, 48, 2048, 1024, bias=True, seed=6)
    assert torch.equal(a.view(torch.int16), b.view(torch.int16)), "same inputs, same launch plan -> same bits"
    from nano_vllm_learn_b200.rowparallel_cuda import poll_error
    assert poll_error() == 0


@pytest.mark.parametrize("n", [33, 64, 65, 100, 128, 200, 256, 512])      # 64- and 128-token UMMA N, 1-4 token blocks
@pytest.mark.parametrize("k,h", [(1024, 256), (960, 256), (2048, 1024)])
def test_gemm_half_tcgen05_tile(local_comm, n, k, h):
    """The swap-AB tcgen05 tile (feature tile 128: A = 128 output features of W, B = 64|128 token rows, accumulator in
    TMEM, operands through the same cp.async ring in the 128-byte-swizzled layout) against fp64, same bar."""
    from nano_vllm_learn_b200 import ops
    ops.rowparallel_set_plan(128, 0, 2)
    assert ops.rowparallel_plan(n, k, h, 2)[0] == 128
    y = _check(ops, local_comm, n, k, h, bias=(n % 2 == 0), seed=n * 3 + k)
    # and bit-level agreement class with the mma.sync tile: both accumulate in fp32 and round once
    ops.rowparallel_set_plan(64, 0, 2)
    y2 = _check(ops, local_comm, n, k, h, bias=(n % 2 == 0), seed=n * 3 + k)
    assert (y.float() - y2.float()).abs().max().item() <= 2 * ULP * y2.float().abs().max().item()


@pytest.mark.parametrize("ksplit", [1, 2, 4, 8])
def test_gemm_half_tcgen05_cluster_ksplit(local_comm, ksplit):
    from nano_vllm_learn_b200 import ops
    ops.rowparallel_set_plan(128, ksplit, 2)
    _check(ops, local_comm, 96, 2048, 1024, bias=True, seed=400 + ksplit)
    if ksplit <= 4:   # 2 token blocks x 8 feature tiles x 8-CTA clusters exceed the co-resident cluster capacity
        _check(ops, local_comm, 130, 2048, 1024, bias=False, seed=500 + ksplit, x_pad=64)
    ops.rowparallel_set_plan(0, 0, 2)
    from nano_vllm_learn_b200.rowparallel_cuda import poll_error
    assert poll_error() == 0
