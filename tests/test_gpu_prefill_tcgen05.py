"""GPU: the tcgen05/TMEM prefill kernel (variant 1) against the oracle, the reference fixtures and the
mma.sync tile kernel on the same inputs."""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import paged_attention_oracle as orc

pytestmark = pytest.mark.gpu
TOL = 2e-2
DEV = "cuda"
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture()
def ops(lib):
    from nano_vllm_learn_b200 import ops as _ops
    _ops.set_prefill_variant(1)
    yield _ops
    _ops.set_prefill_variant(-1)


def _cuda(c):
    return {k: (v.to(DEV) if isinstance(v, torch.Tensor) else v) for k, v in c.items()}


def _maxerr(a, b):
    return (a.float().cpu() - b.float().cpu()).abs().max().item()


@pytest.mark.parametrize("name", list(cases.PREFILL_CASES))
def test_tcgen05_prefill_vs_oracle_and_fixture(ops, name):
    c = cases.make_prefill_case(name)
    want = orc.varlen_prefill_ref(c["q"], c["k"], c["v"], c["cu_seqlens_q"], c["cu_seqlens_k"], c["scale"])
    g = _cuda(c)
    got = ops.flash_attn_varlen_func(g["q"], g["k"], g["v"], c["max_seqlen_q"], g["cu_seqlens_q"], c["max_seqlen_k"],
                                     g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True)
    assert ops.prefill_last_path() == 1, "the tcgen05 kernel did not serve this call"
    assert _maxerr(got, want) <= TOL
    gold = cases.from_bits(np.load(os.path.join(GOLD, f"prefill_{name}.npz"))["out"])
    assert _maxerr(got, gold) <= TOL


@pytest.mark.parametrize("name", list(cases.PREFIX_CASES))
def test_tcgen05_prefix_prefill_vs_oracle(ops, name):
    c = cases.make_prefix_case(name)
    want = orc.varlen_prefill_ref(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], c["scale"],
                                  c["block_tables"])
    g = _cuda(c)
    got = ops.flash_attn_varlen_func(g["q"], g["k_cache"], g["v_cache"], c["max_seqlen_q"], g["cu_seqlens_q"],
                                     c["max_seqlen_k"], g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True,
                                     block_table=g["block_tables"])
    assert ops.prefill_last_path() == 1, "the tcgen05 kernel did not serve this call"
    assert _maxerr(got, want) <= TOL


def test_tcgen05_full_size_vs_mma_kernel(ops):
    """bench.py's prefill step (16 x 1024, 16q/8kv) with strided k/v views: both kernels agree; one
    sequence against the CPU oracle."""
    bsz, L, hq, hkv = 16, 1024, 16, 8
    g = torch.Generator(device=DEV).manual_seed(9)
    total = bsz * L
    q = torch.randn(total, hq, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    fused = torch.randn(total, 2 * hkv * 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    k = fused[:, :hkv * 128].view(total, hkv, 128)
    v = fused[:, hkv * 128:].view(total, hkv, 128)
    cu = torch.arange(0, total + 1, L, dtype=torch.int32, device=DEV)
    scale = 128 ** -0.5
    n1, n0 = ops.prefill_path_count(1), ops.prefill_path_count(0)
    out = ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=scale, causal=True)
    assert ops.prefill_last_path() == 1 and ops.prefill_path_count(1) == n1 + 1
    ops.set_prefill_variant(0)
    ref = ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=scale, causal=True)
    assert ops.prefill_last_path() == 0 and ops.prefill_path_count(0) == n0 + 1
    ops.set_prefill_variant(1)
    assert _maxerr(out, ref) <= 8e-3
    i = 11
    want = orc.varlen_prefill_ref(q[i * L:(i + 1) * L].cpu(), k[i * L:(i + 1) * L].cpu(), v[i * L:(i + 1) * L].cpu(),
                                  torch.tensor([0, L], dtype=torch.int32), torch.tensor([0, L], dtype=torch.int32), scale)
    assert _maxerr(out[i * L:(i + 1) * L], want) <= TOL


def test_tcgen05_ragged_batch_vs_mma_kernel(ops):
    lens = [1, 127, 128, 129, 255, 257, 700, 33]
    hq, hkv = 8, 2
    g = torch.Generator(device=DEV).manual_seed(4)
    total = sum(lens)
    q = torch.randn(total, hq, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    k = torch.randn(total, hkv, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    v = torch.randn(total, hkv, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    cu = torch.tensor([0] + list(np.cumsum(lens)), dtype=torch.int32, device=DEV)
    out = ops.flash_attn_varlen_func(q, k, v, max(lens), cu, max(lens), cu, softmax_scale=128 ** -0.5)
    want = orc.varlen_prefill_ref(q.cpu(), k.cpu(), v.cpu(), cu.cpu(), cu.cpu(), 128 ** -0.5)
    assert _maxerr(out, want) <= TOL


def test_default_variant_is_tcgen05_and_no_timeout_latched(lib):
    """The library default (variant -1) must run the tcgen05 kernel on these shapes, and no mbarrier wait of
    any prefill launch in this process may have timed out (the kernel latches instead of trapping)."""
    from nano_vllm_learn_b200 import ops as _ops
    _ops.set_prefill_variant(-1)
    c = cases.make_prefill_case(next(iter(cases.PREFILL_CASES)))
    g = _cuda(c)
    _ops.flash_attn_varlen_func(g["q"], g["k"], g["v"], c["max_seqlen_q"], g["cu_seqlens_q"], c["max_seqlen_k"],
                                g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True)
    assert _ops.prefill_last_path() == 1
    assert _ops.PREFILL_KERNEL_NAMES[_ops.prefill_last_path()] == "prefill_tcgen05_kernel"
    torch.cuda.synchronize()
    assert _ops.prefill_error() == 0


def test_mbarrier_timeout_surfaces_on_the_product_path_and_does_not_poison_later_launches(ops, lib):
    """a timed-out mbarrier wait used to set a sticky latch nobody on the engine path read, after which
    every later wait of the unit returned early.  Now the waiter also writes a mapped host word; the next
    flash_attn_varlen_func call reads it, clears both words and raises; the call after that is clean."""
    import ctypes
    from nano_vllm_learn_b200 import _lib
    c = cases.make_prefill_case("p_small")
    want = orc.varlen_prefill_ref(c["q"], c["k"], c["v"], c["cu_seqlens_q"], c["cu_seqlens_k"], c["scale"])
    g = _cuda(c)

    def run():
        return ops.flash_attn_varlen_func(g["q"], g["k"], g["v"], c["max_seqlen_q"], g["cu_seqlens_q"],
                                          c["max_seqlen_k"], g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True)
    assert _maxerr(run(), want) <= TOL                 # also arms the host word on this device
    assert ops.prefill_poll_error() == 0
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.nvllm_tc_latch_selftest(7, stream) == 1
    assert lib.nvllm_tc_latch_selftest(0, stream) == 0  # a 1 ms wait on a barrier nobody completes, in the prefill unit
    torch.cuda.synchronize()
    code = ops.prefill_poll_error()
    assert code & 0x80000000, hex(code)
    with pytest.raises(_lib.NvllmError, match="mbarrier wait timed out"):
        run()
    assert ops.prefill_poll_error() == 0               # cleared by the raise: host word ...
    assert _maxerr(run(), want) <= TOL                 # ... and device latch: this launch waits for real again
    assert ops.prefill_last_path() == 1
    torch.cuda.synchronize()
    assert ops.prefill_error() == 0


def test_rowparallel_tile_latch_reports_through_poll_error(lib):
    """The RowParallelLinear unit has its own copy of the latch; it reports through nvllm_rowparallel_poll_error."""
    import ctypes
    from nano_vllm_learn_b200 import _lib, rowparallel_cuda as rp
    assert rp.poll_error() == 0
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.nvllm_tc_latch_selftest(1, stream) == 0
    torch.cuda.synchronize()
    assert rp.poll_error() & 0x80000000
    from nano_vllm_learn_b200 import ops as _ops
    assert _ops.prefill_poll_error() == 0               # the prefill unit's latch is a different word
    with pytest.raises(_lib.NvllmError, match="tcgen05 tile"):
        rp.raise_on_timeout()
    assert rp.poll_error() == 0
    assert lib.nvllm_tc_latch_selftest(1, stream) == 0  # the device latch was cleared too: a second timeout latches again
    torch.cuda.synchronize()
    assert rp.poll_error() & 0x80000000
    assert rp.poll_error(clear=True) & 0x80000000 and rp.poll_error() == 0
