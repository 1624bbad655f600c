"""nvllm_paged_decode_noalloc: split-KV decode whose scratch lives inside the library image.

Some integrations of this library cannot get device memory at all: a serving loop bound over plain FFI has no torch
allocator, and once the engine has given the rest of HBM to the KV cache a raw cudaMalloc may fail (and would stall
the device if it did not).  For them the library carries a fixed 2 MiB scratch array and an entry point that takes
the same arguments as nvllm_paged_decode minus the workspace.  This file checks, through the C ABI only, that the
entry point returns what nvllm_paged_decode returns on the same inputs -- twice in a row with different shapes, since
every call reuses the one array and finds the previous call's bytes in it -- and that it refuses shapes beyond the
array instead of launching.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide the
parity tests.  It passes on its own in a fresh process (first use of the array in the process) and after the rest of
the suite.
"""
import ctypes
import re
import os

import pytest
import torch

import cases

TOL = 2e-2          # the decode bar of tests/test_gpu_parity.py
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _scratch_limit():
    src = open(os.path.join(ROOT, "include", "nvllm_attn.h")).read()
    m = re.search(r"#define\s+NVLLM_PAGED_DECODE_NOALLOC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", src)
    return int(m.group(1)) << int(m.group(2))


def test_noalloc_refuses_before_touching_the_gpu(lib):
    """Argument and size errors come back without any CUDA call, so this runs on a machine without a GPU."""
    limit = _scratch_limit()
    assert limit == 2 << 20

    def call(bs=4, hq=16, hkv=8, d=128, bsz=256, splits=4, out=16):
        return lib.nvllm_paged_decode_noalloc(out, 16, hq * d, 16, 16, bsz * hkv * d, hkv * d, 16, 4, 16, bs, hq, hkv, d,
                                              bsz, 0.1, splits, None)
    assert call(bs=0) == 0
    assert call(splits=0) == 1 and call(splits=-4) == 1       # the balanced grid is not offered here
    assert call(out=None) == 1
    assert call(d=64) == 2 and call(hq=24, hkv=1) == 2 and call(splits=129) == 2
    # the largest geometry that fits and the first that does not: 64 KiB of counters + 516 B per partial row
    rows = (limit - 65536) // 516
    assert rows == 3937
    for bs, hq, s in [(123, 16, 2), (30, 16, 8), (61, 8, 8)]:       # not called here: a call that fits goes to the GPU
        assert lib.nvllm_paged_decode_workspace_bytes(bs, hq, 128, s) <= limit
    too_big = [(124, 16, 2), (31, 16, 8), (64, 16, 4), (512, 16, 2)]
    for bs, hq, s in too_big:
        assert lib.nvllm_paged_decode_workspace_bytes(bs, hq, 128, s) > limit
        assert call(bs=bs, hq=hq, hkv=hq // 2, splits=s) == 5, (bs, hq, s)
    assert lib.nvllm_status_string(5) == b"workspace missing or too small"


def _run_both(lib, c, splits):
    """One call of each entry point on the same device tensors, on torch's current stream."""
    dev = "cuda"
    q, kc, vc = c["q"].to(dev), c["k_cache"].to(dev), c["v_cache"].to(dev)
    lens, tables = c["context_lens"].to(dev), c["block_tables"].to(dev)
    bs, hq, d = q.shape
    hkv, bsz = c["hkv"], c["block_size"]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    head = (q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), kc.stride(0), kc.stride(1), tables.data_ptr(),
            tables.stride(0), lens.data_ptr(), bs, hq, hkv, d, bsz, float(c["scale"]), splits)

    # existing entry point: caller-owned workspace, zero-filled once
    ws_bytes = lib.nvllm_paged_decode_workspace_bytes(bs, hq, d, splits)
    ws = torch.zeros(max(ws_bytes, 16), dtype=torch.uint8, device=dev)
    want = torch.full((bs, hq, d), float("nan"), dtype=torch.bfloat16, device=dev)
    assert lib.nvllm_paged_decode(want.data_ptr(), *head, ws.data_ptr() if ws_bytes else None, ws_bytes, stream) == 0

    # new entry point: no workspace argument; everything it needs was allocated above, by the test
    got = torch.full((bs, hq, d), float("nan"), dtype=torch.bfloat16, device=dev)
    launches = lib.nvllm_launch_count()
    assert lib.nvllm_paged_decode_noalloc(got.data_ptr(), *head, stream) == 0
    assert lib.nvllm_launch_count() == launches + 1
    torch.cuda.synchronize()
    return got.float().cpu(), want.float().cpu()


@pytest.mark.gpu
def test_noalloc_matches_paged_decode_two_calls_in_a_row(lib):
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device")
    # Two different shapes back to back: the second call finds the first one's counters and partials in the array.
    # Then the first shape again with another split count, and a single split (which uses no scratch at all).
    plan = [("ragged_g4", 4), ("ragged_g2_tail", 3), ("ragged_g4", 8), ("g8", 16), ("c1_bs1_ctx512", 2),
            ("ragged_g2_tail", 1)]
    for name, splits in plan:
        c = cases.make_decode_case(name)
        got, want = _run_both(lib, c, splits)
        assert not torch.isnan(got).any(), (name, splits)
        err = (got - want).abs().max().item()
        assert err <= TOL, (name, splits, err)
    # the same inputs twice in a row give the same bits (nothing left behind by call 1 leaks into call 2)
    c = cases.make_decode_case("ragged_g4")
    first, _ = _run_both(lib, c, 4)
    second, want = _run_both(lib, c, 4)
    assert torch.equal(first, second)
    assert (second - want).abs().max().item() <= TOL


@pytest.mark.gpu
def test_noalloc_at_the_size_limit_and_after_dirty_scratch(lib):
    """A geometry that nearly fills the 2 MiB array (bs 30 x 16 q heads x 8 splits = 3840 of 3937 rows), run right after
    smaller calls have left their bytes all over it."""
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device")
    g = torch.Generator().manual_seed(77)
    bs, hq, hkv, bsz, nblk = 30, 16, 8, 256, 64
    lens = torch.randint(1, 2 * bsz + 1, (bs,), generator=g).to(torch.int32)
    lens[0], lens[1] = 1, 2 * bsz
    tables = torch.full((bs, 2), -1, dtype=torch.int32)
    perm = torch.randperm(nblk, generator=g).tolist()
    for b in range(bs):
        for j in range((int(lens[b]) + bsz - 1) // bsz):
            tables[b, j] = perm.pop()
    c = dict(q=torch.randn(bs, hq, 128, generator=g).to(torch.bfloat16),
             k_cache=torch.randn(nblk, bsz, hkv, 128, generator=g).to(torch.bfloat16),
             v_cache=torch.randn(nblk, bsz, hkv, 128, generator=g).to(torch.bfloat16),
             context_lens=lens, block_tables=tables, scale=128 ** -0.5, hq=hq, hkv=hkv, block_size=bsz)
    assert lib.nvllm_paged_decode_workspace_bytes(bs, hq, 128, 8) <= _scratch_limit()
    for splits in (8, 2, 8):
        got, want = _run_both(lib, c, splits)
        assert not torch.isnan(got).any()
        assert (got - want).abs().max().item() <= TOL, splits
