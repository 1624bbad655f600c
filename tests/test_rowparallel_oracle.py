"""CPU: the RowParallelLinear oracle against outputs of the reference class itself (tests/golden/rowparallel.npz,
frozen by tests/golden/make_golden_rowparallel.py under gloo), the host-side plan/size arithmetic of the C ABI, and the
module patch under a world_size-2 gloo group."""
import os
import sys

import numpy as np
import pytest
import torch

import cases
from oracle import rowparallel_oracle as rpo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "rowparallel.npz"))
ULP = 2.0 ** -8   # bf16 has 8 significand bits: one rounding is <= 2^-9 relative, two roundings <= ~2^-8


@pytest.mark.parametrize("name", list(cases.ROWPARALLEL_CASES))
def test_oracle_matches_reference_fixture(name):
    c = cases.make_rowparallel_case(name)
    assert cases.sha(c["x"], c["weight"], c["bias"]) == str(GOLD[f"{name}_sha"]), "fixture inputs differ"
    want = cases.from_bits(GOLD[f"{name}_y"])
    xs, ws = rpo.shard_inputs(c["x"], c["weight"], c["world"])
    got = rpo.rowparallel_ref(xs, ws, c["bias"])
    if c["world"] == 2:
        # one bf16 addition of two bf16 partials: no freedom in the order -> bit-exact
        assert torch.equal(got.view(torch.int16), want.view(torch.int16))
    else:
        # summation order of the collective is implementation-defined: bound, don't match
        scale = want.float().abs().max().item()
        assert (got.float() - want.float()).abs().max().item() <= 4 * ULP * scale
    # and the reference itself is within rounding of the exact contraction (what the kernel targets)
    exact = rpo.rowparallel_exact(xs, ws, c["bias"])
    scale = exact.abs().max().item()
    assert (want.double() - exact).abs().max().item() <= (1 + c["world"]) * ULP * scale


def test_oracle_against_live_reference_if_present():
    if rpo._ref_root() is None:
        pytest.skip("reference tree not on this machine")
    c = cases.make_rowparallel_case("rp_bias_tp2")
    ys = rpo.run_reference_world(c["x"], c["weight"], c["bias"], 2, port=29760 + os.getpid() % 200)
    xs, ws = rpo.shard_inputs(c["x"], c["weight"], 2)
    got = rpo.rowparallel_ref(xs, ws, c["bias"])
    for y in ys:
        assert torch.equal(y.view(torch.int16), got.view(torch.int16))


def test_comm_bytes_and_plan_are_host_arithmetic(lib):
    # Qwen3-0.6B tp=2: o_proj (1024 -> 1024) and down_proj (1536 -> 1024); Qwen3-8B tp=2/8
    for world, k, h in [(2, 1024, 1024), (2, 1536, 1024), (2, 2048, 4096), (8, 512, 4096), (8, 1536, 4096)]:
        b = lib.nvllm_rowparallel_comm_bytes(world, 512, k, h)
        assert b > 4096 and b % 16 == 0 and b < (1 << 30)
        assert lib.nvllm_rowparallel_comm_bytes(world, 16, k, h) <= b
    # unsupported shapes size to 0: k not a multiple of 64, out not a multiple of 64, world out of range
    assert lib.nvllm_rowparallel_comm_bytes(2, 512, 1000, 1024) == 0
    assert lib.nvllm_rowparallel_comm_bytes(2, 512, 1024, 1000) == 0
    assert lib.nvllm_rowparallel_comm_bytes(1, 512, 1024, 1024) == 0
    assert lib.nvllm_rowparallel_comm_bytes(9, 512, 1024, 1024) == 0
    import ctypes
    bh, ks, mode = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
    for n in (1, 16, 64, 128, 129, 512):
        for world, k, h in [(2, 1024, 1024), (2, 2048, 4096), (8, 512, 4096)]:
            assert lib.nvllm_rowparallel_plan(n, k, h, world, ctypes.byref(bh), ctypes.byref(ks), ctypes.byref(mode)) == 0
            tiles = (h // bh.value) * ((n + 127) // 128)
            assert bh.value in (32, 64, 128) and ks.value in (1, 2, 4, 8) and mode.value in (0, 1, 3)
            # the tcgen05 tile (128 features) only from 128 tokens up (64 where k per rank >= 4096)
            # and only on shards of >= 8 Mi elements (small shards would be left with too few CTAs)
            assert (bh.value == 128) == (n >= 128 and h % 128 == 0 and k * h >= (8 << 20))
            assert tiles * ks.value <= 296, "grid must fit one co-resident wave (2 CTAs x 148 SMs)"
            assert k % (ks.value * 64) == 0
    assert lib.nvllm_rowparallel_plan(513, 1024, 1024, 2, None, None, None) == 2
    assert lib.nvllm_rowparallel_plan(0, 1024, 1024, 2, None, None, None) == 2
    # forcing is validated
    assert lib.nvllm_rowparallel_set_plan(48, 0, -1, 0) == 1
    assert lib.nvllm_rowparallel_set_plan(0, 3, -1, 0) == 1
    assert lib.nvllm_rowparallel_set_plan(0, 0, 4, 0) == 1   # 2: local-partial profiling mode, 3: flag-in-data one-shot
    assert lib.nvllm_rowparallel_set_plan(128, 0, -1, 0) == 0
    assert lib.nvllm_rowparallel_plan(16, 1024, 1024, 2, None, None, None) == 2     # the tcgen05 tile stages >= 64 token rows
    assert lib.nvllm_rowparallel_plan(64, 1024, 1024, 2, ctypes.byref(bh), None, None) == 0 and bh.value == 128
    assert lib.nvllm_rowparallel_set_tc_min_tokens(0) == 0
    assert lib.nvllm_rowparallel_set_plan(0, 0, -1, 0) == 0
    assert lib.nvllm_rowparallel_plan(512, 2048, 4096, 2, ctypes.byref(bh), None, None) == 0 and bh.value == 64
    assert lib.nvllm_rowparallel_set_tc_min_tokens(128) == 0
    assert lib.nvllm_rowparallel_plan(512, 2048, 4096, 2, ctypes.byref(bh), None, None) == 0 and bh.value == 128
    assert lib.nvllm_rowparallel_plan(512, 1024, 1024, 2, ctypes.byref(bh), None, None) == 0 and bh.value == 64
    assert lib.nvllm_rowparallel_plan(64, 4096, 4096, 2, ctypes.byref(bh), None, None) == 0 and bh.value == 128
    assert lib.nvllm_rowparallel_plan(64, 2048, 4096, 2, ctypes.byref(bh), None, None) == 0 and bh.value == 64
    assert lib.nvllm_rowparallel_set_plan(64, 2, 1, 0) == 0
    assert lib.nvllm_rowparallel_plan(16, 1024, 1024, 2, ctypes.byref(bh), ctypes.byref(ks), ctypes.byref(mode)) == 0
    assert (bh.value, ks.value, mode.value) == (64, 2, 1)
    assert lib.nvllm_rowparallel_set_plan(0, 0, -1, 0) == 0


def test_launch_argument_validation_without_gpu(lib):
    import ctypes
    ptrs = (ctypes.c_void_p * 2)(4096, 8192)
    f = lib.nvllm_rowparallel_linear_allreduce
    assert f(16, 1024, 16, 1024, 16, 1024, None, 0, 1024, 1024, ptrs, 1 << 26, 0, 2, None) == 0      # n == 0: no-op
    assert f(None, 1024, 16, 1024, 16, 1024, None, 4, 1024, 1024, ptrs, 1 << 26, 0, 2, None) == 1
    assert f(16, 1024, 16, 1024, 16, 1024, None, 4, 1024, 1024, ptrs, 1 << 26, 2, 2, None) == 1     # rank >= world
    assert f(16, 1024, 8, 1024, 16, 1024, None, 4, 1024, 1024, ptrs, 1 << 26, 0, 2, None) == 3      # misaligned x
    assert f(16, 1024, 16, 1028, 16, 1024, None, 4, 1024, 1024, ptrs, 1 << 26, 0, 2, None) == 3     # row stride
    assert f(16, 512, 16, 1024, 16, 1024, None, 4, 1024, 1024, ptrs, 1 << 26, 0, 2, None) == 1      # y stride < out
    assert f(16, 1024, 16, 1024, 16, 1024, None, 600, 1024, 1024, ptrs, 1 << 26, 0, 2, None) == 2   # > 512 tokens
    assert f(16, 1024, 16, 1024, 16, 1024, None, 4, 1024, 1024, ptrs, 4096, 0, 2, None) == 5         # buffer too small


def _patch_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import importlib.util
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        spec = importlib.util.spec_from_file_location("ref_linear", os.path.join(rpo._ref_root(), "nanovllm", "layers", "linear.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        from nano_vllm_learn_b200 import rowparallel_cuda
        rowparallel_cuda.install(mod.RowParallelLinear)
        os.environ["NVLLM_FUSED_ALLREDUCE"] = "1"
        rowparallel_cuda.configure(False)              # another backend: every layer stays on the reference path
        ok0 = not mod.RowParallelLinear(1024, 1024)._nvllm_fused and not rowparallel_cuda._SHAPES
        rowparallel_cuda.configure(True)
        c = cases.make_rowparallel_case("rp_qwen06_down_tp2")
        m = mod.RowParallelLinear(c["k"], c["h"]).to(torch.bfloat16)
        m.weight.weight_loader(m.weight, c["weight"])
        ok = ok0 and m._nvllm_fused and (c["k"] // world, c["h"]) in rowparallel_cuda._SHAPES
        odd = mod.RowParallelLinear(200, 100)          # shape the kernel does not take: stays on the reference path
        ok = ok and not odd._nvllm_fused
        # the engine hands over batches by shard and world size (measured crossover vs cuBLAS + NCCL): 3 MiB shard at
        # 2 ranks -> 512 tokens, the 8B-shape shards -> 128; fewer tokens at 4 and 8 ranks
        ok = ok and m._nvllm_fused_max == 512 and rowparallel_cuda.engine_max_tokens(2048, 4096) == 128 \
            and rowparallel_cuda.engine_max_tokens(6144, 4096) == 128 \
            and rowparallel_cuda.engine_max_tokens(1024, 1024, 4) == 128 and rowparallel_cuda.engine_max_tokens(6144, 4096, 4) == 64 \
            and rowparallel_cuda.engine_max_tokens(1024, 1024, 8) == 64
        # env unset: on at the validated world sizes (2 is one); "0": the reference's collective
        del os.environ["NVLLM_FUSED_ALLREDUCE"]
        rowparallel_cuda.configure(True)
        ok = ok and rowparallel_cuda.enabled() and world in rowparallel_cuda.VALIDATED_WORLDS
        os.environ["NVLLM_FUSED_ALLREDUCE"] = "0"
        rowparallel_cuda.configure(True)
        ok = ok and not rowparallel_cuda.enabled() and not mod.RowParallelLinear(1024, 1024)._nvllm_fused
        os.environ["NVLLM_FUSED_ALLREDUCE"] = "1"
        rowparallel_cuda.configure(True)
        rowparallel_cuda._SHAPES.add((c["k"] // world, c["h"]))
        # CPU tensors never reach the CUDA op: the patched forward must be the reference's own
        with torch.no_grad():
            y = m(c["x"].chunk(world, dim=1)[rank].contiguous())
        xs, ws = rpo.shard_inputs(c["x"], c["weight"], world)
        ok = ok and torch.equal(y.view(torch.int16), rpo.rowparallel_ref(xs, ws).view(torch.int16))
        need = rowparallel_cuda.TPComm.required_bytes(world, sorted(rowparallel_cuda._SHAPES))
        sizes = [None] * world
        dist.all_gather_object(sizes, need)
        ok = ok and len(set(sizes)) == 1 and need > 0     # every rank sizes its buffer identically
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_module_patch_world2_gloo(lib):
    if rpo._ref_root() is None:
        pytest.skip("reference tree not on this machine")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29800 + os.getpid() % 150
    procs = [ctx.Process(target=_patch_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert res == {0: True, 1: True}
