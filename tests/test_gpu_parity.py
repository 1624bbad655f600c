"""GPU parity tests: the sm_100a kernels, called through the C ABI, against the CPU oracle, the
fixtures frozen from the reference, the reference's own kernels on the box (Triton store,
flash-attn) and size-independent properties at BASELINE.json's full sizes.

Bars: cache contents and index tensors bit-exact; attention outputs max-abs <= 2e-2 (bf16, fp32
accumulate) as north_star states.
"""
import os
import sys

import numpy as np
import pytest
import torch

import cases
from oracle import paged_attention_oracle as orc

pytestmark = pytest.mark.gpu

TOL = 2e-2
GOLD = os.path.join(os.path.dirname(__file__), "golden")
DEV = "cuda"


@pytest.fixture(scope="module")
def ops(lib):
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device")
    from nano_vllm_learn_b200 import ops as _ops
    return _ops


def _cuda(c):
    return {k: (v.to(DEV) if isinstance(v, torch.Tensor) else v) for k, v in c.items()}


def _maxerr(a, b):
    return (a.float().cpu() - b.float().cpu()).abs().max().item()


# ------------------------------------------------------------------ store_kvcache: bit-exact
@pytest.mark.parametrize("n,hkv,dup,neg,strided", [(1, 8, 0, 0, True), (37, 8, 0, 0, True), (64, 8, 5, 0, True),
                                                   (130, 2, 0, 7, True), (16, 1, 0, 0, False), (513, 4, 2, 3, True)])
def test_store_bit_exact_vs_oracle(ops, n, hkv, dup, neg, strided):
    c = cases.make_store_case(100 + n, n, hkv, 4, strided_v=strided, dup_slot0=dup, neg=neg)
    kc, vc = c["k_cache"].clone(), c["v_cache"].clone()
    orc.store_kvcache_ref(c["key"], c["value"], kc, vc, c["slot_mapping"])
    g = _cuda(c)
    # keep V a strided view on the device too (as the fused qkv row in qwen3.py:106)
    if strided:
        row = hkv * 128
        fused = torch.zeros(n, 3 * row, dtype=torch.bfloat16, device=DEV)
        fused[:, 2 * row:] = c["value"].reshape(n, row).to(DEV)
        g["value"] = fused[:, 2 * row:].view(n, hkv, 128)
        assert n == 1 or g["value"].stride(0) == 3 * row
    ops.store_kvcache(g["key"], g["value"], g["k_cache"], g["v_cache"], g["slot_mapping"])
    torch.cuda.synchronize()
    assert torch.equal(g["k_cache"].cpu().view(torch.int16), kc.view(torch.int16))
    assert torch.equal(g["v_cache"].cpu().view(torch.int16), vc.view(torch.int16))


def test_store_full_size_roundtrip(ops):
    """Prefill-sized scatter (16384 tokens x 8 kv heads, the default max_num_batched_tokens): every
    stored row reads back bit-identical through its slot; untouched rows keep their old bits."""
    n, hkv, nblk = 16384, 8, 80
    g = torch.Generator(device=DEV).manual_seed(0)
    key = torch.randn(n, hkv, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    fused = torch.randn(n, 3 * hkv * 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    value = fused[:, 2 * hkv * 128:].view(n, hkv, 128)
    k_cache = torch.randn(nblk, 256, hkv, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    v_cache = k_cache.flip(0).clone()
    k0, v0 = k_cache.clone(), v_cache.clone()
    slots = torch.randperm(nblk * 256, generator=g, device=DEV)[:n].to(torch.int32)
    ops.store_kvcache(key, value, k_cache, v_cache, slots)
    torch.cuda.synchronize()
    idx = slots.long()
    assert torch.equal(k_cache.view(-1, hkv, 128)[idx].view(torch.int16), key.view(torch.int16))
    assert torch.equal(v_cache.view(-1, hkv, 128)[idx].view(torch.int16), value.contiguous().view(torch.int16))
    mask = torch.ones(nblk * 256, dtype=torch.bool, device=DEV)
    mask[idx] = False
    assert torch.equal(k_cache.view(-1, hkv, 128)[mask].view(torch.int16), k0.view(-1, hkv, 128)[mask].view(torch.int16))
    assert torch.equal(v_cache.view(-1, hkv, 128)[mask].view(torch.int16), v0.view(-1, hkv, 128)[mask].view(torch.int16))
    # idempotence
    k1 = k_cache.clone()
    ops.store_kvcache(key, value, k_cache, v_cache, slots)
    assert torch.equal(k1.view(torch.int16), k_cache.view(torch.int16))


def test_store_matches_reference_triton_kernel(ops):
    """Bit-exact against the reference's own Triton store kernel (nanovllm/layers/attention.py:19-55)."""
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")
    if not os.path.isdir(os.path.join(root, "nanovllm")):
        pytest.skip("baseline/_ref not present")
    sys.path.insert(0, root)
    try:
        from nanovllm.layers.attention import store_kvcache as ref_store
    except Exception as e:
        pytest.skip(f"reference flash backend not importable: {e}")
    c = _cuda(cases.make_store_case(7, 200, 8, 4, dup_slot0=4))
    k_ref, v_ref = c["k_cache"].clone(), c["v_cache"].clone()
    ref_store(c["key"], c["value"], k_ref, v_ref, c["slot_mapping"])
    ops.store_kvcache(c["key"], c["value"], c["k_cache"], c["v_cache"], c["slot_mapping"])
    torch.cuda.synchronize()
    assert torch.equal(k_ref.view(torch.int16), c["k_cache"].view(torch.int16))
    assert torch.equal(v_ref.view(torch.int16), c["v_cache"].view(torch.int16))


# ------------------------------------------------------------------ paged decode
# splits >= 1: fixed split-KV grid; (0 | -P, min_chunks): the balanced persistent grid with one CTA per SM | P CTAs,
# each owning at least `min_chunks` 16-token chunks.  min_chunks = 1 spreads even these small cases over many CTAs, so
# items are cut into many partials (cross-CTA combine); -1 / -3 / -7 CTAs make one CTA walk many segments.
BALANCED = [(0, 0), (0, 1), (-1, 1), (-7, 2), (-512, 1)]


@pytest.fixture()
def min_chunks(lib):
    def _set(n, warps=0):
        assert lib.nvllm_paged_decode_set_tuning(n, warps) == 0
    yield _set
    lib.nvllm_paged_decode_set_tuning(0, 0)


@pytest.mark.parametrize("name", list(cases.DECODE_CASES))
@pytest.mark.parametrize("ctas,mc", BALANCED)
@pytest.mark.parametrize("warps", [4, 8])       # two 4-warp CTAs per SM (G <= 4) / one 8-warp CTA per SM
def test_decode_balanced_vs_oracle_and_fixture(ops, min_chunks, name, ctas, mc, warps):
    c = cases.make_decode_case(name)
    want = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], c["scale"])
    g = _cuda(c)
    min_chunks(mc, warps)
    for _ in range(2):   # twice: the arrival counters of the shared workspace must have reset themselves
        got = ops.flash_attn_with_kvcache(g["q"].unsqueeze(1), g["k_cache"], g["v_cache"], g["context_lens"],
                                          g["block_tables"], softmax_scale=c["scale"], balanced_ctas=-ctas)
        assert _maxerr(got.squeeze(1), want) <= TOL
    gold = cases.from_bits(np.load(os.path.join(GOLD, f"decode_{name}.npz"))["out"])
    assert _maxerr(got.squeeze(1), gold) <= TOL   # the reference's own sdpa.math output


def test_num_splits_zero_means_automatic_like_flash_attn(ops, lib, monkeypatch):
    """flash-attn's `num_splits=0` is 'choose for me'; it must not select the (slower) balanced grid, which
    has its own keyword.  The plan cache follows NVLLM_DECODE_PLAN."""
    monkeypatch.delenv("NVLLM_DECODE_PLAN", raising=False)
    c = cases.make_decode_case("ragged_g4")
    g = _cuda(c)
    args = (g["q"], g["k_cache"], g["v_cache"], g["context_lens"], g["block_tables"])
    auto = ops.flash_attn_with_kvcache(*args, softmax_scale=c["scale"])
    n = lib.nvllm_launch_count()
    zero = ops.flash_attn_with_kvcache(*args, softmax_scale=c["scale"], num_splits=0)
    assert lib.nvllm_launch_count() == n + 1 and torch.equal(zero, auto)
    planned = ops.plan_num_splits(len(c["context_lens"]), c["hq"], c["hkv"], g["block_tables"].size(1) * 256)
    assert planned >= 1
    assert torch.equal(ops.flash_attn_with_kvcache(*args, softmax_scale=c["scale"], num_splits=planned), auto)
    with pytest.raises(ValueError):
        ops.flash_attn_with_kvcache(*args, softmax_scale=c["scale"], num_splits=-3)
    with pytest.raises(AssertionError):
        ops.flash_attn_with_kvcache(*args, softmax_scale=c["scale"], num_splits=2, balanced_ctas=0)
    bal = ops.flash_attn_with_kvcache(*args, softmax_scale=c["scale"], balanced_ctas=0)
    assert _maxerr(bal, auto) <= 4e-3
    # the cached plan follows the environment without clearing the cache by hand
    assert ops.plan_num_splits(64, 16, 8, 4096) == 1
    monkeypatch.setenv("NVLLM_DECODE_PLAN", "balanced")
    assert ops.plan_num_splits(64, 16, 8, 4096) == 0
    monkeypatch.delenv("NVLLM_DECODE_PLAN")
    assert ops.plan_num_splits(64, 16, 8, 4096) == 1


def test_decode_default_plan_is_the_fixed_grid(ops, monkeypatch):
    monkeypatch.delenv("NVLLM_DECODE_PLAN", raising=False)
    monkeypatch.setattr(ops, "_PLAN_CACHE", {})
    assert ops.plan_num_splits(64, 16, 8, 4096) == 1
    assert ops.plan_num_splits(1, 32, 8, 32768) == 16 and ops.plan_num_splits(2, 32, 8, 32768) == 8
    assert ops.plan_num_splits(1, 4, 1, 40960) == 64
    monkeypatch.setenv("NVLLM_DECODE_PLAN", "balanced")   # opt-in: the persistent balanced grid
    monkeypatch.setattr(ops, "_PLAN_CACHE", {})
    assert ops.plan_num_splits(64, 16, 8, 4096) == 0
    monkeypatch.setattr(ops, "_PLAN_CACHE", {})


@pytest.mark.parametrize("name", list(cases.DECODE_CASES))
@pytest.mark.parametrize("splits", [1, 2, 3, 8])
def test_decode_vs_oracle_and_fixture(ops, name, splits):
    c = cases.make_decode_case(name)
    want = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], c["scale"])
    g = _cuda(c)
    got = ops.flash_attn_with_kvcache(g["q"].unsqueeze(1), g["k_cache"], g["v_cache"], g["context_lens"],
                                      g["block_tables"], softmax_scale=c["scale"], num_splits=splits)
    assert got.shape == (len(c["context_lens"]), 1, c["hq"], 128)
    assert _maxerr(got.squeeze(1), want) <= TOL
    gold = cases.from_bits(np.load(os.path.join(GOLD, f"decode_{name}.npz"))["out"])
    assert _maxerr(got.squeeze(1), gold) <= TOL   # the reference's own sdpa.math output


@pytest.mark.parametrize("name", list(cases.DECODE_CASES))
@pytest.mark.parametrize("splits,cluster", [(2, 2), (4, 4), (8, 8), (16, 16), (8, 2), (16, 4), (12, 4), (16, 0), (6, 0)])
def test_decode_cluster_merge_vs_oracle_and_fixture(ops, lib, name, splits, cluster):
    """One-wave grids: the split CTAs of a (sequence, kv head) merge through the cluster leader's shared memory; one
    workspace partial per cluster (splits / cluster of them), none when the cluster holds every split."""
    c = cases.make_decode_case(name)
    want = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], c["scale"])
    g = _cuda(c)
    assert lib.nvllm_paged_decode_set_cluster(3) != 0
    assert lib.nvllm_paged_decode_set_cluster(cluster) == 0
    try:
        for _ in range(2):   # twice: the arrival counters must have reset themselves
            got = ops.flash_attn_with_kvcache(g["q"], g["k_cache"], g["v_cache"], g["context_lens"], g["block_tables"],
                                              softmax_scale=c["scale"], num_splits=splits)
            assert _maxerr(got, want) <= TOL
        lib.nvllm_paged_decode_set_cluster(1)
        flat = ops.flash_attn_with_kvcache(g["q"], g["k_cache"], g["v_cache"], g["context_lens"], g["block_tables"],
                                           softmax_scale=c["scale"], num_splits=splits)
        assert _maxerr(got, flat) <= 4e-3      # same partition of the tokens, a different merge tree
    finally:
        lib.nvllm_paged_decode_set_cluster(0)
    gold = cases.from_bits(np.load(os.path.join(GOLD, f"decode_{name}.npz"))["out"])
    assert _maxerr(got, gold) <= TOL


def test_decode_all_variants(ops, lib):
    c = cases.make_decode_case("ragged_g4")
    want = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], c["scale"])
    g = _cuda(c)
    try:
        for v in range(lib.nvllm_paged_decode_num_variants()):
            assert lib.nvllm_paged_decode_set_variant(v) == 0
            for splits in (1, 4):
                got = ops.flash_attn_with_kvcache(g["q"], g["k_cache"], g["v_cache"], g["context_lens"],
                                                  g["block_tables"], softmax_scale=c["scale"], num_splits=splits)
                assert _maxerr(got, want) <= TOL, (v, splits)
    finally:
        lib.nvllm_paged_decode_set_variant(-1)


def test_decode_edges(ops, min_chunks):
    c = cases.make_decode_case("ragged_g2_tail")
    g = _cuda(c)
    min_chunks(2)
    # ctx == 0 rows (zero-padded graph rows) -> zeros, other rows unchanged; padding entries may be 0 or -1
    lens = g["context_lens"].clone()
    lens[1] = 0
    tables = g["block_tables"].clone()
    tables[tables < 0] = 0
    for splits in (1, 4, 0, -5):
        kw = dict(num_splits=splits) if splits >= 1 else dict(balanced_ctas=-splits)   # <= 0: the balanced grid
        got = ops.flash_attn_with_kvcache(g["q"], g["k_cache"], g["v_cache"], lens, tables,
                                          softmax_scale=c["scale"], **kw)
        ref = ops.flash_attn_with_kvcache(g["q"], g["k_cache"], g["v_cache"], g["context_lens"], g["block_tables"],
                                          softmax_scale=c["scale"], **kw)
        assert torch.count_nonzero(got[1]) == 0
        if splits >= 1:    # the fixed grid gives every (sequence, kv head) its own CTAs: other rows keep their bits
            assert torch.equal(got[0], ref[0]) and torch.equal(got[2:], ref[2:])
        else:              # the balanced grid cuts the live work elsewhere once a row is empty
            assert _maxerr(got[0], ref[0]) <= 4e-3 and _maxerr(got[2:], ref[2:]) <= 4e-3
    # every row empty (a fully padded captured batch): zeros from the balanced grid too
    got = ops.flash_attn_with_kvcache(g["q"], g["k_cache"], g["v_cache"], torch.zeros_like(lens), tables,
                                      softmax_scale=c["scale"], balanced_ctas=0)
    assert torch.count_nonzero(got) == 0
    # bs == 0
    empty = ops.flash_attn_with_kvcache(g["q"][:0], g["k_cache"], g["v_cache"], lens[:0], tables[:0],
                                        softmax_scale=c["scale"])
    assert empty.shape == (0, c["hq"], 128)
    # strided q rows (a slice of a wider activation)
    wide = torch.zeros(len(lens), c["hq"] * 128 + 256, dtype=torch.bfloat16, device=DEV)
    wide[:, :c["hq"] * 128] = g["q"].reshape(len(lens), -1)
    qv = wide[:, :c["hq"] * 128].view(len(lens), c["hq"], 128)
    got = ops.flash_attn_with_kvcache(qv, g["k_cache"], g["v_cache"], g["context_lens"], g["block_tables"],
                                      softmax_scale=c["scale"], num_splits=1)
    ref = ops.flash_attn_with_kvcache(g["q"], g["k_cache"], g["v_cache"], g["context_lens"], g["block_tables"],
                                      softmax_scale=c["scale"], num_splits=1)
    assert torch.equal(got, ref)


def _full_decode_inputs(bs, ctx_list, hq, hkv, seed=0, extra_blocks=8):
    g = torch.Generator(device=DEV).manual_seed(seed)
    need = [(c + 255) // 256 for c in ctx_list]
    nblk = sum(need) + extra_blocks
    k_cache = torch.randn(nblk, 256, hkv, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    v_cache = torch.randn(nblk, 256, hkv, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    q = torch.randn(bs, hq, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    perm = torch.randperm(nblk, generator=g, device=DEV).to(torch.int32)
    tables = torch.full((bs, max(need)), -1, dtype=torch.int32, device=DEV)
    pos = 0
    for b, n in enumerate(need):
        tables[b, :n] = perm[pos:pos + n]
        pos += n
    lens = torch.tensor(ctx_list, dtype=torch.int32, device=DEV)
    return q, k_cache, v_cache, lens, tables


def test_decode_full_size_properties(ops):
    """bench_my's worst point: bs=64, ctx 4096, Qwen3-0.6B heads (16 q / 8 kv).  1.07 GB of K/V."""
    bs, hq, hkv = 64, 16, 8
    ctx = [4096 - (i % 7) for i in range(bs)]
    q, kc, vc, lens, tables = _full_decode_inputs(bs, ctx, hq, hkv)
    scale = 128 ** -0.5
    o1 = ops.flash_attn_with_kvcache(q, kc, vc, lens, tables, softmax_scale=scale, num_splits=1)
    # (i) split invariance (planned splits and an odd count)
    for kw in (dict(num_splits=None), dict(num_splits=0), dict(balanced_ctas=0), dict(balanced_ctas=37),
               dict(num_splits=2), dict(num_splits=5)):
        os_ = ops.flash_attn_with_kvcache(q, kc, vc, lens, tables, softmax_scale=scale, **kw)
        assert _maxerr(os_, o1) <= 4e-3, kw
    # the balanced grid is deterministic and page-placement invariant as well
    ob = ops.flash_attn_with_kvcache(q, kc, vc, lens, tables, softmax_scale=scale, balanced_ctas=0)
    assert torch.equal(ob, ops.flash_attn_with_kvcache(q, kc, vc, lens, tables, softmax_scale=scale, balanced_ctas=0))
    # (ii) page-permutation invariance: move every page elsewhere, remap the table -> identical bits
    nblk = kc.shape[0]
    perm = torch.randperm(nblk, device=DEV)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(nblk, device=DEV)
    kc2, vc2 = kc[perm].contiguous(), vc[perm].contiguous()      # new page j holds old page perm[j]
    t2 = tables.clone()
    live = tables >= 0
    t2[live] = inv[tables[live].long()].to(torch.int32)
    o2 = ops.flash_attn_with_kvcache(q, kc2, vc2, lens, t2, softmax_scale=scale, num_splits=1)
    assert torch.equal(o1, o2)
    assert torch.equal(ob, ops.flash_attn_with_kvcache(q, kc2, vc2, lens, t2, softmax_scale=scale, balanced_ctas=0))
    # (iii) linearity in V (K fixed -> identical softmax weights): out(V) + out(V') == out(V + V') up to bf16 rounding
    vb = torch.randn_like(vc)
    oa = ops.flash_attn_with_kvcache(q, kc, vb, lens, tables, softmax_scale=scale, num_splits=1)
    osum = ops.flash_attn_with_kvcache(q, kc, (vc.float() + vb.float()).to(torch.bfloat16), lens, tables,
                                       softmax_scale=scale, num_splits=1)
    assert _maxerr(osum, o1.float() + oa.float()) <= TOL
    # (iv) three sampled sequences against the CPU oracle
    for b in (0, 31, 63):
        want = orc.paged_decode_ref(q[b:b + 1].cpu(), kc.cpu(), vc.cpu(), lens[b:b + 1].cpu(), tables[b:b + 1].cpu(), scale)
        assert _maxerr(o1[b:b + 1], want) <= TOL


def test_decode_balanced_ragged_batch(ops):
    """Ragged live contexts incl. empty rows, 1-token rows and page-boundary lengths, on a wide zero-padded table (the
    captured-graph layout): the balanced grid (whose ranges cut items anywhere) against the one-CTA-per-item kernel
    and, for sampled rows, the CPU oracle.  8B-shape heads at tp=2 (16 q / 4 kv)."""
    ctx = [0, 1, 15, 16, 17, 255, 256, 257, 4096, 513, 0, 3000, 31, 1024, 2047, 2049] * 3 + [8191, 7, 0, 777]
    bs = len(ctx)
    q, kc, vc, lens, tables = _full_decode_inputs(bs, [max(x, 1) for x in ctx], 16, 4, seed=11)
    lens = torch.tensor(ctx, dtype=torch.int32, device=DEV)
    wide = torch.zeros(bs, 40, dtype=torch.int32, device=DEV)
    wide[:, :tables.shape[1]] = tables.clamp(min=0)
    scale = 128 ** -0.5
    ref = ops.flash_attn_with_kvcache(q, kc, vc, lens, wide, softmax_scale=scale, num_splits=1)
    for ctas in (0, -148, -61, -9):
        got = ops.flash_attn_with_kvcache(q, kc, vc, lens, wide, softmax_scale=scale, balanced_ctas=-ctas)
        assert _maxerr(got, ref) <= 4e-3, ctas
        assert torch.count_nonzero(got[lens == 0]) == 0
    for b in (1, 8, 19, 48):
        want = orc.paged_decode_ref(q[b:b + 1].cpu(), kc.cpu(), vc.cpu(), lens[b:b + 1].cpu(), wide[b:b + 1].cpu(), scale)
        assert _maxerr(got[b:b + 1], want) <= TOL


def test_decode_long_context_split(ops):
    """configs[4]: Qwen3-8B-shape (32 q / 8 kv), bs=2, ctx 32768, table width 160 (max_model_len 40960)."""
    q, kc, vc, lens, tables = _full_decode_inputs(2, [32768, 30001], 32, 8, seed=3)
    wide = torch.zeros(2, 160, dtype=torch.int32, device=DEV)          # graph-style zero padding
    wide[:, :tables.shape[1]] = tables.clamp(min=0)
    scale = 128 ** -0.5
    o_plan = ops.flash_attn_with_kvcache(q, kc, vc, lens, wide, softmax_scale=scale)      # planned split count
    o_one = ops.flash_attn_with_kvcache(q, kc, vc, lens, tables, softmax_scale=scale, num_splits=1)
    assert _maxerr(o_plan, o_one) <= 4e-3
    o_fixed = ops.flash_attn_with_kvcache(q, kc, vc, lens, wide, softmax_scale=scale, num_splits=9)
    assert _maxerr(o_fixed, o_one) <= 4e-3
    want = orc.paged_decode_ref(q[1:2].cpu(), kc.cpu(), vc.cpu(), lens[1:2].cpu(), tables[1:2].cpu(), scale)
    assert _maxerr(o_plan[1:2], want) <= TOL


def test_decode_vs_flash_attn(ops):
    try:
        from flash_attn import flash_attn_with_kvcache as fa
    except Exception as e:
        pytest.skip(f"flash-attn not importable: {e}")
    q, kc, vc, lens, tables = _full_decode_inputs(16, [513 + 97 * i for i in range(16)], 16, 8, seed=5)
    scale = 128 ** -0.5
    ref = fa(q.unsqueeze(1), kc, vc, cache_seqlens=lens, block_table=tables.clamp(min=0), softmax_scale=scale,
             causal=True)
    got = ops.flash_attn_with_kvcache(q.unsqueeze(1), kc, vc, lens, tables, softmax_scale=scale)
    assert _maxerr(got, ref) <= TOL


def test_decode_cuda_graph_replay(ops):
    """Capture once with zeroed metadata (as capture_cudagraph does), replay with live metadata."""
    c = cases.make_decode_case("ragged_g4")
    g = _cuda(c)
    bs = len(c["context_lens"])
    width = 16
    q_buf = torch.zeros_like(g["q"])
    lens_buf = torch.zeros(bs, dtype=torch.int32, device=DEV)
    tab_buf = torch.zeros(bs, width, dtype=torch.int32, device=DEV)
    out_buf = torch.zeros_like(g["q"])
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        out_buf.copy_(ops.flash_attn_with_kvcache(q_buf, g["k_cache"], g["v_cache"], lens_buf, tab_buf,
                                                  softmax_scale=c["scale"]))
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out_buf.copy_(ops.flash_attn_with_kvcache(q_buf, g["k_cache"], g["v_cache"], lens_buf, tab_buf,
                                                  softmax_scale=c["scale"]))
    want = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], c["scale"])
    for _ in range(2):
        q_buf.copy_(g["q"])
        lens_buf.copy_(g["context_lens"])
        tab_buf.zero_()
        tab_buf[:, :g["block_tables"].shape[1]] = g["block_tables"]
        graph.replay()
        torch.cuda.synchronize()
        assert _maxerr(out_buf, want) <= TOL
        # shorter contexts on the same graph
        lens_buf.copy_(torch.clamp(g["context_lens"] // 2, min=1))
        graph.replay()
        torch.cuda.synchronize()
        half = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], torch.clamp(c["context_lens"] // 2, min=1),
                                    c["block_tables"], c["scale"])
        assert _maxerr(out_buf, half) <= TOL


# ------------------------------------------------------------------ varlen prefill
@pytest.mark.parametrize("name", list(cases.PREFILL_CASES))
def test_prefill_vs_oracle_and_fixture(ops, name):
    c = cases.make_prefill_case(name)
    want = orc.varlen_prefill_ref(c["q"], c["k"], c["v"], c["cu_seqlens_q"], c["cu_seqlens_k"], c["scale"])
    g = _cuda(c)
    got = ops.flash_attn_varlen_func(g["q"], g["k"], g["v"], c["max_seqlen_q"], g["cu_seqlens_q"], c["max_seqlen_k"],
                                     g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True)
    assert _maxerr(got, want) <= TOL
    gold = cases.from_bits(np.load(os.path.join(GOLD, f"prefill_{name}.npz"))["out"])
    assert _maxerr(got, gold) <= TOL


@pytest.mark.parametrize("name", list(cases.PREFIX_CASES))
def test_prefix_prefill_vs_oracle(ops, name):
    c = cases.make_prefix_case(name)
    want = orc.varlen_prefill_ref(c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["cu_seqlens_k"], c["scale"],
                                  c["block_tables"])
    g = _cuda(c)
    got = ops.flash_attn_varlen_func(g["q"], g["k_cache"], g["v_cache"], c["max_seqlen_q"], g["cu_seqlens_q"],
                                     c["max_seqlen_k"], g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True,
                                     block_table=g["block_tables"])
    assert _maxerr(got, want) <= TOL


def test_prefix_prefill_vs_flash_attn(ops):
    """The reference's flash backend on the box is the ground truth for the prefix-cache case."""
    try:
        from flash_attn import flash_attn_varlen_func as fa
    except Exception as e:
        pytest.skip(f"flash-attn not importable: {e}")
    for name in cases.PREFIX_CASES:
        c = cases.make_prefix_case(name)
        g = _cuda(c)
        ref = fa(g["q"], g["k_cache"], g["v_cache"], max_seqlen_q=c["max_seqlen_q"], cu_seqlens_q=g["cu_seqlens_q"],
                 max_seqlen_k=c["max_seqlen_k"], cu_seqlens_k=g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True,
                 block_table=g["block_tables"].clamp(min=0))
        got = ops.flash_attn_varlen_func(g["q"], g["k_cache"], g["v_cache"], c["max_seqlen_q"], g["cu_seqlens_q"],
                                         c["max_seqlen_k"], g["cu_seqlens_k"], softmax_scale=c["scale"], causal=True,
                                         block_table=g["block_tables"])
        assert _maxerr(got, ref) <= TOL, name


def test_prefill_full_size_consistency(ops):
    """bench.py's prefill step: 16 sequences x 1024 tokens, 16 q / 8 kv heads.  The last query of every
    sequence must agree with a paged decode over the same keys (cross-kernel), sampled rows with the oracle."""
    bsz, L, hq, hkv = 16, 1024, 16, 8
    g = torch.Generator(device=DEV).manual_seed(9)
    total = bsz * L
    q = torch.randn(total, hq, 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    fused = torch.randn(total, 2 * hkv * 128, generator=g, device=DEV, dtype=torch.float32).to(torch.bfloat16)
    k = fused[:, :hkv * 128].view(total, hkv, 128)           # strided views like the fused qkv output
    v = fused[:, hkv * 128:].view(total, hkv, 128)
    cu = torch.arange(0, total + 1, L, dtype=torch.int32, device=DEV)
    scale = 128 ** -0.5
    out = ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=scale, causal=True)
    # store everything into a paged cache, then decode the last token of each sequence
    nblk = bsz * L // 256
    kc = torch.zeros(nblk, 256, hkv, 128, dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros_like(kc)
    perm = torch.randperm(nblk, generator=g, device=DEV)
    tables = perm.view(bsz, L // 256).to(torch.int32)
    slots = (tables.long().repeat_interleave(256, dim=1) * 256 + torch.arange(256, device=DEV).repeat(L // 256)).view(-1)
    ops.store_kvcache(k, v, kc, vc, slots.to(torch.int32))
    lens = torch.full((bsz,), L, dtype=torch.int32, device=DEV)
    q_last = q.view(bsz, L, hq, 128)[:, -1].contiguous()
    dec = ops.flash_attn_with_kvcache(q_last, kc, vc, lens, tables, softmax_scale=scale, num_splits=1)
    assert _maxerr(dec, out.view(bsz, L, hq, 128)[:, -1]) <= 8e-3
    # paged prefill over the same cache (block_tables set, cached == 0) equals the dense prefill
    out_paged = ops.flash_attn_varlen_func(q, kc, vc, L, cu, L, cu, softmax_scale=scale, causal=True, block_table=tables)
    assert _maxerr(out_paged, out) <= 1e-6
    # oracle on one sequence
    i = 5
    want = orc.varlen_prefill_ref(q[i * L:(i + 1) * L].cpu(), k[i * L:(i + 1) * L].cpu(), v[i * L:(i + 1) * L].cpu(),
                                  torch.tensor([0, L], dtype=torch.int32), torch.tensor([0, L], dtype=torch.int32), scale)
    assert _maxerr(out[i * L:(i + 1) * L], want) <= TOL


def test_prefill_zero_length_sequence(ops):
    """A fully cached prompt gives Lq == 0 for that sequence: no rows, neighbours intact."""
    c = cases.make_prefix_case("px_one_token")
    g = _cuda(c)
    cu_q = torch.tensor([0, 0, 64], dtype=torch.int32, device=DEV)
    cu_k = torch.tensor([0, 256, 256 + 576], dtype=torch.int32, device=DEV)
    q = g["q"][1:65]
    got = ops.flash_attn_varlen_func(q, g["k_cache"], g["v_cache"], 64, cu_q, 576, cu_k, softmax_scale=c["scale"],
                                     causal=True, block_table=g["block_tables"])
    want = orc.varlen_prefill_ref(q.cpu(), c["k_cache"], c["v_cache"], cu_q.cpu(), cu_k.cpu(), c["scale"],
                                  c["block_tables"])
    assert _maxerr(got, want) <= TOL


# ------------------------------------------------------------------ the Attention module (drop-in boundary)
def test_attention_module_prefill_then_decode(ops):
    from nano_vllm_learn_b200 import set_context, reset_context, get_context
    from nano_vllm_learn_b200 import attention_cuda
    from nano_vllm_learn_b200.attention_cuda import Attention
    attention_cuda.bind_context(get_context)   # the engine e2e test may have bound nanovllm's context in this process
    hq, hkv, nblk = 16, 8, 12
    attn = Attention(hq, 128, 128 ** -0.5, hkv)
    assert attn.k_cache.numel() == 0 and attn.v_cache.numel() == 0
    g = torch.Generator().manual_seed(42)
    lens = [300, 77]
    total = sum(lens)
    q = torch.randn(total, hq * 128, generator=g).to(torch.bfloat16)
    k = torch.randn(total, hkv * 128, generator=g).to(torch.bfloat16)
    v = torch.randn(total, hkv * 128, generator=g).to(torch.bfloat16)
    cu = torch.tensor([0, 300, 377], dtype=torch.int32)
    tables = torch.tensor([[7, 2], [9, -1]], dtype=torch.int32)
    slots = torch.cat([torch.arange(7 * 256, 8 * 256), torch.arange(2 * 256, 2 * 256 + 44),
                       torch.arange(9 * 256, 9 * 256 + 77)]).to(torch.int32)
    # warmup mode: no cache bound -> attention only (attention.py:72,84)
    set_context(True, cu.to(DEV), cu.to(DEV), 300, 300, slots.to(DEV), None, None)
    o = attn(q.to(DEV), k.to(DEV), v.to(DEV))
    want = orc.varlen_prefill_ref(q.view(total, hq, 128), k.view(total, hkv, 128), v.view(total, hkv, 128), cu, cu,
                                  128 ** -0.5)
    assert o.shape == (total, hq * 128) and _maxerr(o.view(total, hq, 128), want) <= TOL
    # bind a cache, prefill again: the store must land bit-exact
    kv = torch.zeros(2, nblk, 256, hkv, 128, dtype=torch.bfloat16, device=DEV)
    attn.k_cache, attn.v_cache = kv[0], kv[1]
    o = attn(q.to(DEV), k.to(DEV), v.to(DEV))
    kc = torch.zeros(nblk, 256, hkv, 128, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    orc.store_kvcache_ref(k.view(total, hkv, 128), v.view(total, hkv, 128), kc, vc, slots)
    assert torch.equal(kv[0].cpu().view(torch.int16), kc.view(torch.int16))
    assert torch.equal(kv[1].cpu().view(torch.int16), vc.view(torch.int16))
    # one decode step for both sequences
    qd = torch.randn(2, hq * 128, generator=g).to(torch.bfloat16)
    kd = torch.randn(2, hkv * 128, generator=g).to(torch.bfloat16)
    vd = torch.randn(2, hkv * 128, generator=g).to(torch.bfloat16)
    dslots = torch.tensor([2 * 256 + 44, 9 * 256 + 77], dtype=torch.int32)
    ctx = torch.tensor([301, 78], dtype=torch.int32)
    set_context(False, slot_mapping=dslots.to(DEV), context_lens=ctx.to(DEV), block_tables=tables.to(DEV))
    od = attn(qd.to(DEV), kd.to(DEV), vd.to(DEV))
    orc.store_kvcache_ref(kd.view(2, hkv, 128), vd.view(2, hkv, 128), kc, vc, dslots)
    want = orc.paged_decode_ref(qd.view(2, hq, 128), kc, vc, ctx, tables, 128 ** -0.5)
    assert od.shape == (2, hq * 128) and _maxerr(od.view(2, hq, 128), want) <= TOL
    assert torch.equal(kv[0].cpu().view(torch.int16), kc.view(torch.int16))
    reset_context()
