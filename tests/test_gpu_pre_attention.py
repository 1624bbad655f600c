"""GPU: fused q/k RMSNorm + RoPE + store kernel vs the oracle and the fixture frozen
from the reference's RMSNorm / RotaryEmbedding modules.  V rows are pure data movement (bit-exact); q/k
are fp32 arithmetic rounded to bf16 three times (norm, weight, rope): the fp32 sum-of-squares is reduced in
a different order, so an intermediate may round the other way — at most 2 bf16 ulp on < 1 % of elements."""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import paged_attention_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _ulp_check(got, want, max_frac=0.01):
    got, want = got.float().cpu(), want.float().cpu()
    diff = (got - want).abs()
    # RoPE combines two elements (x1*cos - x2*sin), so the error scale of an output is the magnitude of its
    # inputs, not of the (possibly cancelled) result: 2 bf16 ulp of the head row's largest element
    ulp = want.abs().amax(dim=-1, keepdim=True) * 2.0 ** -6 + 1e-30
    assert (diff <= ulp).all(), f"more than 2 bf16 ulp (of the row scale) off: max diff {diff.max().item()}"
    assert (diff > 0).float().mean().item() <= max_frac


@pytest.mark.parametrize("n,hq,hkv", [(37, 16, 8), (1, 16, 8), (64, 4, 1), (513, 8, 2)])
def test_fused_pre_attention(lib, n, hq, hkv):
    from nano_vllm_learn_b200 import ops
    c = cases.make_pre_attention_case(seed=30 + n, n=n, hq=hq, hkv=hkv)
    cs = cases.rope_cache(c["rope_base"], c["max_position"])
    q_ref, k_ref, v_ref = orc.qk_norm_rope_ref(c["qkv"], c["positions"], c["q_weight"], c["k_weight"], c["eps"], cs,
                                               hq, hkv)
    kc = torch.randn(c["nblk"], 256, hkv, 128).to(torch.bfloat16)
    vc = torch.randn(c["nblk"], 256, hkv, 128).to(torch.bfloat16)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    orc.store_kvcache_ref(k_ref, v_ref, kc_ref, vc_ref, c["slots"])
    kc_d, vc_d = kc.to(DEV), vc.to(DEV)
    q, k = ops.qk_norm_rope_store(c["qkv"].to(DEV), c["positions"].to(DEV), c["q_weight"].to(DEV),
                                  c["k_weight"].to(DEV), c["eps"], cs.to(DEV), hq, hkv, 128, kc_d, vc_d,
                                  c["slots"].to(DEV), want_k=True)
    torch.cuda.synchronize()
    _ulp_check(q.view(n, hq, 128), q_ref)
    _ulp_check(k.view(n, hkv, 128), k_ref)
    assert torch.equal(vc_d.cpu().view(torch.int16), vc_ref.view(torch.int16))           # V: bit-exact
    _ulp_check(kc_d, kc_ref)
    # the cache holds exactly the k rows the kernel returned
    flat = kc_d.view(-1, hkv * 128)[c["slots"].long().to(DEV)]
    assert torch.equal(flat.view(torch.int16), k.view(torch.int16))
    # no cache bound (warmup): outputs only
    q2, k2 = ops.qk_norm_rope_store(c["qkv"].to(DEV), c["positions"].to(DEV), c["q_weight"].to(DEV),
                                    c["k_weight"].to(DEV), c["eps"], cs.to(DEV), hq, hkv, 128)
    assert torch.equal(q2, q) and torch.equal(k2, k)


def test_fused_pre_attention_vs_reference_fixture(lib):
    from nano_vllm_learn_b200 import ops
    c = cases.make_pre_attention_case()
    g = np.load(os.path.join(GOLD, "pre_attention.npz"))
    cs = cases.rope_cache(c["rope_base"], c["max_position"])
    q, k = ops.qk_norm_rope_store(c["qkv"].to(DEV), c["positions"].to(DEV), c["q_weight"].to(DEV),
                                  c["k_weight"].to(DEV), c["eps"], cs.to(DEV), c["hq"], c["hkv"], 128)
    _ulp_check(q, cases.from_bits(g["q"]))
    _ulp_check(k, cases.from_bits(g["k"]))


def test_attention_forward_fused_all_modes(lib):
    """Attention.forward_fused (integration edit A5) in its three modes — dense prefill, paged-prefix prefill,
    decode — against the oracle chain norm/rope -> store -> attention, using the reference's module shapes."""
    from types import SimpleNamespace
    from nano_vllm_learn_b200 import set_context, reset_context, get_context, attention_cuda
    from nano_vllm_learn_b200.attention_cuda import Attention
    attention_cuda.bind_context(get_context)
    hq, hkv, d, nblk = 8, 2, 128, 8
    scale = d ** -0.5
    g = torch.Generator().manual_seed(77)
    cs = cases.rope_cache(1000000.0, 4096)
    qw = (1 + 0.1 * torch.randn(d, generator=g)).to(torch.bfloat16)
    kw = (1 + 0.1 * torch.randn(d, generator=g)).to(torch.bfloat16)
    q_norm = SimpleNamespace(weight=qw.to(DEV), eps=1e-6)
    k_norm = SimpleNamespace(weight=kw.to(DEV), eps=1e-6)
    rot = SimpleNamespace(cos_sin_cache=cs.to(DEV))
    attn = Attention(hq, d, scale, hkv)
    kv = torch.zeros(2, nblk, 256, hkv, d, dtype=torch.bfloat16, device=DEV)
    attn.k_cache, attn.v_cache = kv[0], kv[1]
    kc = torch.zeros(nblk, 256, hkv, d, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)

    def run(qkv, positions):
        return attn.forward_fused(qkv.to(DEV), positions.to(DEV), q_norm, k_norm, rot)

    # (1) dense prefill of two sequences (300 and 40 tokens): blocks [5, 2] and [6]
    lens = [300, 40]
    n = sum(lens)
    qkv = (torch.randn(n, (hq + 2 * hkv) * d, generator=g) * 1.2).to(torch.bfloat16)
    pos = torch.cat([torch.arange(l) for l in lens])
    cu = torch.tensor([0, 300, 340], dtype=torch.int32)
    slots = torch.cat([torch.arange(5 * 256, 6 * 256), torch.arange(2 * 256, 2 * 256 + 44),
                       torch.arange(6 * 256, 6 * 256 + 40)]).to(torch.int32)
    set_context(True, cu.to(DEV), cu.to(DEV), 300, 300, slots.to(DEV), None, None)
    o = run(qkv, pos)
    qr, kr, vr = orc.qk_norm_rope_ref(qkv, pos, qw, kw, 1e-6, cs, hq, hkv)
    orc.store_kvcache_ref(kr, vr, kc, vc, slots)
    want = orc.varlen_prefill_ref(qr, kr, vr, cu, cu, scale)
    assert (o.view(n, hq, d).float().cpu() - want.float()).abs().max().item() <= 2e-2
    assert torch.equal(kv[1].cpu().view(torch.int16), vc.view(torch.int16))
    _ulp_check(kv[0].cpu(), kc)

    # (2) prefix-cache prefill: a third sequence re-uses block 5 (256 cached tokens) and adds 70 new ones in block 7
    qkv2 = (torch.randn(70, (hq + 2 * hkv) * d, generator=g) * 1.2).to(torch.bfloat16)
    pos2 = torch.arange(256, 326)
    cu_q = torch.tensor([0, 70], dtype=torch.int32)
    cu_k = torch.tensor([0, 326], dtype=torch.int32)
    slots2 = torch.arange(7 * 256, 7 * 256 + 70).to(torch.int32)
    tab2 = torch.tensor([[5, 7]], dtype=torch.int32)
    set_context(True, cu_q.to(DEV), cu_k.to(DEV), 70, 326, slots2.to(DEV), None, tab2.to(DEV))
    o2 = run(qkv2, pos2)
    qr2, kr2, vr2 = orc.qk_norm_rope_ref(qkv2, pos2, qw, kw, 1e-6, cs, hq, hkv)
    orc.store_kvcache_ref(kr2, vr2, kc, vc, slots2)
    want2 = orc.varlen_prefill_ref(qr2, kc, vc, cu_q, cu_k, scale, tab2)
    assert (o2.view(70, hq, d).float().cpu() - want2.float()).abs().max().item() <= 2e-2

    # (3) one decode step for the first two sequences
    qkv3 = (torch.randn(2, (hq + 2 * hkv) * d, generator=g) * 1.2).to(torch.bfloat16)
    pos3 = torch.tensor([300, 40])
    slots3 = torch.tensor([2 * 256 + 44, 6 * 256 + 40], dtype=torch.int32)
    ctx = torch.tensor([301, 41], dtype=torch.int32)
    tab3 = torch.tensor([[5, 2], [6, -1]], dtype=torch.int32)
    set_context(False, slot_mapping=slots3.to(DEV), context_lens=ctx.to(DEV), block_tables=tab3.to(DEV))
    o3 = run(qkv3, pos3)
    qr3, kr3, vr3 = orc.qk_norm_rope_ref(qkv3, pos3, qw, kw, 1e-6, cs, hq, hkv)
    orc.store_kvcache_ref(kr3, vr3, kc, vc, slots3)
    want3 = orc.paged_decode_ref(qr3, kc, vc, ctx, tab3, scale)
    assert (o3.view(2, hq, d).float().cpu() - want3.float()).abs().max().item() <= 2e-2
    assert torch.equal(kv[1].cpu().view(torch.int16), vc.view(torch.int16))
    reset_context()
