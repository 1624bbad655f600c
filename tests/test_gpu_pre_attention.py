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
