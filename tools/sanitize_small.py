"""Smallest end-to-end pass over every kernel of the library, for `compute-sanitizer --tool memcheck`:
store, fused pre-attention, decode (single + split), both prefill kernels (dense + paged prefix)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import cases  # noqa: E402
from nano_vllm_learn_b200 import ops  # noqa: E402

dev = "cuda"


def cu(c):
    return {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in c.items()}


c = cu(cases.make_store_case(1, 37, 8, 2, dup_slot0=2, neg=2))
ops.store_kvcache(c["key"], c["value"], c["k_cache"], c["v_cache"], c["slot_mapping"])
p = cases.make_pre_attention_case()
cs = cases.rope_cache(p["rope_base"], p["max_position"]).to(dev)
kc = torch.zeros(p["nblk"], 256, p["hkv"], 128, dtype=torch.bfloat16, device=dev)
vc = torch.zeros_like(kc)
ops.qk_norm_rope_store(p["qkv"].to(dev), p["positions"].to(dev), p["q_weight"].to(dev), p["k_weight"].to(dev), p["eps"],
                       cs, p["hq"], p["hkv"], 128, kc, vc, p["slots"].to(dev))
d = cu(cases.make_decode_case("ragged_g4"))
for splits in (1, 3):
    ops.flash_attn_with_kvcache(d["q"], d["k_cache"], d["v_cache"], d["context_lens"], d["block_tables"],
                                softmax_scale=d["scale"], num_splits=splits)
for variant in (0, 1):
    ops.set_prefill_variant(variant)
    f = cu(cases.make_prefill_case("p_small"))
    ops.flash_attn_varlen_func(f["q"], f["k"], f["v"], f["max_seqlen_q"], f["cu_seqlens_q"], f["max_seqlen_k"],
                               f["cu_seqlens_k"], softmax_scale=f["scale"])
    x = cu(cases.make_prefix_case("px_shared512"))
    ops.flash_attn_varlen_func(x["q"], x["k_cache"], x["v_cache"], x["max_seqlen_q"], x["cu_seqlens_q"],
                               x["max_seqlen_k"], x["cu_seqlens_k"], softmax_scale=x["scale"],
                               block_table=x["block_tables"])
ops.set_prefill_variant(-1)
torch.cuda.synchronize()
print("sanitize_small: all kernels ran")
