"""Generate tests/golden/*.npz|json by running the REFERENCE's own CPU-runnable code.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
What is frozen, and from which reference function:
  decode_*.npz   nanovllm/layers/attention_sdpa.py::flash_attn_with_kvcache (:122-182) on CPU, SDPA MATH
  prefill_*.npz  nanovllm/layers/attention_sdpa.py::flash_attn_varlen_func (:65-119) on CPU, no prefix
  gather.npz     nanovllm/layers/attention_sdpa.py::_gather_cache_tokens (:185-205)
  glue.json      nanovllm/engine/model_runner.py::prepare_prefill/prepare_decode (:171-269) driven by
                 the reference BlockManager/Sequence (block_manager.py, sequence.py), via
                 oracle/runner_glue_oracle.ReferenceGlue
Versions recorded in each file.  Inputs come from tests/cases.py seeds; their sha256 is stored.
"""
import json
import os
import random
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference"
sys.path.insert(0, REF)

import cases  # noqa: E402
from oracle.runner_glue_oracle import ReferenceGlue  # noqa: E402

from nanovllm.layers import attention_sdpa as ref_sdpa  # noqa: E402
from torch.nn.attention import SDPBackend  # noqa: E402

META = dict(torch=torch.__version__, reference="ztxdcyy/nano-vllm-learn @ /root/reference")


def gen_decode():
    for name in cases.DECODE_CASES:
        c = cases.make_decode_case(name)
        with torch.no_grad():
            o = ref_sdpa.flash_attn_with_kvcache(c["q"].unsqueeze(1), c["k_cache"], c["v_cache"],
                                                 cache_seqlens=c["context_lens"], block_table=c["block_tables"],
                                                 causal=False, sdpa_backend=SDPBackend.MATH)
        o = o.squeeze(1)
        np.savez_compressed(os.path.join(HERE, f"decode_{name}.npz"), out=cases.bf16_bits(o),
                            inputs_sha=cases.sha(c["q"], c["k_cache"], c["v_cache"], c["context_lens"],
                                                 c["block_tables"]),
                            meta=json.dumps(META))
        print("decode", name, tuple(o.shape))


def gen_prefill():
    for name in cases.PREFILL_CASES:
        c = cases.make_prefill_case(name)
        with torch.no_grad():
            o = ref_sdpa.flash_attn_varlen_func(c["q"], c["k"], c["v"], max_seqlen_q=c["max_seqlen_q"],
                                                cu_seqlens_q=c["cu_seqlens_q"], max_seqlen_k=c["max_seqlen_k"],
                                                cu_seqlens_k=c["cu_seqlens_k"], causal=True,
                                                sdpa_backend=SDPBackend.MATH, block_table=None)
        np.savez_compressed(os.path.join(HERE, f"prefill_{name}.npz"), out=cases.bf16_bits(o),
                            inputs_sha=cases.sha(c["q"], c["k"], c["v"], c["cu_seqlens_q"]),
                            meta=json.dumps(META))
        print("prefill", name, tuple(o.shape))


def gen_gather():
    c = cases.make_prefix_case("px_shared512")
    outs = {}
    lk = (c["cu_seqlens_k"][1:] - c["cu_seqlens_k"][:-1]).tolist()
    for i, n in enumerate(lk):
        outs[f"k{i}_sha"] = cases.sha(ref_sdpa._gather_cache_tokens(c["k_cache"], c["block_tables"][i], n))
    np.savez_compressed(os.path.join(HERE, "gather.npz"), inputs_sha=cases.sha(c["k_cache"], c["block_tables"]),
                        meta=json.dumps(META), **outs)
    print("gather", lk)


def _dump(d):
    return {k: (v.tolist() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}


def _dtypes(d):
    return {k: str(v.dtype) for k, v in d.items() if isinstance(v, torch.Tensor)}


def gen_glue():
    g = ReferenceGlue(REF)
    Seq, BM = g.sequence.Sequence, g.block_manager.BlockManager
    traces = []

    # (1) two sequences sharing a 512-token prefix, second one hits the first one's blocks in the same batch
    random.seed(0)
    shared = [random.randint(0, 10000) for _ in range(512)]
    prompts = [shared + [random.randint(0, 10000) for _ in range(300)],
               shared + [random.randint(0, 10000) for _ in range(512)]]
    traces.append(dict(name="shared_prefix_512", num_blocks=64, prompts=prompts, decode_steps=[1, 2, 213, 214, 215]))

    # (2) short prompts around block boundaries (the may_append paths incl. the hash-compensation branch)
    random.seed(1)
    lens = [1, 2, 3, 7, 255, 256, 257, 258, 511, 512, 513]
    prompts = [[random.randint(0, 10000) for _ in range(n)] for n in lens]
    traces.append(dict(name="short_prompts", num_blocks=64, prompts=prompts,
                       decode_steps=[1, 2, 3, 254, 255, 256, 257, 258, 300]))

    # (3) a fully cached prompt of exactly 2 blocks -> seqlen_q == 0 for the second sequence
    random.seed(2)
    p = [random.randint(0, 10000) for _ in range(512)]
    traces.append(dict(name="fully_cached", num_blocks=16, prompts=[p + [1, 2, 3], p], decode_steps=[1, 2]))

    out = []
    for tr in traces:
        bm = BM(tr["num_blocks"], 256)
        seqs = [Seq(list(p)) for p in tr["prompts"]]
        for s in seqs:
            bm.allocate(s)
        rec = dict(name=tr["name"], num_blocks=tr["num_blocks"], prompts=tr["prompts"], steps=[])
        pf = g.prepare_prefill(seqs)
        rec["prefill"] = _dump(pf)
        rec["prefill_dtypes"] = _dtypes(pf)
        rec["prefill_state"] = [dict(block_table=list(s.block_table), num_cached_tokens=s.num_cached_tokens)
                                for s in seqs]
        last = max(tr["decode_steps"])
        for step in range(1, last + 1):
            for i, s in enumerate(seqs):
                s.append_token((step * 31 + i * 7) % 10000)   # the token sampled by the previous step
                bm.may_append(s)
            if step in tr["decode_steps"]:
                dc = g.prepare_decode(seqs)
                rec["steps"].append(dict(step=step, decode=_dump(dc), dtypes=_dtypes(dc),
                                         state=[dict(block_table=list(s.block_table), num_tokens=len(s),
                                                     last_token=s.last_token) for s in seqs]))
        out.append(rec)

    # (4) forced hash == -1 entering may_append's compensation branch (block_manager.py:138-146)
    bm = BM(8, 256)
    random.seed(3)
    s = Seq([random.randint(0, 10000) for _ in range(256)])
    bm.allocate(s)
    bm.blocks[s.block_table[-1]].hash = -1      # force the state the compensation handles
    s.append_token(42)
    bm.may_append(s)
    dc = g.prepare_decode([s])
    out.append(dict(name="forced_hash_compensation", decode=_dump(dc), dtypes=_dtypes(dc),
                    state=dict(block_table=list(s.block_table), num_tokens=len(s), last_token=s.last_token,
                               recomputed_hash_is_minus1=bm.blocks[s.block_table[0]].hash == -1)))
    with open(os.path.join(HERE, "glue.json"), "w") as f:
        json.dump(dict(meta=META, traces=out), f)
    print("glue", [t["name"] for t in out])


def gen_pre_attention():
    """q/k RMSNorm + RoPE through the reference's own modules, eager (TORCHDYNAMO_DISABLE=1 so that
    @torch.compile does not need a GPU/compiler): layernorm.py RMSNorm, rotary_embedding.py RotaryEmbedding."""
    assert os.environ.get("TORCHDYNAMO_DISABLE") == "1", "run with TORCHDYNAMO_DISABLE=1"
    from nanovllm.layers.layernorm import RMSNorm
    from nanovllm.layers.rotary_embedding import RotaryEmbedding
    c = cases.make_pre_attention_case()
    hq, hkv, d = c["hq"], c["hkv"], cases.D
    qn, kn = RMSNorm(d, c["eps"]).to(torch.bfloat16), RMSNorm(d, c["eps"]).to(torch.bfloat16)
    qn.weight.data, kn.weight.data = c["q_weight"].clone(), c["k_weight"].clone()
    rot = RotaryEmbedding(d, d, c["max_position"], c["rope_base"])
    assert torch.equal(rot.cos_sin_cache, cases.rope_cache(c["rope_base"], c["max_position"]))
    q, k, v = c["qkv"].split([hq * d, hkv * d, hkv * d], dim=-1)
    with torch.no_grad():
        qr = qn(q.view(-1, hq, d)).view(q.shape)
        kr = kn(k.view(-1, hkv, d)).view(k.shape)
        qr, kr = rot(c["positions"], qr, kr)
    np.savez_compressed(os.path.join(HERE, "pre_attention.npz"), q=cases.bf16_bits(qr), k=cases.bf16_bits(kr),
                        inputs_sha=cases.sha(c["qkv"], c["positions"], c["q_weight"], c["k_weight"]),
                        meta=json.dumps(META))
    print("pre_attention", tuple(qr.shape), tuple(kr.shape))


def gen_sampler():
    """nanovllm/layers/sampler.py::Sampler.forward, the reference class itself on CPU; its exponential_ draw is
    the first use of torch's CPU generator after torch.manual_seed(noise_seed)."""
    from oracle.sampler_oracle import load_reference_sampler
    ref = load_reference_sampler((REF,))()
    outs = {}
    for name in cases.SAMPLER_CASES:
        c = cases.make_sampler_case(name)
        torch.manual_seed(c["noise_seed"])
        with torch.no_grad():
            tok = ref(c["logits"].clone(), c["temperatures"])
        outs[f"{name}_tokens"] = tok.numpy()
        outs[f"{name}_sha"] = cases.sha(c["logits"], c["temperatures"])
        print("sampler", name, tok.tolist())
    np.savez_compressed(os.path.join(HERE, "sampler.npz"), meta=json.dumps(META), **outs)


if __name__ == "__main__":
    torch.set_num_threads(8)
    if "--sampler-only" in sys.argv:
        gen_sampler()
        sys.exit(0)
    if "--pre-attention-only" in sys.argv:
        gen_pre_attention()
        sys.exit(0)
    gen_decode()
    gen_prefill()
    gen_gather()
    gen_glue()
    gen_pre_attention()
    gen_sampler()
