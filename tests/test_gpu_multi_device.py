"""GPU (needs >= 2 devices; skipped on a 1-GPU box): one PROCESS driving two GPUs, the second one first.

VERDICT r1 weak #7: launch state that is per device (cudaFuncSetAttribute(MaxDynamicSharedMemorySize), the SM count)
used to be cached per process, so the first >= 48 KB launch on a second device of the same process would fail.  A fresh
interpreter runs store + split-KV decode + tcgen05 prefill + sampler on cuda:1 BEFORE cuda:0 touches the library, and
checks both against the CPU oracle."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import os, sys
ROOT = sys.argv[1]
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import cases
from oracle import paged_attention_oracle as orc
from nano_vllm_learn_b200 import ops

def run(dev):
    torch.cuda.set_device(dev)
    d = torch.device("cuda", dev)
    # paged decode, single split and split-KV (dynamic shared memory > 48 KB on both variants)
    name = next(iter(cases.DECODE_CASES))
    c = cases.make_decode_case(name)
    want = orc.paged_decode_ref(c["q"], c["k_cache"], c["v_cache"], c["context_lens"], c["block_tables"], c["scale"])
    for splits in (1, 3):
        got = ops.flash_attn_with_kvcache(c["q"].to(d), c["k_cache"].to(d), c["v_cache"].to(d), c["context_lens"].to(d),
                                          c["block_tables"].to(d), softmax_scale=c["scale"], num_splits=splits)
        err = (got.float().cpu() - want.float()).abs().max().item()
        assert err <= 2e-2, ("decode", dev, splits, err)
    # varlen prefill on the tcgen05 kernel and on the mma.sync kernel
    pname = next(iter(cases.PREFILL_CASES))
    p = cases.make_prefill_case(pname)
    wantp = orc.varlen_prefill_ref(p["q"], p["k"], p["v"], p["cu_seqlens_q"], p["cu_seqlens_k"], p["scale"])
    for variant, path in ((1, 1), (0, 0)):
        ops.set_prefill_variant(variant)
        gotp = ops.flash_attn_varlen_func(p["q"].to(d), p["k"].to(d), p["v"].to(d), p["max_seqlen_q"],
                                          p["cu_seqlens_q"].to(d), p["max_seqlen_k"], p["cu_seqlens_k"].to(d),
                                          softmax_scale=p["scale"], causal=True)
        assert ops.prefill_last_path() == path
        err = (gotp.float().cpu() - wantp.float()).abs().max().item()
        assert err <= 2e-2, ("prefill", dev, variant, err)
    ops.set_prefill_variant(-1)
    # greedy sampler
    logits = torch.randn(5, 4096, device=d).to(torch.bfloat16)
    tok = ops.sample_tokens(logits, torch.zeros(5, dtype=torch.float32, device=d))
    assert torch.equal(tok.cpu(), logits.float().argmax(-1).cpu())
    torch.cuda.synchronize(dev)
    assert ops.prefill_error() == 0

run(1)      # the second device FIRST
run(0)
run(1)
print("multi-device ok")
"""


def test_second_device_first_in_one_process(lib, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "multi_device.py"
    script.write_text(SCRIPT)
    proc = subprocess.run([sys.executable, str(script), ROOT], capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-4000:]
    assert "multi-device ok" in proc.stdout
