"""Generate tests/golden/lm_head_sampler.npz by running the REFERENCE's own ParallelLMHead + Sampler.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_lm_head.py
What is frozen: for every case of tests/cases.LM_HEAD_CASES, the token ids rank 0 obtains from
nanovllm/layers/embed_head.py::ParallelLMHead.forward (:61-71, local shard GEMM + dist.gather + torch.cat) followed by
nanovllm/layers/sampler.py::Sampler.forward (:10-18) right after torch.manual_seed(noise_seed), one CPU process per rank
under torch.distributed/gloo (oracle/sampler_oracle.run_reference_lm_head_world), with the reference's own weight_loader
sharding the lm_head.  Inputs are regenerated from seeds; their sha256 is stored.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402
from oracle import sampler_oracle as so  # noqa: E402


def main():
    out = {}
    for i, name in enumerate(cases.LM_HEAD_CASES):
        c = cases.make_lm_head_case(name)
        tok, logits = so.run_reference_lm_head_world(c["x"], c["weight"], c["temperatures"], c["noise_seed"], c["world"],
                                                     is_prefill=c["cu_seqlens_q"] is not None,
                                                     cu_seqlens_q=c["cu_seqlens_q"], port=29770 + i)
        out[f"{name}_tokens"] = tok.numpy()
        out[f"{name}_sha"] = cases.sha(c["x"], c["weight"], c["temperatures"])
        print(name, tuple(logits.shape), tok.tolist())
    out["meta"] = json.dumps(dict(torch=torch.__version__, backend="gloo",
                                  reference="ztxdcyy/nano-vllm-learn @ /root/reference"))
    np.savez_compressed(os.path.join(HERE, "lm_head_sampler.npz"), **out)


if __name__ == "__main__":
    main()
