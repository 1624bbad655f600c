"""Generate tests/golden/vocab_embedding.npz by running the REFERENCE's own VocabParallelEmbedding.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_embedding.py
What is frozen: for every case of tests/cases.EMBEDDING_CASES, the output of
nanovllm/layers/embed_head.py::VocabParallelEmbedding.forward (:35-43) on every rank, one CPU process per rank under
torch.distributed/gloo (oracle/embedding_oracle.run_reference_world), with the reference's own weight_loader sharding
the table.  Inputs are regenerated from seeds; their sha256 is stored.
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
from oracle import embedding_oracle as eo  # noqa: E402


def main():
    out = {}
    for i, name in enumerate(cases.EMBEDDING_CASES):
        c = cases.make_embedding_case(name)
        ys = eo.run_reference_world(c["x"], c["weight"], c["world"], port=29810 + i)
        for r in range(1, c["world"]):
            assert torch.equal(ys[0], ys[r]), "ranks disagree"
        out[f"{name}_y"] = cases.bf16_bits(ys[0])
        out[f"{name}_sha"] = cases.sha(c["x"], c["weight"])
        print(name, tuple(ys[0].shape))
    out["meta"] = json.dumps(dict(torch=torch.__version__, backend="gloo",
                                  reference="ztxdcyy/nano-vllm-learn @ /root/reference"))
    np.savez_compressed(os.path.join(HERE, "vocab_embedding.npz"), **out)


if __name__ == "__main__":
    main()
