"""Generate tests/golden/rowparallel.npz by running the REFERENCE's own RowParallelLinear.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_rowparallel.py
What is frozen: for every case of tests/cases.ROWPARALLEL_CASES, the output of
nanovllm/layers/linear.py::RowParallelLinear.forward (:185-190) on every rank, one CPU process per rank under
torch.distributed/gloo (oracle/rowparallel_oracle.run_reference_world), with the reference's own weight_loader
sharding W.  Inputs are regenerated from seeds; their sha256 is stored.
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
from oracle import rowparallel_oracle as rpo  # noqa: E402


def main():
    out = {}
    for i, name in enumerate(cases.ROWPARALLEL_CASES):
        c = cases.make_rowparallel_case(name)
        ys = rpo.run_reference_world(c["x"], c["weight"], c["bias"], c["world"], port=29740 + i)
        for r in range(1, c["world"]):
            assert torch.equal(ys[0].view(torch.int16), ys[r].view(torch.int16)), "ranks disagree"
        out[f"{name}_y"] = cases.bf16_bits(ys[0])
        out[f"{name}_sha"] = cases.sha(c["x"], c["weight"], c["bias"])
        print(name, tuple(ys[0].shape))
    out["meta"] = json.dumps(dict(torch=torch.__version__, backend="gloo",
                                  reference="ztxdcyy/nano-vllm-learn @ /root/reference"))
    np.savez_compressed(os.path.join(HERE, "rowparallel.npz"), **out)


if __name__ == "__main__":
    main()
