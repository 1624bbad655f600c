"""CPU checks of the vocab-parallel embedding oracle (oracle/embedding_oracle.py) against the reference's own outputs
(tests/golden/vocab_embedding.npz: VocabParallelEmbedding.forward under gloo, one process per rank) -- SURVEY §8f
rank 2 (nanovllm/layers/embed_head.py:35-43)."""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import embedding_oracle as eo

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "vocab_embedding.npz"))


@pytest.mark.parametrize("name", list(cases.EMBEDDING_CASES))
def test_restatement_reproduces_reference(name):
    c = cases.make_embedding_case(name)
    assert cases.sha(c["x"], c["weight"]) == str(GOLD[f"{name}_sha"]), "fixture inputs differ"
    want = cases.from_bits(GOLD[f"{name}_y"])
    got = eo.vocab_embedding_ref(c["x"], c["weight"], c["world"])
    assert got.dtype == torch.bfloat16 and torch.equal(got, want)
    # what the fused kernel returns: the owner's row itself (the sum only adds zero rows to it)
    assert torch.equal(c["weight"][c["x"]], want)


def test_live_reference_world_matches_fixture():
    if eo._ref_root() is None:
        pytest.skip("reference source not present (neither /root/reference nor baseline/_ref)")
    name = "emb_tp4_ragged"
    c = cases.make_embedding_case(name)
    ys = eo.run_reference_world(c["x"], c["weight"], c["world"], port=29830)
    for y in ys:
        assert torch.equal(y, cases.from_bits(GOLD[f"{name}_y"]))
