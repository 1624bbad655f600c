"""CPU checks of the vocabulary-parallel lm_head + sampler oracle (oracle/sampler_oracle.py) against the reference's own
outputs (tests/golden/lm_head_sampler.npz: ParallelLMHead.forward + Sampler.forward run under gloo, one process per
rank) -- SURVEY §8f rank 4, second half (nanovllm/layers/embed_head.py:61-71, engine/model_runner.py:311)."""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import sampler_oracle as so

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "lm_head_sampler.npz"))


def _shards(c):
    last = None if c["cu_seqlens_q"] is None else (c["cu_seqlens_q"][1:] - 1).long()
    return so.lm_head_shards_ref(c["x"], c["weight"], c["world"], last)


@pytest.mark.parametrize("name", list(cases.LM_HEAD_CASES))
def test_gather_restatement_reproduces_reference_tokens(name):
    c = cases.make_lm_head_case(name)
    assert cases.sha(c["x"], c["weight"], c["temperatures"]) == str(GOLD[f"{name}_sha"]), "fixture inputs differ"
    logits = so.lm_head_gather_ref(_shards(c))
    noise = so.reference_noise(logits.shape, c["noise_seed"])
    got = so.sampler_ref(logits, c["temperatures"], noise)
    assert got.tolist() == GOLD[f"{name}_tokens"].tolist()


@pytest.mark.parametrize("name", list(cases.LM_HEAD_CASES))
def test_per_shard_pairs_merge_to_the_same_tokens(name):
    """The kernel's algorithm: one (key, global index) pair per rank and row, best pair wins -- same tokens as
    gathering the logits first."""
    c = cases.make_lm_head_case(name)
    shards = _shards(c)
    noise = so.reference_noise((shards[0].shape[0], c["weight"].shape[0]), c["noise_seed"])
    got = so.sharded_sampler_ref(shards, c["temperatures"], noise)
    assert got.tolist() == GOLD[f"{name}_tokens"].tolist()


def test_cross_shard_tie_takes_the_lowest_global_index():
    c = cases.make_lm_head_case("vp_decode_tp2")
    logits = so.lm_head_gather_ref(_shards(c))
    vocab = logits.shape[1]
    top = int(GOLD["vp_decode_tp2_tokens"][0])
    assert top < vocab // 2 and float(logits[0, top]) == float(logits[0, vocab - 3])   # the tie spans both shards


def test_live_reference_world_matches_fixture():
    if so._ref_root() is None:
        pytest.skip("reference source not present (neither /root/reference nor baseline/_ref)")
    name = "vp_prefill_tp2"
    c = cases.make_lm_head_case(name)
    tok, logits = so.run_reference_lm_head_world(c["x"], c["weight"], c["temperatures"], c["noise_seed"], c["world"],
                                                 is_prefill=True, cu_seqlens_q=c["cu_seqlens_q"], port=29790)
    assert tok.tolist() == GOLD[f"{name}_tokens"].tolist()
    assert torch.equal(logits, so.lm_head_gather_ref(_shards(c)).float())
