"""GPU (>= 2 devices): the reference ENGINE with tensor_parallel_size = 2 on the cuda backend -- spawned ranks, NCCL
process group, shm RPC, CUDA graphs, the head-sharded attention kernels (Hq/tp, Hkv/tp heads per rank) -- with and
without vocabulary-parallel sampling (INTEGRATION.md edit A8) and the fused embedding exchange (A9).  The model path is identical in both runs, so the greedy
token ids must be identical token for token: the 16-byte pair exchange returns exactly what dist.gather + torch.cat +
Sampler.forward on rank 0 return (nanovllm/layers/embed_head.py:66-70, engine/model_runner.py:311)."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tp2_engine_tokens_identical_with_vocab_parallel_sampling(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    if not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "nanovllm")):
        pytest.skip("baseline/_ref not present")
    out = str(tmp_path / "tokens.json")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check_tp_tokens.py"), "--tensor-parallel-size", "2",
                        "--layers", "2", "--steps", "24", "--arms", "nccl", "vp", "emb", "--port", "2470", "--out", out],
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, (p.stdout + p.stderr)[-1500:]
    rec = json.load(open(out))
    assert rec["vp"]["identical_to_nccl"] is True and rec["vp"]["sampler"].endswith("sampler_cuda")
    assert rec["emb"]["identical_to_nccl"] is True            # fused VocabParallelEmbedding (edit A9): a bit-exact copy
    assert len(rec["tokens"]["vp"]) == 6 and all(len(t) == 24 for t in rec["tokens"]["vp"])
    # with vocabulary-parallel sampling every rank holds the sampled ids on its device, and with seq_id in the
    # pickled Sequence state (runner_glue.install_sequence_identity) worker ranks recognise an unchanged batch: rank 1
    # must advance decode steps on the device like rank 0, not re-stage each one from the host.
    stats = {r["rank"]: r for r in rec["vp"]["glue_stats"]}
    assert set(stats) == {0, 1}
    assert stats[1]["fast_steps"] > 0 and stats[1]["fast_steps"] == stats[0]["fast_steps"], stats
    assert stats[1]["staged_steps"] == stats[0]["staged_steps"], stats
