"""Token identity of the reference ENGINE under tensor parallelism across the optional fusions of the cuda backend.

    python tools/check_tp_tokens.py --tensor-parallel-size 2 [--layers 4] [--out tokens.json]

Runs the UNMODIFIED reference engine (baseline/_ref + the INTEGRATION.md edits applied to a /tmp copy) on the same
greedy requests once per arm, each arm in its own process:
    nccl      cuda backend, reference collectives: F.linear + dist.all_reduce, lm_head dist.gather + rank-0 sampler
    vp        + vocabulary-parallel sampling (edit A8) -- the model path is untouched, so the token ids must be
              IDENTICAL to `nccl` (logits are the same bits; only where the arg-max is taken changes)
    emb       + fused VocabParallelEmbedding (edit A9) -- a bit-exact copy of the owner's rows: IDENTICAL again
    fused     + fused RowParallelLinear all-reduce (edit A7) as well -- hidden states differ from NCCL's by one bf16
              rounding per layer, so tokens are compared as a match fraction of the first generated tokens
Prints one JSON line; exit code 0 iff `vp` and `emb` == `nccl` token for token.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ARMS = {
    "nccl": {"NVLLM_FUSED_ALLREDUCE": "0", "NVLLM_VOCAB_PARALLEL_SAMPLER": "0", "NVLLM_FUSED_EMBEDDING": "0"},
    "vp": {"NVLLM_FUSED_ALLREDUCE": "0", "NVLLM_VOCAB_PARALLEL_SAMPLER": "1", "NVLLM_FUSED_EMBEDDING": "0"},
    "emb": {"NVLLM_FUSED_ALLREDUCE": "0", "NVLLM_VOCAB_PARALLEL_SAMPLER": "1", "NVLLM_FUSED_EMBEDDING": "1"},
    "fused": {"NVLLM_FUSED_ALLREDUCE": "1", "NVLLM_VOCAB_PARALLEL_SAMPLER": "1", "NVLLM_FUSED_EMBEDDING": "1"},
}


def child(args):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_my_b200 as bm
    ns = argparse.Namespace(model_path="", shape="0.6b", layers=args.layers)
    bm.setup_paths(ns)
    import torch
    from nanovllm import LLM, SamplingParams
    os.environ["NVLLM_PORT"] = str(args.port)
    import glob
    import tempfile
    stats_dir = tempfile.mkdtemp(prefix="nvllm_glue_stats_")
    os.environ["NVLLM_GLUE_STATS_DIR"] = stats_dir     # inherited by the spawned worker ranks
    g = torch.Generator().manual_seed(11)
    prompts = [torch.randint(0, 10000, (n,), generator=g).tolist() for n in (300, 17, 64, 129, 255, 5)][:args.bs]
    llm = LLM(ns.model_path, enforce_eager=False, max_model_len=1024, max_num_seqs=16, attn_backend="cuda",
              tensor_parallel_size=args.tensor_parallel_size, max_num_batched_tokens=4096)
    try:
        sps = [SamplingParams(temperature=0.0, ignore_eos=True, max_tokens=args.steps) for _ in prompts]
        llm.generate([list(p) for p in prompts[:2]], sps[:2], use_tqdm=False)      # warm-up (compiles)
        torch.cuda.synchronize()
        t0 = time.time()
        out = llm.generate([list(p) for p in prompts], sps, use_tqdm=False)
        torch.cuda.synchronize()
        wall = time.time() - t0
        rec = {"tokens": [o["token_ids"] for o in out], "wall_s": round(wall, 3),
               "sampler": type(llm.model_runner.sampler).__module__}
    finally:
        llm.exit()
    time.sleep(0.5)
    # decode-path counters of every rank (runner_glue's exit hook): device-side steps / host-staged steps
    rec["glue_stats"] = sorted((json.load(open(p)) for p in glob.glob(os.path.join(stats_dir, "glue_rank*.json"))),
                               key=lambda r: r["rank"])
    with open(args.child_out, "w") as f:
        json.dump(rec, f)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tensor-parallel-size", type=int, default=2)
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--steps", type=int, default=48)
    ap.add_argument("--bs", type=int, default=6)
    ap.add_argument("--arms", nargs="+", default=list(ARMS))
    ap.add_argument("--port", type=int, default=2460)
    ap.add_argument("--timeout", type=int, default=240)
    ap.add_argument("--out", default="")
    ap.add_argument("--child-out", default="")
    args = ap.parse_args()
    if args.child_out:
        return child(args)
    res = {}
    for i, arm in enumerate(args.arms):
        path = f"/tmp/nvllm_tp_tokens_{os.getpid()}_{arm}.json"
        env = dict(os.environ, **ARMS[arm])
        cmd = [sys.executable, os.path.abspath(__file__), "--tensor-parallel-size", str(args.tensor_parallel_size),
               "--layers", str(args.layers), "--steps", str(args.steps), "--bs", str(args.bs), "--port", str(args.port + i),
               "--child-out", path]
        try:
            p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=args.timeout)
            res[arm] = json.load(open(path)) if p.returncode == 0 and os.path.exists(path) else \
                {"failed": f"rc={p.returncode}", "log_tail": (p.stdout + p.stderr)[-600:]}
        except subprocess.TimeoutExpired:
            res[arm] = {"failed": "timeout"}
        time.sleep(2.0)
    line = {"tp": args.tensor_parallel_size, "layers": args.layers, "steps": args.steps, "bs": args.bs}
    base = res.get("nccl", {}).get("tokens")
    for arm, r in res.items():
        row = {k: v for k, v in r.items() if k != "tokens"}
        if base is not None and "tokens" in r:
            row["identical_to_nccl"] = r["tokens"] == base
            tot = sum(len(t) for t in base)
            row["match_fraction"] = round(sum(a == b for x, y in zip(r["tokens"], base) for a, b in zip(x, y)) / tot, 4)
            row["first_token_identical"] = [t[0] for t in r["tokens"]] == [t[0] for t in base]
        line[arm] = row
    print(json.dumps(line), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(dict(line, tokens={a: r.get("tokens") for a, r in res.items()}), f)
    ok = base is not None and all(res[a].get("tokens") == base for a in ("vp", "emb") if a in res)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
