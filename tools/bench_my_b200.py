"""bench_my on the B200 box: the reference engine end to end, `cuda` backend beside `flash`/`triton`.

Mirrors the measurement of the reference's bench_my.py (:28-76): seed(0), `randint(0, 10000)` prompts,
temperature 0, ignore_eos, output_len == input_len, tok/s = bs * output_len / wall time of
`llm.generate`, one engine per backend reused across the (bs, len) grid.  The engine is the
UNMODIFIED reference (baseline/_ref) plus the binding edits of INTEGRATION.md applied to a copy
(tools/make_integration.py); the `flash` column therefore runs the reference's stock flash-attn path.

    python tools/bench_my_b200.py --attn-backend flash cuda --batch-sizes 16 32 64 --input-lens 512 1024 2048
    python tools/bench_my_b200.py --tensor-parallel-size 2 ...
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import sys
import time
from random import randint, seed

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def setup_paths(args):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_integration
    dst = make_integration.DEFAULT_DST + f"_{os.getpid()}"
    make_integration.make(os.path.join(ROOT, "baseline", "_ref"), dst)
    sys.path.insert(0, dst)
    # spawned tensor-parallel workers import the engine and the backend by name
    os.environ["PYTHONPATH"] = os.pathsep.join([dst, ROOT, os.environ.get("PYTHONPATH", "")])
    if not args.model_path:
        import make_synthetic_model
        args.model_path = f"/tmp/nvllm_models/qwen3-{args.shape}" + (f"-L{args.layers}" if args.layers else "")
        if not os.path.exists(os.path.join(args.model_path, "model.safetensors")):
            make_synthetic_model.make(args.shape, args.model_path, args.layers)


def run_single(llm, SamplingParams, torch, input_len, num_seqs, output_len, shared_prefix=0, temperature=0.0):
    # shared_prefix > 0: BASELINE configs[2] — every prompt starts with the same tokens, so all but the first
    # sequence hit the prefix cache (cu_seqlens_k > cu_seqlens_q, block_tables set in prefill).  Keep it shorter
    # than the prompt: a fully cached prompt trips the reference's Lq == 0 quirk.
    shared = [randint(0, 10000) for _ in range(shared_prefix)]
    prompts = [shared + [randint(0, 10000) for _ in range(input_len - shared_prefix)] for _ in range(num_seqs)]
    sps = [SamplingParams(temperature=temperature, ignore_eos=True, max_tokens=output_len) for _ in range(num_seqs)]
    torch.cuda.synchronize()
    t0 = time.time()
    llm.generate(prompts, sps, use_tqdm=False)
    torch.cuda.synchronize()
    return (time.time() - t0) * 1000


def bench_backend(args, backend, port):
    import torch
    from nanovllm import LLM, SamplingParams
    os.environ["NVLLM_PORT"] = str(port)
    seed(0)
    res = {}
    llm = LLM(args.model_path, enforce_eager=args.enforce_eager, max_model_len=args.max_model_len,
              attn_backend=backend, tensor_parallel_size=args.tensor_parallel_size,
              max_num_batched_tokens=max(16384, args.max_model_len))
    try:
        llm.generate(["warmup"], SamplingParams(max_tokens=1))
        if not args.no_shape_warmup:
            # one untimed short run per batch size (same for every backend): torch.compile re-specialises the
            # reference's RMSNorm/RoPE/SiLU kernels on first sight of a shape, which otherwise lands in the
            # first timed point of each backend (visible as a 2x outlier in profiles/r01_bench_my_tp1_*).
            for bs in args.batch_sizes:
                run_single(llm, SamplingParams, torch, args.input_lens[0], bs, 8, args.shared_prefix, args.temperature)
        for bs in args.batch_sizes:
            for n in args.input_lens:
                out_len = args.output_len or n
                try:
                    ms = run_single(llm, SamplingParams, torch, n, bs, out_len, args.shared_prefix, args.temperature)
                    res[f"{bs}x{n}"] = dict(bs=bs, input_len=n, output_len=out_len, ms=round(ms, 1),
                                            tok_s=round(bs * out_len / (ms / 1000), 1))
                except torch.cuda.OutOfMemoryError:
                    res[f"{bs}x{n}"] = dict(bs=bs, input_len=n, output_len=out_len, ms=None, tok_s=None)
                    torch.cuda.empty_cache()
                print(json.dumps({"backend": backend, "tp": args.tensor_parallel_size, **res[f"{bs}x{n}"]}), flush=True)
    finally:
        try:
            llm.exit()
        except Exception:
            pass
        del llm
        gc.collect()
        torch.cuda.empty_cache()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model-path", default="")
    ap.add_argument("--shape", default="0.6b", choices=["0.6b", "8b"])
    ap.add_argument("--layers", type=int, default=None)
    ap.add_argument("--attn-backend", nargs="+", default=["flash", "cuda"], choices=["flash", "sdpa", "sdpa.math", "triton", "cuda"])
    ap.add_argument("--input-lens", type=int, nargs="+", default=[512, 1024, 2048])
    ap.add_argument("--batch-sizes", type=int, nargs="+", default=[16, 32, 64])
    ap.add_argument("--output-len", type=int, default=0, help="0 = same as input (bench_my)")
    ap.add_argument("--max-model-len", type=int, default=4096)
    ap.add_argument("--tensor-parallel-size", type=int, default=1)
    ap.add_argument("--enforce-eager", action="store_true")
    ap.add_argument("--shared-prefix", type=int, default=0, help="tokens shared by all prompts (prefix-cache hits)")
    ap.add_argument("--temperature", type=float, default=0.0)
    ap.add_argument("--no-shape-warmup", action="store_true", help="measure exactly like bench_my.py (first point includes compiles)")
    ap.add_argument("--port", type=int, default=2333, help="NCCL rendezvous port of the first engine (+1 per backend)")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    setup_paths(args)
    results = {}
    for i, backend in enumerate(args.attn_backend):
        results[backend] = bench_backend(args, backend, args.port + i)
    print("\nbench_my (tok/s, output_len=input_len), tp=%d" % args.tensor_parallel_size)
    for bs in args.batch_sizes:
        print(f"bs={bs:<4d}" + "".join(f"{b:>12s}" for b in args.attn_backend))
        for n in args.input_lens:
            row = f"{n:>7d}"
            for b in args.attn_backend:
                v = results[b].get(f"{bs}x{n}", {}).get("tok_s")
                row += f"{v:12.0f}" if v else "         OOM"
            print(row)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(dict(tp=args.tensor_parallel_size, shape=args.shape, layers=args.layers,
                           shared_prefix=args.shared_prefix, max_model_len=args.max_model_len,
                           results=results), f, indent=1)


if __name__ == "__main__":
    main()
