"""Fused token sampler (csrc/sampler.cu) vs the reference Sampler module on the same GPU.

    python tools/bench_sampler.py [--out gpurun_out/sampler.jsonl]

Per batch size: CUDA-event time of `Sampler.forward` (reference class from baseline/_ref, eager torch) and
of `nvllm_sample_tokens`, greedy (temperature 0, bench_my's setting) and sampled (temperature 0.8), Qwen3
vocabulary, bf16 logits.  Logits tensors rotate through a pool larger than L2.  Algorithmic bytes per launch
= bs * vocab * 2 (each logit read once) + bs * 12.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, pool, iters):
    for i in range(5):
        fn(pool[i % len(pool)])
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(iters):
        fn(pool[i % len(pool)])
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) * 1e3 / iters   # us


def load_reference_sampler():
    """The unmodified reference class from baseline/_ref/nanovllm/layers/sampler.py (file import: the nanovllm
    package itself pulls flash-attn/triton)."""
    import importlib.util
    path = os.path.join(ROOT, "baseline", "_ref", "nanovllm", "layers", "sampler.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("_ref_sampler", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.Sampler


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--vocab", type=int, default=151936)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--dist", choices=["normal3", "peaked"], default="normal3")
    ap.add_argument("--bs", type=int, nargs="*", default=[1, 8, 16, 64, 256, 512])
    args = ap.parse_args()
    from nano_vllm_learn_b200.sampler_cuda import Sampler
    ref_cls = load_reference_sampler()
    ref = ref_cls() if ref_cls is not None else None
    mine = Sampler()
    dev = "cuda"
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs", 6551.0))
    lines = []
    for bs in args.bs:
        nbytes = bs * args.vocab * 2
        n_pool = max(2, min(64, (256 << 20) // nbytes + 1))
        g = torch.Generator(device=dev).manual_seed(bs)
        pool = [(torch.randn(bs, args.vocab, generator=g, device=dev) * 3).to(torch.bfloat16) for _ in range(n_pool)]
        if args.dist == "peaked":
            # closer to a trained model's logits: a narrow bulk and a few dozen tokens far above it
            for x in pool:
                x.mul_(0.5)
                idx = torch.randint(0, args.vocab, (bs, 32), generator=g, device=dev)
                x.scatter_(1, idx, (14 + 8 * torch.rand(bs, 32, generator=g, device=dev)).to(torch.bfloat16))
        for temp in (0.0, 0.8):
            t = torch.full((bs,), temp, device=dev)
            us_mine = timed(lambda x: mine(x, t), pool, args.iters)
            us_ref = timed(lambda x: ref(x, t), pool, args.iters) if ref is not None else None
            if temp == 0.0 and ref is not None:
                assert torch.equal(mine(pool[0], t), ref(pool[0].clone(), t))
            line = dict(bs=bs, vocab=args.vocab, dist=args.dist, temperature=temp, fused_us=round(us_mine, 2),
                        reference_us=None if us_ref is None else round(us_ref, 2),
                        speedup=None if us_ref is None else round(us_ref / us_mine, 2),
                        algorithmic_bytes=nbytes + bs * 12,
                        fused_gbs=round((nbytes + bs * 12) / us_mine / 1e3, 1),
                        frac_of_measured_hbm=round((nbytes + bs * 12) / us_mine / 1e3 / hbm, 3))
            print(json.dumps(line), flush=True)
            lines.append(line)
    if args.out:
        with open(args.out, "w") as f:
            for line in lines:
                f.write(json.dumps(line) + "\n")


if __name__ == "__main__":
    main()
