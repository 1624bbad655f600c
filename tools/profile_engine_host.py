"""cProfile of the engine's decode loop on the box (host-side cost of scheduler / glue / sampler)."""
import cProfile
import io
import os
import pstats
import sys
from random import randint, seed

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench_my_b200 as bm  # noqa: E402


class A:
    model_path = ""
    shape = "0.6b"
    layers = None


def main():
    backend = sys.argv[1] if len(sys.argv) > 1 else "cuda"
    bs = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    args = A()
    bm.setup_paths(args)
    import torch
    from nanovllm import LLM, SamplingParams
    llm = LLM(args.model_path, enforce_eager=False, max_model_len=4096, attn_backend=backend)
    seed(0)
    llm.generate(["warmup"], SamplingParams(max_tokens=1))
    prompts = [[randint(0, 10000) for _ in range(512)] for _ in range(bs)]
    sp = [SamplingParams(temperature=0.0, ignore_eos=True, max_tokens=64) for _ in range(bs)]
    llm.generate(prompts, sp, use_tqdm=False)   # compile warm-up
    # fresh prompts: re-submitting fully cached 512-token prompts hits the reference's Lq == 0 quirk
    prompts = [[randint(0, 10000) for _ in range(512)] for _ in range(bs)]
    sp = [SamplingParams(temperature=0.0, ignore_eos=True, max_tokens=400) for _ in range(bs)]
    pr = cProfile.Profile()
    torch.cuda.synchronize()
    pr.enable()
    llm.generate(prompts, sp, use_tqdm=False)
    torch.cuda.synchronize()
    pr.disable()
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
    print(s.getvalue())
    llm.exit()


if __name__ == "__main__":
    main()
