"""Kernel-level timeline of the engine's captured decode graph (torch.profiler / CUPTI): which kernels a
decode step launches, their durations and the idle gaps between them."""
import os
import sys
from collections import defaultdict
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
    plen = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
    args = A()
    bm.setup_paths(args)
    import torch
    from torch.profiler import ProfilerActivity, profile
    from nanovllm import LLM, SamplingParams
    llm = LLM(args.model_path, enforce_eager=False, max_model_len=4096, attn_backend=backend)
    seed(0)
    llm.generate(["warmup"], SamplingParams(max_tokens=1))
    prompts = [[randint(0, 10000) for _ in range(plen)] for _ in range(bs)]
    for p, sp in zip(prompts, [SamplingParams(temperature=0.0, ignore_eos=True, max_tokens=64) for _ in range(bs)]):
        llm.add_request(p, sp)
    for _ in range(12):          # prefill + a few decode steps outside the profile
        llm.step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(8):
            llm.step()
        torch.cuda.synchronize()
    evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA),
                 key=lambda e: e.time_range.start)
    agg = defaultdict(lambda: [0, 0.0])
    busy = 0.0
    for e in evs:
        d = e.time_range.end - e.time_range.start
        agg[e.name[:90]][0] += 1
        agg[e.name[:90]][1] += d
        busy += d
    span = evs[-1].time_range.end - evs[0].time_range.start
    print(f"backend={backend} bs={bs} prompt={plen}: 8 decode steps, span {span/8:.1f} us/step, "
          f"kernel-busy {busy/8:.1f} us/step, {len(evs)/8:.0f} kernels/step")
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:22]:
        print(f"{t/8:9.1f} us/step  n/step={n/8:6.1f}  avg={t/n:7.2f} us  {name}")
    llm.exit()


if __name__ == "__main__":
    main()
