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
    for p, sp in zip(prompts, [SamplingParams(temperature=0.0, ignore_eos=True, max_tokens=128) for _ in range(bs)]):
        llm.add_request(p, sp)
    for _ in range(12):          # prefill + a few decode steps outside the profile
        llm.step()
    torch.cuda.synchronize()
    N = 40
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(N):
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
    # the host's share of the idle time: from the end of a step's last kernel (the sampler; any D2H copy of the ids
    # follows it) to the first kernel of the next step's graph -- scheduler, glue, launch.  The rest of the idle time
    # are the gaps between the ~260 small kernels INSIDE the captured graph, which no host code can shorten.
    gaps = []
    for i, e in enumerate(evs[:-1]):
        if "sample_tokens_kernel" in e.name or "ArgMaxOps" in e.name:
            nxt = next((x for x in evs[i + 1:] if not x.name.startswith("Memcpy") and not x.name.startswith("Memset")
                        and "sample_tokens" not in x.name and "ArgMaxOps" not in x.name and "where" not in x.name), None)
            if nxt is not None:
                gaps.append(nxt.time_range.start - e.time_range.end)
    if gaps:
        gaps.sort()
        print(f"between-step gap (sampler end -> next step's first kernel): median {gaps[len(gaps) // 2]:.1f} us, "
              f"mean {sum(gaps) / len(gaps):.1f} us over {len(gaps)} steps")
    glue = getattr(llm.model_runner, "_nvllm_glue", None)
    print(f"backend={backend} bs={bs} prompt={plen}: {N} decode steps, span {span/N:.1f} us/step, "
          f"kernel-busy {busy/N:.1f} us/step, {len(evs)/N:.0f} kernels/step, "
          f"idle {100 * (span - busy) / span:.1f} % of the step; glue decode steps on device / staged: "
          f"{getattr(glue, 'fast_steps', None)} / {getattr(glue, 'staged_steps', None)}")
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:22]:
        print(f"{t/N:9.1f} us/step  n/step={n/N:6.1f}  avg={t/n:7.2f} us  {name}")
    llm.exit()


if __name__ == "__main__":
    main()
