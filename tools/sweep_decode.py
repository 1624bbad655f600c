"""Decode-kernel tuning sweep on one GPU: variants x split counts x shapes, CUDA-event timed with an
L2 flush between launches.  Prints one JSON line per point (collected under gpurun_out/)."""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from nano_vllm_learn_b200 import _lib, ops  # noqa: E402

D, BLOCK = 128, 256


def make(bs, ctx, hq, hkv, dev, width=None, seed=0):
    g = torch.Generator(device=dev).manual_seed(seed)
    pages = (ctx + BLOCK - 1) // BLOCK
    nblk = bs * pages
    kc = torch.empty(nblk, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
    vc = torch.empty(nblk, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
    q = torch.empty(bs, hq, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
    tab = torch.randperm(nblk, generator=g, device=dev).to(torch.int32).view(bs, pages)
    if width and width > pages:
        wide = torch.zeros(bs, width, dtype=torch.int32, device=dev)
        wide[:, :pages] = tab
        tab = wide
    lens = torch.full((bs,), ctx, dtype=torch.int32, device=dev)
    nbytes = bs * ctx * hkv * D * 4 + 2 * bs * hq * D * 2 + 4 * bs + 4 * bs * pages
    return q, kc, vc, lens, tab, nbytes


def share_table(sets):
    """One block table and context_lens for every cache copy: the layers of a model share the step's table while each
    owns its cache, so the metadata is L2-resident when the K/V bytes are not."""
    t, lens = sets[0][4], sets[0][3]
    return [(q, kc, vc, lens, t, nb) for q, kc, vc, _, _, nb in sets]


def timeit(fn, flush, iters=12):
    """fn(i) runs on the i-th copy of the inputs; copies rotate so the working set exceeds L2 and no
    dirty-line flush pollutes the timed launch."""
    for i in range(3):
        fn(i)
    ts = []
    for i in range(iters):
        torch.cuda._sleep(300_000)   # keep the GPU busy while the host queues the launch: host latency is not timed
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        fn(i)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts), min(ts)


def timeit_graph(fn, ncopy, reps=4, iters=6):
    """Per-launch device time inside a CUDA graph of ncopy*reps back-to-back launches rotating over the input copies
    (how the engine runs the kernel: L layers per replay, each on its own cache).  No host launch latency, no event
    quantisation per launch; adjacent launches may overlap tail and ramp exactly as they do in the engine."""
    n = ncopy * reps
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for i in range(ncopy):
            fn(i)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(n):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3 / n)
    return statistics.median(ts), min(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="c2")
    ap.add_argument("--variants", default="all")
    ap.add_argument("--flash", action="store_true", help="also time flash-attn's kernel on the same inputs")
    ap.add_argument("--balanced", default="", help="comma list of ctas:min_chunks[:warps] for the balanced grid, e.g. 0:64,0:16:8,-296:64:4")
    ap.add_argument("--clusters", default="0", help="comma list of cluster sizes of the split-KV grid (0 auto, 1 off, 2/4/8/16)")
    ap.add_argument("--splits", default="", help="override the per-shape split list of the fixed grid (comma list; 'none' skips it)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    lib = _lib.load()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    shapes = {
        # name: (bs, ctx, hq, hkv, table width or None, split list)
        "c1": [(1, 512, 16, 8, 16, [0, 1, 2, 4, 8])],
        "c2": [(bs, ctx, 16, 8, 16, [0, 1, 2, 4, 8, 16]) for bs in (16, 32, 64) for ctx in (1024, 2048, 4096)],
        "c4": [(128, 4096, 32 // tp, 8 // tp, 16, [0, 1, 2, 4, 8]) for tp in (1, 2, 4, 8)],
        "c5": [(bs, 32768, 32 // tp, 8 // tp, 160, [0, 8, 16, 32, 64]) for bs in (1, 4) for tp in (1, 8)],
        "c5all": [(bs, 32768, 32 // tp, 8 // tp, 160, [0, 16]) for bs in (1, 2, 4) for tp in (1, 8)],
        "head": [(64, 4096, 16, 8, 16, [0, 1])],
        "c5bs1": [(1, 32768, 32, 8, 160, [0, 12, 14, 16, 18, 20, 24, 32]), (2, 32768, 32, 8, 160, [0, 8, 9, 10, 12, 16]),
                  (1, 32768, 4, 1, 160, [0, 48, 64, 96, 128])],
        "c4tp1": [(128, 4096, 32, 8, 16, [0])],
        "c4tp8": [(128, 4096, 4, 1, 16, [0])],
    }
    os.environ.pop("NVLLM_DECODE_PLAN", None)    # "planned" below = the split-KV plan; the balanced grid is explicit
    sel = []
    for s in args.shapes.split(","):
        sel += shapes[s]
    nvar = lib.nvllm_paged_decode_num_variants()
    variants = list(range(nvar)) if args.variants == "all" else [int(v) for v in args.variants.split(",")]   # -1 = auto
    fa = None
    if args.flash:
        try:
            from flash_attn import flash_attn_with_kvcache as fa
        except Exception as e:
            print(json.dumps({"flash": f"unavailable: {e}"}))
    for bs, ctx, hq, hkv, width, split_list in sel:
        one = bs * ctx * hkv * D * 4
        ncopy = max(1, min(16, -(-(1024 << 20) // one)))
        sets = share_table([make(bs, ctx, hq, hkv, dev, width, seed=c) for c in range(ncopy)])
        q, kc, vc, lens, tab, nbytes = sets[0]
        out = torch.empty_like(q)
        if fa is not None:
            f = lambda i: fa(sets[i % ncopy][0].unsqueeze(1), sets[i % ncopy][1], sets[i % ncopy][2], cache_seqlens=lens,
                             block_table=sets[i % ncopy][4], softmax_scale=D ** -0.5, causal=True)
            med, mn = timeit(f, flush)
            gmed, gmn = timeit_graph(f, ncopy)
            print(json.dumps(dict(impl="flash", bs=bs, ctx=ctx, hq=hq, hkv=hkv, us=round(med, 2), us_min=round(mn, 2),
                                  gbs=round(nbytes / med / 1e3, 1), graph_us=round(gmed, 2),
                                  graph_gbs=round(nbytes / gmed / 1e3, 1))), flush=True)
        for spec in [x for x in args.balanced.split(",") if x]:
            ctas, mc, warps = (int(t) for t in (spec.split(":") + ["0"])[:3])
            lib.nvllm_paged_decode_set_tuning(mc, warps)
            ws = torch.zeros(max(ops.decode_workspace_bytes(bs, hq, D, min(ctas, 0)), 16), dtype=torch.uint8, device=dev)
            f = lambda i: ops.flash_attn_with_kvcache(sets[i % ncopy][0], sets[i % ncopy][1], sets[i % ncopy][2], lens,
                                                      sets[i % ncopy][4], softmax_scale=D ** -0.5, balanced_ctas=max(-ctas, 0),
                                                      out=out, workspace=ws)
            med, mn = timeit(f, flush)
            gmed, gmn = timeit_graph(f, ncopy)
            print(json.dumps(dict(impl="cuda-balanced", ctas=ctas, min_chunks=mc, warps=warps, bs=bs, ctx=ctx, hq=hq, hkv=hkv,
                                  us=round(med, 2), us_min=round(mn, 2), gbs=round(nbytes / med / 1e3, 1),
                                  graph_us=round(gmed, 2), graph_gbs=round(nbytes / gmed / 1e3, 1))), flush=True)
        lib.nvllm_paged_decode_set_tuning(0, 0)
        if args.splits == "none":
            split_list = []
        elif args.splits:
            split_list = [int(x) for x in args.splits.split(",")]
        for v in variants:
            lib.nvllm_paged_decode_set_variant(v)
            done = set()
            for s in split_list:
                ns = s if s > 0 else ops.plan_num_splits(bs, hq, hkv, tab.size(1) * BLOCK)
                if ctx // ns < 32 or (s > 0 and ns in done):
                    continue
                done.add(ns)
                ws = torch.zeros(max(ops.decode_workspace_bytes(bs, hq, D, ns), 16), dtype=torch.uint8, device=dev)
                f = lambda i: ops.flash_attn_with_kvcache(sets[i % ncopy][0], sets[i % ncopy][1], sets[i % ncopy][2], lens,
                                                          sets[i % ncopy][4], softmax_scale=D ** -0.5, num_splits=ns,
                                                          out=out, workspace=ws)
                for cl in [int(x) for x in args.clusters.split(",")]:
                    if ns == 1 and cl != 0:
                        continue
                    assert lib.nvllm_paged_decode_set_cluster(cl) == 0
                    med, mn = timeit(f, flush)
                    gmed, gmn = timeit_graph(f, ncopy)
                    print(json.dumps(dict(impl="cuda", variant=v, bs=bs, ctx=ctx, hq=hq, hkv=hkv, splits=ns, cluster=cl,
                                          planned=(s == 0), us=round(med, 2), us_min=round(mn, 2),
                                          gbs=round(nbytes / med / 1e3, 1), graph_us=round(gmed, 2),
                                          graph_gbs=round(nbytes / gmed / 1e3, 1))), flush=True)
                lib.nvllm_paged_decode_set_cluster(0)
        del q, kc, vc, sets
        torch.cuda.empty_cache()
    lib.nvllm_paged_decode_set_variant(-1)


if __name__ == "__main__":
    main()
