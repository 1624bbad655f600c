"""Prefill-kernel timing on one GPU: tcgen05 kernel vs mma.sync kernel vs flash-attn, CUDA-event timed.
Prints one JSON line per (shape, impl) with TFLOP/s (causal FLOPs)."""
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


def flops(lq_list, cached_list, hq):
    return 4 * hq * D * sum(lq * c + lq * (lq + 1) // 2 for lq, c in zip(lq_list, cached_list))


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return statistics.median(ts), min(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--flash", action="store_true")
    ap.add_argument("--shape-idx", type=int, default=-1)
    ap.add_argument("--impls", default="tcgen05,mma.sync")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--emu", default="", help="comma list of exp2-on-FMA settings of the tcgen05 kernel (pairs of 8; "
                    "needs the tools library: python nano-vllm-learn_b200/build.py --tools)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    if args.emu:
        _lib.load_tools()      # before anything loads the product library
    # (name, batch, Lq, cached, Hq, Hkv)
    shapes = [("bench.py step 16x1024 0.6B", 16, 1024, 0, 16, 8),
              ("C3 prefix 16x(512 new + 512 cached) 0.6B", 16, 512, 512, 16, 8),
              ("1x8192 0.6B", 1, 8192, 0, 16, 8),
              ("4x4096 8B-shape", 4, 4096, 0, 32, 8),
              ("64x256 0.6B", 64, 256, 0, 16, 8),
              ("warmup 1x4096 0.6B", 1, 4096, 0, 16, 8)]
    if args.shape_idx >= 0:
        shapes = [shapes[args.shape_idx]]
    fa = None
    if args.flash:
        try:
            from flash_attn import flash_attn_varlen_func as fa
        except Exception as e:
            print(json.dumps({"flash": f"unavailable: {e}"}))
    for name, b, lq, cached, hq, hkv in shapes:
        g = torch.Generator(device=dev).manual_seed(0)
        lk = lq + cached
        q = torch.empty(b * lq, hq, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
        cu_q = torch.arange(0, b * lq + 1, lq, dtype=torch.int32, device=dev)
        cu_k = torch.arange(0, b * lk + 1, lk, dtype=torch.int32, device=dev)
        fl = flops([lq] * b, [cached] * b, hq)
        scale = D ** -0.5
        if cached:
            pages = (lk + BLOCK - 1) // BLOCK
            nblk = b * pages
            k = torch.empty(nblk, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            v = torch.empty(nblk, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            tab = torch.randperm(nblk, generator=g, device=dev).to(torch.int32).view(b, pages)
        else:
            k = torch.empty(b * lk, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            v = torch.empty(b * lk, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            tab = None
        out = torch.empty_like(q)
        impls = [(n, v) for n, v in (("tcgen05", 1), ("mma.sync", 0)) if n in args.impls.split(",")]
        for label, variant in impls:
            ops.set_prefill_variant(variant)
            f = lambda: ops.flash_attn_varlen_func(q, k, v, lq, cu_q, lk, cu_k, softmax_scale=scale, block_table=tab, out=out)
            med, mn = timeit(f, args.iters)
            print(json.dumps(dict(shape=name, impl=label, us=round(med, 1), us_min=round(mn, 1),
                                  tflops=round(fl / med / 1e6, 1), gflop=round(fl / 1e9, 2))), flush=True)
        if args.emu:
            tl = _lib.load_tools()
            ops.set_prefill_variant(1)
            f = lambda: ops.flash_attn_varlen_func(q, k, v, lq, cu_q, lk, cu_k, softmax_scale=scale, block_table=tab, out=out)
            base = None
            for e in [int(x) for x in args.emu.split(",")]:
                tl.nvllm_varlen_prefill_set_exp2_poly(e)
                med, mn = timeit(f, args.iters)
                res = f().float()
                if base is None:
                    base = res
                print(json.dumps(dict(shape=name, impl="tcgen05", exp2_poly_pairs_of_8=e, us=round(med, 1), us_min=round(mn, 1),
                                      tflops=round(fl / med / 1e6, 1),
                                      max_abs_diff_vs_first=round((res - base).abs().max().item(), 6))), flush=True)
            tl.nvllm_varlen_prefill_set_exp2_poly(-1)
        ops.set_prefill_variant(-1)
        if fa is not None:
            f = lambda: fa(q, k, v, max_seqlen_q=lq, cu_seqlens_q=cu_q, max_seqlen_k=lk, cu_seqlens_k=cu_k,
                           softmax_scale=scale, causal=True, block_table=tab)
            med, mn = timeit(f)
            print(json.dumps(dict(shape=name, impl="flash-attn", us=round(med, 1), us_min=round(mn, 1),
                                  tflops=round(fl / med / 1e6, 1), gflop=round(fl / 1e9, 2))), flush=True)
        del q, k, v
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
