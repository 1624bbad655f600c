"""GEMM part of the fused RowParallelLinear kernel ALONE on one GPU (exchange mode 2: the rank's partial goes straight
to y, no pushes, no waits) beside cuBLAS F.linear on the same shard — how well the kernel streams W.

    python tools/bench_rowparallel_local.py [--out gpurun_out/rowparallel_local.jsonl]
    NVLLM_RP_BK=64 python tools/bench_rowparallel_local.py ...      # 128-byte staged rows instead of 256
    python tools/bench_rowparallel_local.py --ncu                   # a few eager launches of one shape for ncu

Weights rotate over enough distinct copies to exceed the 126 MB L2 ("hbm" rows, how a decoder's layers see them) or
stay on one copy ("l2" rows, what a chained microbenchmark sees).  Both arms are captured in CUDA graphs that walk
every copy once and timed with CUDA events on the launching stream.  Device time only; this is a tuning tool, not bench.py.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)



def make_comm(nbytes):
    from nano_vllm_learn_b200 import _lib
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    lib = _lib.load()
    p = ctypes.c_void_p()
    _lib.check(lib.nvllm_tp_comm_alloc(nbytes, ctypes.byref(p), None), "tp_comm_alloc")
    # "world 2" whose two inboxes are the same local buffer: mode 2 never touches them
    return TPComm(0, 2, nbytes, p.value, [p.value, p.value], "local")


def timed_graph(fn, iters, chain):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        fn()
    for _ in range(3):
        gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / (iters * chain)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--ncu", action="store_true")
    ap.add_argument("--tile", type=int, default=0, help="force the feature tile: 32 | 64 (mma.sync) | 128 (tcgen05)")
    ap.add_argument("--tokens", type=int, nargs="+", default=[1, 16, 32, 64, 128, 256])
    ap.add_argument("--only", default="", help="substring filter on the shape name")
    a = ap.parse_args()
    from nano_vllm_learn_b200 import ops
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    torch.cuda.set_device(0)
    dev = torch.device("cuda")
    shapes = [("qwen3-0.6b o_proj tp2", 1024, 1024), ("qwen3-0.6b down_proj tp2", 1536, 1024),
              ("qwen3-8b o_proj tp2", 2048, 4096), ("qwen3-8b down_proj tp2", 6144, 4096),
              ("qwen3-8b down_proj tp8", 1536, 4096)]
    comm = make_comm(TPComm.required_bytes(2, [(k, h) for _, k, h in shapes]))
    ops.rowparallel_set_plan(a.tile, 0, 2)
    g = torch.Generator().manual_seed(0)
    if a.ncu:
        k, h, n = 6144, 4096, a.tokens[0] if a.tile == 128 else 16
        w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16).to(dev)
        x = torch.randn(n, k, generator=g).to(torch.bfloat16).to(dev)
        y = torch.empty(n, h, dtype=torch.bfloat16, device=dev)
        for _ in range(6):
            ops.rowparallel_linear_allreduce(x, w, None, comm, out=y)
        torch.cuda.synchronize()
        return
    rows = []
    for name, k, h in shapes:
        if a.only and a.only not in name:
            continue
        wbytes = k * h * 2
        ncopies = max(2, -(-(320 << 20) // wbytes))
        ncopies = min(ncopies, 160)
        ws = [(torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16).to(dev) for _ in range(min(ncopies, 4))]
        while len(ws) < ncopies:          # distinct memory, repeated values
            ws.append(ws[len(ws) % 4].clone())
        for n in a.tokens:
            x = torch.randn(n, k, generator=g).to(torch.bfloat16).to(dev)
            y = torch.empty(n, h, dtype=torch.bfloat16, device=dev)
            got = ops.rowparallel_linear_allreduce(x, ws[0], None, comm).float()
            want = F.linear(x, ws[0]).float()
            err = (got - want).abs().max().item()
            scale = want.abs().max().item()
            row = dict(shape=name, k_per_rank=k, out=h, n=n, bk=os.environ.get("NVLLM_RP_BK", "auto"),
                       plan=ops.rowparallel_plan(n, k, h, 2), max_abs_diff_vs_cublas=err, scale=scale)
            chain = max(8, ncopies)      # one graph walks every copy once: > 320 MB of W between reuses
            for res, pick in (("hbm", lambda i: ws[i % ncopies]), ("l2", lambda i: ws[0])):
                def fused():
                    for i in range(chain):
                        ops.rowparallel_linear_allreduce(x, pick(i), None, comm, out=y)

                def cublas():
                    for i in range(chain):
                        torch.mm(x, pick(i).t(), out=y)
                us_f, us_c = timed_graph(fused, a.iters, chain), timed_graph(cublas, a.iters, chain)
                row[f"us_fused_{res}"] = round(us_f, 2)
                row[f"us_cublas_{res}"] = round(us_c, 2)
                row[f"w_gbs_fused_{res}"] = round(wbytes / us_f / 1e3, 1)
                row[f"w_gbs_cublas_{res}"] = round(wbytes / us_c / 1e3, 1)
            rows.append(row)
            print(json.dumps(row), flush=True)
    ops.rowparallel_set_plan(0, 0, -1)
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        with open(a.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
