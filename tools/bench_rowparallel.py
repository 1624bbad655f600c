"""Fused RowParallelLinear (GEMM + all-reduce, one launch) beside the reference's F.linear + NCCL all_reduce.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29555 \
        tools/bench_rowparallel.py [--check-only] [--out gpurun_out/rowparallel.jsonl]

One process per GPU (the reference's layout, llm_engine.py:29-36): NCCL process group, TPComm over CUDA IPC.
Correctness first (both against the float64 contraction and against the NCCL path), then per-call device time with
both arms captured in CUDA graphs of CHAIN back-to-back calls (how the decode path runs them, 2 per layer), CUDA
events on the launching stream, max over ranks.  The reference arm is exactly RowParallelLinear.forward
(nanovllm/layers/linear.py:185-190).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHAIN = 56   # 28 layers x (o_proj + down_proj)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check-only", action="store_true")
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from nano_vllm_learn_b200 import ops
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    dev = torch.device("cuda")
    # (name, in_features all ranks, out_features)
    shapes = [("qwen3-0.6b o_proj", 2048, 1024), ("qwen3-0.6b down_proj", 3072, 1024),
              ("qwen3-8b o_proj", 4096, 4096), ("qwen3-8b down_proj", 12288, 4096)]
    shapes = [s for s in shapes if (s[1] // world) % 64 == 0]
    comm = TPComm.from_process_group(TPComm.required_bytes(world, [(k // world, h) for _, k, h in shapes]))
    ops.rowparallel_set_plan(timeout_ms=10000)
    if rank == 0:
        print(f"transport: {comm.how}, {comm.nbytes / 2**20:.1f} MiB per rank, world {world}", flush=True)

    def reference(x, w):
        y = F.linear(x, w)
        dist.all_reduce(y)
        return y

    ok = True
    rows = []
    for name, k, h in shapes:
        for n in ([1, 16, 100, 512] if a.check_only else [1, 16, 32, 64, 128, 256, 512]):
            g = torch.Generator().manual_seed(1000 + n + k)          # same on every rank
            x = torch.randn(n, k, generator=g).to(torch.bfloat16)
            w = (torch.randn(h, k, generator=g) * k ** -0.5).to(torch.bfloat16)
            xr = x.chunk(world, 1)[rank].contiguous().to(dev)
            wr = w.chunk(world, 1)[rank].contiguous().to(dev)
            y = ops.rowparallel_linear_allreduce(xr, wr, None, comm)
            yref = reference(xr, wr)
            torch.cuda.synchronize()
            exact = x.double() @ w.double().T
            scale = exact.abs().max().item()
            e_mine = (y.cpu().double() - exact).abs().max().item()
            e_ref = (yref.cpu().double() - exact).abs().max().item()
            gathered = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(gathered, y)
            same = all(torch.equal(gathered[0].view(torch.int16), t.view(torch.int16)) for t in gathered)
            err_word = comm.status()[1]
            good = same and err_word == 0 and e_mine <= 2.0 ** -8 * scale
            ok = ok and good
            row = dict(shape=name, k_per_rank=k // world, out=h, n=n, world=world, plan=ops.rowparallel_plan(n, k // world, h, world),
                       err_fused=e_mine, err_reference=e_ref, scale=scale, ranks_identical=same, timeout_word=err_word)
            if not a.check_only:
                for arm, fn in (("fused", lambda: ops.rowparallel_linear_allreduce(xr, wr, None, comm)),
                                ("reference", lambda: reference(xr, wr))):
                    for _ in range(3):
                        fn()
                    torch.cuda.synchronize()
                    dist.barrier()
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr):
                        for _ in range(CHAIN):
                            fn()
                    for _ in range(3):
                        gr.replay()
                    torch.cuda.synchronize()
                    dist.barrier()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(a.iters):
                        gr.replay()
                    e1.record()
                    torch.cuda.synchronize()
                    t = torch.tensor([e0.elapsed_time(e1) * 1000 / (a.iters * CHAIN)], device=dev)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    row[f"us_{arm}"] = round(t.item(), 2)
                    del gr
                wbytes = h * (k // world) * 2
                row["w_stream_gbs_fused"] = round(wbytes / row["us_fused"] / 1e3, 1)
                row["speedup"] = round(row["us_reference"] / row["us_fused"], 2)
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        if a.out:
            os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
            with open(a.out, "w") as f:
                for r in rows:
                    f.write(json.dumps(r) + "\n")
        print("ROWPARALLEL_IPC_OK" if flag.item() == 1 else "ROWPARALLEL_IPC_FAILED", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
