#!/usr/bin/env python
"""bench.py — the hot path's headline metric on B200: paged-decode attention GB/s.

    python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
    python bench.py --impl reference ...                     (reference sdpa.math path on host CPU cores)

One *step* = one decode step's worth of the attention hot path for the workload: for each of the
model's L layers, `store_kvcache` of the new token of every sequence followed by the paged GQA decode
over that layer's own K/V cache.  Workload at N=1 (BASELINE.json configs[1], worst point of the
bench_my sweep, SURVEY §8d "C2"): Qwen3-0.6B shape (28 layers, 16 q / 8 kv heads, head_dim 128,
block 256), bs=64, ctx=4096.  Each layer has its own 1.07 GB cache (30 GB total), so the working
set is far beyond the 126 MB L2 and no flush is needed between iterations.

N > 1: one process per GPU, every rank runs the same per-GPU workload on its own shard of
(sequence, kv-head) units; attention needs no collective -> "scaling": "weak".
The JSON line also carries `tp_strong`: the Qwen3-8B-shape tensor-parallel point (configs[3],
bs=128 ctx 4096, 32q/8kv heads split N ways) measured on rank 0's shard.

value   = algorithmic bytes of all ranks' steps / max-over-ranks device time (CUDA events).
e2e     = same metric through the public `Attention.forward` + Context API with HOST inputs:
          per layer, q/k/v are copied from pinned host memory and the output is read back.
roofline= the paged_decode kernel alone: algorithmic bytes per launch / mean launch duration
          measured with CUDA events around every decode launch of the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

D = 128
BLOCK = 256
METRIC = "paged-decode attn GB/s"
UNIT = "GB/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    # workload overrides (defaults = the contract workload); used for ncu captures and sweeps
    ap.add_argument("--layers", type=int, default=28)
    ap.add_argument("--bs", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=4096)
    ap.add_argument("--hq", type=int, default=16)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--splits", type=int, default=0, help="0 = library plan")
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-tp", action="store_true")
    ap.add_argument("--no-prefill", action="store_true")
    ap.add_argument("--cpu-sample-bs", type=int, default=2)
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def decode_bytes(bs, ctx_list, hq, hkv):
    kv = sum(ctx_list) * hkv * D * 2 * 2
    qo = 2 * bs * hq * D * 2
    idx = 4 * bs + 4 * sum((c + BLOCK - 1) // BLOCK for c in ctx_list)
    return kv + qo + idx


def store_bytes(n, hkv):
    return n * (4 * hkv * D * 2 + 4)


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.proc = None
        self.path = None
        self.gpu = gpu_index

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.out = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.gpu)], stdout=self.out,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.out.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        with open(self.path) as f:
            for line in f:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    mx.append(float(parts[2]))
                except ValueError:
                    continue
                for name, val in zip(names, parts[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------- cuda arm
class Workload:
    """Per-rank synthetic decode workload resident in HBM."""

    def __init__(self, dev, layers, bs, ctx, hq, hkv, seed):
        g = torch.Generator(device=dev).manual_seed(seed)
        self.layers, self.bs, self.hq, self.hkv = layers, bs, hq, hkv
        # ctx_b includes the token stored this step (model_runner.py:252)
        self.ctx_list = [ctx] * bs
        pages = (ctx + BLOCK - 1) // BLOCK
        nblk = bs * pages
        self.kv = torch.empty(2, layers, nblk, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev)
        for l in range(layers):  # N(0,1) fill, layer by layer to bound temporaries
            self.kv[0, l].normal_(generator=g)
            self.kv[1, l].normal_(generator=g)
        perm = torch.randperm(nblk, generator=g, device=dev).to(torch.int32)   # non-contiguous pages
        self.tables = perm.view(bs, pages).contiguous()
        self.lens = torch.tensor(self.ctx_list, dtype=torch.int32, device=dev)
        last = self.tables[:, -1].long() * BLOCK + (ctx - (pages - 1) * BLOCK) - 1
        self.slots = last.to(torch.int32)
        self.q = torch.randn(layers, bs, hq * D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
        # k contiguous, v a strided view of a fused row (qwen3.py:106)
        self.kvrow = torch.randn(layers, bs, 2 * hkv * D, generator=g, device=dev, dtype=torch.float32).to(torch.bfloat16)
        self.bytes_decode = decode_bytes(bs, self.ctx_list, hq, hkv)
        self.bytes_store = store_bytes(bs, hkv)
        self.bytes_step = layers * (self.bytes_decode + self.bytes_store)

    def k_new(self, l):
        return self.kvrow[l, :, :self.hkv * D].view(self.bs, self.hkv, D)

    def v_new(self, l):
        return self.kvrow[l, :, self.hkv * D:].view(self.bs, self.hkv, D)


def run_cuda(args):
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the cuda arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    from nano_vllm_learn_b200 import _lib, ops, set_context, reset_context
    from nano_vllm_learn_b200.attention_cuda import Attention
    lib = _lib.load()
    if args.variant >= 0:
        assert lib.nvllm_paged_decode_set_variant(args.variant) == 0

    wl = Workload(dev, args.layers, args.bs, args.ctx, args.hq, args.hkv, seed=rank)
    scale = D ** -0.5
    splits = args.splits if args.splits > 0 else ops.plan_num_splits(wl.bs, wl.hkv, wl.tables.size(1) * BLOCK)
    out = torch.empty(wl.bs, wl.hq, D, dtype=torch.bfloat16, device=dev)
    ws_bytes = ops.decode_workspace_bytes(wl.bs, wl.hq, D, splits)
    ws = torch.zeros(max(ws_bytes, 16), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()

    def step(events=None):
        for l in range(wl.layers):
            ops.store_kvcache(wl.k_new(l), wl.v_new(l), wl.kv[0, l], wl.kv[1, l], wl.slots)
            if events is not None:
                e0 = torch.cuda.Event(enable_timing=True)
                e1 = torch.cuda.Event(enable_timing=True)
                e0.record(stream)
            ops.flash_attn_with_kvcache(wl.q[l].view(wl.bs, wl.hq, D), wl.kv[0, l], wl.kv[1, l], wl.lens, wl.tables,
                                        softmax_scale=scale, num_splits=splits, out=out, workspace=ws)
            if events is not None:
                e1.record(stream)
                events.append((e0, e1))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    lib.nvllm_reset_launch_count()
    events = []
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record(stream)
    for _ in range(args.steps):
        step(events)
    t1.record(stream)
    barrier()
    launches = int(lib.nvllm_launch_count())
    clocks = sampler.stop() if rank == 0 else None
    ms_total = t0.elapsed_time(t1)
    kern_ms = [a.elapsed_time(b) for a, b in events]
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * wl.bytes_step / (ms_per_step * 1e-3) / 1e9

    peak, peak_src = peaks()
    kern_us = statistics.mean(kern_ms) * 1e3
    achieved = wl.bytes_decode / (kern_us * 1e-6) / 1e9
    roofline = {"bound": "hbm", "kernel": "paged_decode_kernel" + ("+combine" if splits > 1 else ""),
                "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                "frac_of_nominal_8TBs": round(achieved / 8000.0, 4), "peak_source": peak_src,
                "us_per_launch": round(kern_us, 2), "algorithmic_bytes_per_launch": wl.bytes_decode,
                "num_splits": splits, "traffic": load_traffic()}

    # ---- e2e through the public API with host buffers ---------------------------------------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, wl, dev, Attention, set_context, reset_context, scale, dist)
        if dist is not None:
            tt = torch.tensor([e2e.pop("_ms")], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            e2e_ms = float(tt.item())
        else:
            e2e_ms = e2e.pop("_ms")
        e2e["value"] = round(world * wl.bytes_step / (e2e_ms * 1e-3) / 1e9, 1)
        e2e["ms_per_step"] = round(e2e_ms, 4)
        e2e["h2d_bytes_per_step"] *= world
        e2e["d2h_bytes_per_step"] *= world

    tp_strong = None
    if not args.no_tp:
        del wl.kv
        torch.cuda.empty_cache()
        tp_strong = run_tp_point(args, dev, world, rank, ops, dist)

    prefill = None
    if rank == 0 and not args.no_prefill:
        prefill = run_prefill_points(dev, ops)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args, sample_bs=args.cpu_sample_bs, iters=None)

    if dist is not None:
        dist.barrier()
    if rank == 0:
        line = {"metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"Qwen3-0.6B-shape decode attention hot path (store_kvcache + paged GQA decode) "
                                       f"x {wl.layers} layers, bs={wl.bs}, ctx={args.ctx}, {wl.hq}q/{wl.hkv}kv heads, "
                                       f"head_dim 128, block 256 (bench_my worst point, BASELINE configs[1])",
                           "per_gpu_algorithmic_bytes_per_step": wl.bytes_step,
                           "parallelism": f"{world} rank(s), (sequence, kv-head) shards, no collective",
                           "l2": "per-layer caches total %.1f GB >> 126 MB L2; no flush needed" %
                                 (wl.layers * wl.bs * ((args.ctx + BLOCK - 1) // BLOCK) * BLOCK * wl.hkv * D * 4 / 1e9),
                           "timing": "CUDA events on the launch stream, max over ranks"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
                "cpu_baseline": cpu, "tp_strong": tp_strong, "prefill": prefill}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def load_traffic():
    """dram bytes per launch of the decode kernel from the committed ncu summary, if present."""
    path = os.path.join(ROOT, "profiles", "decode_traffic.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                return json.load(f).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def run_e2e(args, wl, dev, Attention, set_context, reset_context, scale, dist):
    """Public-API path: Attention.forward(q, k, v) + Context, inputs in pinned host memory."""
    attn = [Attention(wl.hq, D, scale, wl.hkv) for _ in range(wl.layers)]
    for l, a in enumerate(attn):
        a.k_cache, a.v_cache = wl.kv[0, l], wl.kv[1, l]
    h_q = wl.q.cpu().pin_memory()                       # [L, bs, Hq*D]
    h_kv = wl.kvrow.cpu().pin_memory()                  # [L, bs, 2*Hkv*D]  (k | v fused row)
    h_slots = wl.slots.cpu().pin_memory()
    h_lens = wl.lens.cpu().pin_memory()
    h_tab = wl.tables.cpu().pin_memory()
    h_out = torch.empty(wl.layers, wl.bs, wl.hq * D, dtype=torch.bfloat16).pin_memory()
    h2d = (h_q.numel() + h_kv.numel()) * 2 + (h_slots.numel() + h_lens.numel() + h_tab.numel()) * 4
    d2h = h_out.numel() * 2
    kd = wl.hkv * D

    # Copies overlap compute on their own streams: the step's inputs are staged host->device on `up` (one event
    # per layer), the kernels run on the current stream, each layer's output goes device->host on `down`.
    # Every byte still moves inside the timed region; double-buffered device staging (2 x L slots) keeps a
    # step's uploads from overwriting inputs the previous step is still reading.
    cur = torch.cuda.current_stream()
    up, down = torch.cuda.Stream(), torch.cuda.Stream()
    d_q = torch.empty((2,) + tuple(h_q.shape), dtype=h_q.dtype, device=dev)
    d_kv = torch.empty((2,) + tuple(h_kv.shape), dtype=h_kv.dtype, device=dev)
    d_o = torch.empty((2, wl.layers, wl.bs, wl.hq * D), dtype=torch.bfloat16, device=dev)
    d_slots, d_lens, d_tab = (torch.empty_like(t, device=dev) for t in (h_slots, h_lens, h_tab))
    ev_in = [[torch.cuda.Event() for _ in range(wl.layers)] for _ in range(2)]
    ev_out = [[torch.cuda.Event() for _ in range(wl.layers)] for _ in range(2)]
    ev_meta, ev_done = torch.cuda.Event(), [torch.cuda.Event(), torch.cuda.Event()]
    state = {"n": 0}

    def step():
        par = state["n"] & 1
        state["n"] += 1
        up.wait_event(ev_done[par])            # slot `par` was last read two steps ago
        with torch.cuda.stream(up):
            d_slots.copy_(h_slots, non_blocking=True)
            d_lens.copy_(h_lens, non_blocking=True)
            d_tab.copy_(h_tab, non_blocking=True)
            ev_meta.record(up)
            for l in range(wl.layers):
                d_q[par, l].copy_(h_q[l], non_blocking=True)
                d_kv[par, l].copy_(h_kv[l], non_blocking=True)
                ev_in[par][l].record(up)
        cur.wait_event(ev_meta)
        set_context(False, slot_mapping=d_slots, context_lens=d_lens, block_tables=d_tab)
        for l in range(wl.layers):
            cur.wait_event(ev_in[par][l])
            kvr = d_kv[par, l]
            d_o[par, l].copy_(attn[l](d_q[par, l], kvr[:, :kd], kvr[:, kd:]))
            ev_out[par][l].record(cur)
            down.wait_event(ev_out[par][l])
            with torch.cuda.stream(down):
                h_out[l].copy_(d_o[par, l], non_blocking=True)
        ev_done[par].record(cur)
        cur.wait_stream(down)                  # the step is over when its last output byte is on the host
        reset_context()

    with torch.inference_mode():
        for _ in range(max(3, args.warmup)):
            step()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(args.steps):
            step()
        t1.record()
        torch.cuda.synchronize()
    return {"_ms": t0.elapsed_time(t1) / args.steps, "unit": UNIT, "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h,
            "api": "nano_vllm_learn_b200.attention_cuda.Attention.forward + Context, pinned host q/k/v in, output out; "
                   "H2D / compute / D2H on three streams"}


def run_tp_point(args, dev, world, rank, ops, dist):
    """configs[3]: Qwen3-8B-shape decode, bs=128 ctx=4096, 32q/8kv heads split `world` ways (strong scaling)."""
    if 8 % world != 0:
        return None
    hq, hkv, bs, ctx, layers = 32 // world, 8 // world, 128, 4096, 4
    wl = Workload(dev, layers, bs, ctx, hq, hkv, seed=100 + rank)
    scale = D ** -0.5
    splits = ops.plan_num_splits(bs, hkv, wl.tables.size(1) * BLOCK)
    out = torch.empty(bs, hq, D, dtype=torch.bfloat16, device=dev)
    ws = torch.zeros(max(ops.decode_workspace_bytes(bs, hq, D, splits), 16), dtype=torch.uint8, device=dev)

    def one(l):
        ops.flash_attn_with_kvcache(wl.q[l].view(bs, hq, D), wl.kv[0, l], wl.kv[1, l], wl.lens, wl.tables,
                                    softmax_scale=scale, num_splits=splits, out=out, workspace=ws)

    for _ in range(3):
        one(0)
    times = []
    for i in range(12):   # rotate over 4 distinct caches (>= 1.07 GB in total even at tp=8): nothing is re-read from L2
        torch.cuda._sleep(300_000)   # GPU busy while the host queues the launch: host launch latency is not timed
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record()
        one(i % layers)
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    us = statistics.median(times) * 1e3
    t = torch.tensor([us], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    us = float(t.item())
    peak, _ = peaks()
    gbs = wl.bytes_decode / (us * 1e-6) / 1e9
    return {"workload": f"Qwen3-8B-shape decode op, bs=128 ctx=4096, tp={world}: {hq}q/{hkv}kv heads per GPU",
            "us_per_launch_max_over_ranks": round(us, 2), "per_gpu_GBps": round(gbs, 1),
            "aggregate_GBps": round(gbs * world, 1), "frac_of_measured_peak": round(gbs / peak, 4),
            "num_splits": splits, "l2": "launches rotate over 4 distinct caches, working set >> 126 MB L2"}


def run_prefill_points(dev, ops):
    """Secondary line: the tcgen05 varlen-prefill kernel against the tensor roofline.
    FLOPs = 4*Hq*128*sum_i[Lq_i*cached_i + Lq_i(Lq_i+1)/2]; peak = measured bf16 burst."""
    peak = 1590.0
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            peak = float(json.load(f)["bf16_tflops"])
    out = []
    for name, b, lq, cached, hq, hkv in (("bench.py prefill step: 16 x 1024, 16q/8kv", 16, 1024, 0, 16, 8),
                                         ("prefix cache: 16 x (512 new + 512 cached), 16q/8kv", 16, 512, 512, 16, 8),
                                         ("Qwen3-8B-shape: 4 x 4096, 32q/8kv", 4, 4096, 0, 32, 8)):
        g = torch.Generator(device=dev).manual_seed(1)
        lk = lq + cached
        q = torch.empty(b * lq, hq, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
        cu_q = torch.arange(0, b * lq + 1, lq, dtype=torch.int32, device=dev)
        cu_k = torch.arange(0, b * lk + 1, lk, dtype=torch.int32, device=dev)
        if cached:
            pages = (lk + BLOCK - 1) // BLOCK
            k = torch.empty(b * pages, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            v = torch.empty(b * pages, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            tab = torch.randperm(b * pages, generator=g, device=dev).to(torch.int32).view(b, pages)
        else:
            k = torch.empty(b * lk, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            v = torch.empty(b * lk, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
            tab = None
        o = torch.empty_like(q)
        fl = 4 * hq * D * b * (lq * cached + lq * (lq + 1) // 2)
        fn = lambda: ops.flash_attn_varlen_func(q, k, v, lq, cu_q, lk, cu_k, softmax_scale=D ** -0.5, block_table=tab, out=o)
        for _ in range(3):
            fn()
        ts = []
        for _ in range(10):
            a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a_.record()
            fn()
            b_.record()
            torch.cuda.synchronize()
            ts.append(a_.elapsed_time(b_))
        us = statistics.median(ts) * 1e3
        out.append({"workload": name, "kernel": "prefill_tcgen05_kernel", "us": round(us, 1),
                    "achieved": round(fl / us / 1e6, 1), "peak": peak, "unit": "TFLOP/s", "bound": "tensor",
                    "frac": round(fl / us / 1e6 / peak, 4)})
        del q, k, v, o
    return out


# --------------------------------------------------------------------------------------- CPU arms
def _reference_sdpa():
    """The reference's own sdpa.math decode (attention_sdpa.py:122-182), imported UNMODIFIED from
    baseline/_ref (the pip --target install of /root/reference)."""
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(os.path.join(ref_root, "nanovllm")):
        sys.path.insert(0, ref_root)
        try:
            from nanovllm.layers import attention_sdpa
            from torch.nn.attention import SDPBackend
            return (lambda q, kc, vc, lens, tab: attention_sdpa.flash_attn_with_kvcache(
                q.unsqueeze(1), kc, vc, cache_seqlens=lens, block_table=tab, causal=False,
                sdpa_backend=SDPBackend.MATH)), "reference"
        except Exception as e:  # pragma: no cover - depends on the box
            sys.stderr.write(f"bench.py: reference import failed ({e}); using the oracle port\n")
    from oracle import paged_attention_oracle as orc
    return (lambda q, kc, vc, lens, tab: orc.paged_decode_ref(q, kc, vc, lens, tab, D ** -0.5)), "port"


def cpu_sample(args, sample_bs):
    g = torch.Generator().manual_seed(0)
    pages = (args.ctx + BLOCK - 1) // BLOCK
    nblk = sample_bs * pages
    kc = torch.randn(nblk, BLOCK, args.hkv, D, generator=g).to(torch.bfloat16)
    vc = torch.randn(nblk, BLOCK, args.hkv, D, generator=g).to(torch.bfloat16)
    q = torch.randn(sample_bs, args.hq, D, generator=g).to(torch.bfloat16)
    tab = torch.randperm(nblk, generator=g).to(torch.int32).view(sample_bs, pages)
    lens = torch.full((sample_bs,), args.ctx, dtype=torch.int32)
    nbytes = decode_bytes(sample_bs, [args.ctx] * sample_bs, args.hq, args.hkv)
    return q, kc, vc, lens, tab, nbytes


def cpu_baseline(args, sample_bs, iters):
    fn, kind = _reference_sdpa()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    q, kc, vc, lens, tab, nbytes = cpu_sample(args, sample_bs)
    with torch.no_grad():
        for _ in range(3):
            fn(q, kc, vc, lens, tab)
        times = []
        t_end = time.perf_counter() + 12.0
        while (iters is None and time.perf_counter() < t_end and len(times) < 200) or (iters and len(times) < iters):
            t0 = time.perf_counter()
            fn(q, kc, vc, lens, tab)
            times.append(time.perf_counter() - t0)
    med = statistics.median(times)
    return {"value": round(nbytes / med / 1e9, 3), "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
            "sample": f"one layer's paged decode op, {sample_bs} of the {args.bs} sequences at ctx={args.ctx}, "
                      f"{args.hq}q/{args.hkv}kv heads, median of {len(times)} calls ({med * 1e3:.1f} ms each)",
            "ms_per_call": round(med * 1e3, 3), "host_cpu_count": cores}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    fn, kind = _reference_sdpa()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    q, kc, vc, lens, tab, nbytes = cpu_sample(args, args.cpu_sample_bs)
    with torch.no_grad():
        for _ in range(max(3, args.warmup)):
            fn(q, kc, vc, lens, tab)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn(q, kc, vc, lens, tab)
        dt = (time.perf_counter() - t0) / args.steps
    val = round(nbytes / dt / 1e9, 3)
    sample = (f"each step = one layer's paged decode op on {args.cpu_sample_bs} of the {args.bs} sequences at "
              f"ctx={args.ctx} ({args.hq}q/{args.hkv}kv heads), reference attention_sdpa.flash_attn_with_kvcache "
              f"(torch SDPA MATH) on host CPU tensors")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"Qwen3-0.6B-shape decode attention hot path, bs={args.bs}, ctx={args.ctx}, "
                                   f"{args.hq}q/{args.hkv}kv heads, head_dim 128, block 256 (bounded CPU sample)"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
                             "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_cuda(a)
