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
    ap.add_argument("--splits", type=int, default=None,
                    help="default: library plan; 0 / -P = balanced persistent grid (P CTAs); >= 1 = fixed split-KV grid")
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-tp", action="store_true")
    ap.add_argument("--no-prefill", action="store_true")
    ap.add_argument("--cpu-sample-bs", type=int, default=2)
    ap.add_argument("--no-flash", action="store_true", help="skip the flash-attn comparator legs (vs_flash)")
    ap.add_argument("--no-long-ctx", action="store_true")
    ap.add_argument("--no-fused-allreduce", action="store_true")
    ap.add_argument("--no-tp-e2e", action="store_true", help="skip the bounded end-to-end engine run (bench_my protocol)")
    ap.add_argument("--tp-e2e-timeout", type=int, default=420)
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


def time_back_to_back(fn, n_sets, launches=60, warm=6):
    """Per-launch device time of `fn(i)` (i rotates over `n_sets` input copies whose total exceeds L2): `launches`
    launches queued back to back on the current stream with a CUDA event between neighbours, ONE synchronise at the
    end -- the host runs ahead of the GPU, so no launch latency and no single-launch event overhead is timed.
    Returns min / median / max / mean in microseconds."""
    for i in range(warm):
        fn(i % n_sets)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(launches + 1)]
    torch.cuda._sleep(2_000_000)          # ~1 ms of GPU work: the queue is full before the first timed launch starts
    ev[0].record()
    for i in range(launches):
        fn(i % n_sets)
        ev[i + 1].record()
    torch.cuda.synchronize()
    us = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(launches))
    return {"min": round(us[0], 2), "median": round(statistics.median(us), 2), "max": round(us[-1], 2),
            "mean": round(sum(us) / len(us), 2), "launches": launches}


def time_in_graph(fn, n_sets, reps=2, iters=5):
    """Per-launch device time inside ONE CUDA graph of n_sets*reps launches (how the engine replays a decode step:
    one launch per layer, each on its own cache)."""
    n = n_sets * reps
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(n_sets):
            fn(i)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(n):
            fn(i % n_sets)
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3 / n)
    del g
    return round(statistics.median(ts), 2)


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
    splits = args.splits if args.splits is not None else ops.plan_num_splits(wl.bs, wl.hq, wl.hkv, wl.tables.size(1) * BLOCK)
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
    kname = "paged_decode_balanced_kernel" if splits <= 0 else "paged_decode_kernel" + ("+combine" if splits > 1 else "")
    roofline = {"bound": "hbm", "kernel": kname,
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

    cfg_layers, cfg_bs, cfg_hq, cfg_hkv, bytes_step = wl.layers, wl.bs, wl.hq, wl.hkv, wl.bytes_step
    vs_flash = None
    if rank == 0 and not args.no_flash:
        vs_flash = run_vs_flash_headline(wl, ops, scale, splits, out, ws)

    tp_strong = None
    del wl.kv
    torch.cuda.empty_cache()
    if not args.no_tp:
        tp_strong = run_tp_point(args, dev, world, rank, ops, dist)

    long_ctx = None
    if rank == 0 and not args.no_long_ctx:
        long_ctx = run_long_ctx(args, dev, ops)

    prefill = None
    if rank == 0 and not args.no_prefill:
        prefill = run_prefill_points(dev, ops, with_flash=not args.no_flash)

    fused = None
    if dist is not None and not args.no_fused_allreduce:
        try:
            fused = run_fused_allreduce(args, dev, rank, world, ops, dist)
        except Exception as e:   # a comparator leg must not take the headline line down with it
            fused = {"failed": f"{type(e).__name__}: {e}"[:300]}

    vp_sampler = None
    if dist is not None and not args.no_fused_allreduce:
        try:
            vp_sampler = run_vocab_parallel_sampler(args, dev, rank, world, ops, dist)
        except Exception as e:
            vp_sampler = {"failed": f"{type(e).__name__}: {e}"[:300]}

    fused_emb = None
    if dist is not None and not args.no_fused_allreduce:
        try:
            fused_emb = run_fused_embedding(args, dev, rank, world, ops, dist)
        except Exception as e:
            fused_emb = {"failed": f"{type(e).__name__}: {e}"[:300]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args, sample_bs=args.cpu_sample_bs, iters=None)

    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    del wl
    torch.cuda.empty_cache()
    if rank != 0:
        return

    tp_e2e = None
    if not args.no_tp_e2e and os.environ.get("NVLLM_BENCH_TP_E2E", "1") != "0":
        time.sleep(3.0 if world > 1 else 0.0)    # the other ranks are exiting and releasing their GPUs
        try:
            tp_e2e = run_tp_e2e(args, world)
        except Exception as e:
            tp_e2e = {"failed": f"{type(e).__name__}: {e}"[:300]}

    line = {"metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"Qwen3-0.6B-shape decode attention hot path (store_kvcache + paged GQA decode) "
                                   f"x {cfg_layers} layers, bs={cfg_bs}, ctx={args.ctx}, {cfg_hq}q/{cfg_hkv}kv heads, "
                                   f"head_dim 128, block 256 (bench_my worst point, BASELINE configs[1])",
                       "per_gpu_algorithmic_bytes_per_step": bytes_step,
                       "parallelism": f"{world} rank(s), (sequence, kv-head) shards, no collective",
                       "l2": "per-layer caches total %.1f GB >> 126 MB L2; no flush needed" %
                             (cfg_layers * cfg_bs * ((args.ctx + BLOCK - 1) // BLOCK) * BLOCK * cfg_hkv * D * 4 / 1e9),
                       "timing": "CUDA events on the launch stream, max over ranks"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline,
            "cpu_baseline": cpu, "vs_flash": vs_flash, "tp_strong": tp_strong, "long_ctx": long_ctx,
            "prefill": prefill, "fused_allreduce": fused, "vocab_parallel_sampler": vp_sampler,
            "fused_embedding": fused_emb, "tp_e2e": tp_e2e}
    # The line is ~25 KB and a log that keeps only its tail loses the legs printed first (missing #3): the
    # LAST key repeats, in under 1.5 KB, the one or two numbers of every leg that the claims in DESIGN.md rest on.
    line["digest"] = leg_digest(line)
    print(json.dumps(line), flush=True)


def leg_digest(line):
    """Compact copy of the comparator / scaling / parity numbers of every leg (values only, no prose)."""
    def pick(d, *keys):
        return {k: d[k] for k in keys if isinstance(d, dict) and k in d} if isinstance(d, dict) else None
    out = {"roofline_frac": (line.get("roofline") or {}).get("frac"),
           "e2e": (line.get("e2e") or {}).get("value")}
    out["vs_flash"] = pick(line.get("vs_flash"), "decode_us", "flash_attn_decode_us", "speedup", "max_abs_diff")
    out["tp_strong"] = pick(line.get("tp_strong"), "us_per_launch_max_over_ranks", "per_gpu_GBps", "frac_of_measured_peak")
    if isinstance(line.get("long_ctx"), list):
        out["long_ctx"] = [[r.get("bs"), r.get("heads", "")[:8], r.get("graph_us_per_launch"), r.get("GBps"),
                            r.get("flash_attn_graph_us_per_launch")] for r in line["long_ctx"] if isinstance(r, dict)]
    if isinstance(line.get("prefill"), list):
        out["prefill"] = [[r.get("us"), r.get("achieved"), r.get("frac"), r.get("flash_attn_us")]
                          for r in line["prefill"] if isinstance(r, dict)]
    for leg in ("fused_allreduce", "vocab_parallel_sampler", "fused_embedding"):
        v = line.get(leg)
        if isinstance(v, dict):
            rows = v.get("rows") or v.get("shapes") or []
            out[leg] = {"parity_ok": v.get("parity_ok"),
                        "speedup_min_max": [min((r.get("speedup", 0) for r in rows if isinstance(r, dict)), default=None),
                                            max((r.get("speedup", 0) for r in rows if isinstance(r, dict)), default=None)]}
    e = line.get("tp_e2e")
    if isinstance(e, dict):
        out["tp_e2e"] = {"arms": {a.get("arm"): a.get("tok_s") for a in e.get("arms", []) if isinstance(a, dict)},
                         "cuda_over_flash": e.get("cuda_over_flash")}
    return out


def decode_kernel_digest():
    """sha256 over the sources the decode kernel is compiled from; profiles/decode_traffic.json records it."""
    import hashlib
    h = hashlib.sha256()
    csrc = os.path.join(ROOT, "nano-vllm-learn_b200", "csrc")
    for name in ("common.cuh", "paged_decode.cu", "paged_decode_params.cuh"):
        with open(os.path.join(csrc, name), "rb") as f:
            h.update(name.encode())
            h.update(f.read())
    return h.hexdigest()


def load_traffic():
    """dram bytes per launch of the decode kernel from the committed ncu summary -- None when the summary was taken on
    other kernel sources than the ones in this tree (tools/make_decode_traffic.py writes the digest)."""
    path = os.path.join(ROOT, "profiles", "decode_traffic.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                rec = json.load(f)
            if rec.get("kernel_source_sha256") != decode_kernel_digest():
                return None
            return rec.get("dram_bytes_per_launch")
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


def _decode_sets(dev, bs, ctx, hq, hkv, width, min_total_bytes, max_sets, seed):
    """Independent (q, k_cache, v_cache) copies so that consecutive launches never re-read K/V from L2; ONE block table and
    context_lens shared by all of them, as the layers of a model share the step's table while each owns its cache."""
    pages = (ctx + BLOCK - 1) // BLOCK
    one = bs * pages * BLOCK * hkv * D * 2 * 2
    n_sets = max(2, min(max_sets, -(-min_total_bytes // one)))
    sets = []
    g = torch.Generator(device=dev).manual_seed(seed)
    nblk = bs * pages
    tab = torch.randperm(nblk, generator=g, device=dev).to(torch.int32).view(bs, pages)
    if width and width > pages:     # captured-graph layout: zero padded to ceil(max_model_len / 256) columns
        wide = torch.zeros(bs, width, dtype=torch.int32, device=dev)
        wide[:, :pages] = tab
        tab = wide
    for _ in range(n_sets):
        kc = torch.empty(nblk, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
        vc = torch.empty(nblk, BLOCK, hkv, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
        q = torch.empty(bs, hq, D, dtype=torch.bfloat16, device=dev).normal_(generator=g)
        sets.append((q, kc, vc, tab))
    lens = torch.full((bs,), ctx, dtype=torch.int32, device=dev)
    return sets, lens, decode_bytes(bs, [ctx] * bs, hq, hkv)


def _time_decode_shape(dev, ops, bs, ctx, hq, hkv, width, seed, flash=None, min_total=1 << 30, max_sets=16):
    sets, lens, nbytes = _decode_sets(dev, bs, ctx, hq, hkv, width, min_total, max_sets, seed)
    n_sets = len(sets)
    scale = D ** -0.5
    splits = ops.plan_num_splits(bs, hq, hkv, sets[0][3].size(1) * BLOCK)
    out = torch.empty(bs, hq, D, dtype=torch.bfloat16, device=dev)
    ws = torch.zeros(max(ops.decode_workspace_bytes(bs, hq, D, splits), 16), dtype=torch.uint8, device=dev)

    def one(i):
        q, kc, vc, tab = sets[i]
        ops.flash_attn_with_kvcache(q, kc, vc, lens, tab, softmax_scale=scale, num_splits=splits, out=out, workspace=ws)

    t = time_back_to_back(one, n_sets)
    res = {"us": t, "graph_us_per_launch": time_in_graph(one, n_sets), "algorithmic_bytes": nbytes, "num_splits": splits,
           "input_copies_rotated": n_sets}
    if flash is not None:
        def fl(i):
            q, kc, vc, tab = sets[i]
            flash(q.unsqueeze(1), kc, vc, cache_seqlens=lens, block_table=tab, softmax_scale=scale, causal=True)
        try:
            res["flash_us"] = time_back_to_back(fl, n_sets)
            res["flash_graph_us_per_launch"] = time_in_graph(fl, n_sets)
        except Exception as e:   # comparator only
            res["flash_us"] = f"failed: {type(e).__name__}: {e}"[:200]
    del sets, out, ws
    torch.cuda.empty_cache()
    return res


def _flash_decode():
    try:
        from flash_attn import flash_attn_with_kvcache
        return flash_attn_with_kvcache
    except Exception:
        return None


def run_tp_point(args, dev, world, rank, ops, dist):
    """configs[3]: Qwen3-8B-shape decode, bs=128 ctx=4096, 32q/8kv heads split `world` ways (strong scaling).
    60 launches back to back rotating over >= 1 GiB of distinct caches; min / median / max, max over ranks."""
    if 8 % world != 0:
        return None
    hq, hkv, bs, ctx = 32 // world, 8 // world, 128, 4096
    r = _time_decode_shape(dev, ops, bs, ctx, hq, hkv, 16, seed=100 + rank, flash=None if args.no_flash else _flash_decode(),
                           min_total=2 << 30, max_sets=8)
    vals = [r["us"]["median"], r["us"]["min"], r["us"]["max"], r["graph_us_per_launch"]]
    t = torch.tensor(vals, dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    med, mn, mx, gr = (float(x) for x in t.tolist())
    peak, _ = peaks()
    gbs = r["algorithmic_bytes"] / (med * 1e-6) / 1e9
    out = {"workload": f"Qwen3-8B-shape decode op, bs=128 ctx=4096, tp={world}: {hq}q/{hkv}kv heads per GPU",
           "us_per_launch_max_over_ranks": round(med, 2), "us_min": round(mn, 2), "us_max": round(mx, 2),
           "graph_us_per_launch": round(gr, 2), "per_gpu_GBps": round(gbs, 1), "aggregate_GBps": round(gbs * world, 1),
           "frac_of_measured_peak": round(gbs / peak, 4), "frac_of_nominal_8TBs": round(gbs / 8000.0, 4),
           "num_splits": r["num_splits"],
           "timing": f"median of {r['us']['launches']} back-to-back launches (CUDA event between neighbours), max over ranks",
           "l2": f"launches rotate over {r['input_copies_rotated']} distinct caches (>= 2 GiB), nothing is re-read from L2"}
    if isinstance(r.get("flash_us"), dict):
        out["flash_attn_us_rank0"] = r["flash_us"]["median"]
    return out


def run_long_ctx(args, dev, ops):
    """configs[4]: Qwen3-8B-shape long-context decode, ctx 32768, table width 160 (max_model_len 40960), bs 1/2/4 at
    tp=1 (32q/8kv) and the tp=8 shard (4q/1kv).  Small batches: the launch is short, so both the back-to-back event
    timing and the in-graph per-launch time are reported; launches rotate over >= 1 GiB of distinct caches."""
    peak, _ = peaks()
    flash = None if args.no_flash else _flash_decode()
    rows = []
    # last row: BASELINE configs[0] (the reference's CPU-runnable case, bs=1 ctx=512 16q/8kv, 2 pages) on the GPU, so
    # the CPU milliseconds of BASELINE.md §2 have a GPU figure beside them (launch-latency bound: 2.1 MB of traffic)
    for bs, hq, hkv, tp, ctx, width in ((1, 32, 8, 1, 32768, 160), (2, 32, 8, 1, 32768, 160), (4, 32, 8, 1, 32768, 160),
                                        (1, 4, 1, 8, 32768, 160), (4, 4, 1, 8, 32768, 160), (1, 16, 8, 1, 512, 2)):
        r = _time_decode_shape(dev, ops, bs, ctx, hq, hkv, width, seed=500 + bs + tp, flash=flash)
        us = r["us"]["median"]
        gus = r["graph_us_per_launch"]
        # headline of the row = the in-graph per-launch time (decode runs inside CUDA graphs in the engine, and a 12-30 us
        # launch issued from Python is host-bound); the stream-issued figure is kept beside it
        row = {"bs": bs, "ctx": ctx, "heads": f"{hq}q/{hkv}kv (tp={tp} shard)", "graph_us_per_launch": gus,
               "GBps": round(r["algorithmic_bytes"] / gus / 1e3, 1),
               "frac_of_measured_peak": round(r["algorithmic_bytes"] / gus / 1e3 / peak, 4),
               "stream_us": us, "stream_us_min": r["us"]["min"], "stream_us_max": r["us"]["max"],
               "stream_GBps": round(r["algorithmic_bytes"] / us / 1e3, 1),
               "algorithmic_bytes": r["algorithmic_bytes"], "num_splits": r["num_splits"]}
        if isinstance(r.get("flash_us"), dict):
            row["flash_attn_us"] = r["flash_us"]["median"]
            row["flash_attn_graph_us_per_launch"] = r.get("flash_graph_us_per_launch")
        if r["algorithmic_bytes"] * r["input_copies_rotated"] < (126 << 20):
            row["l2"] = "all rotated copies together fit the 126 MB L2: a latency figure, not an HBM one"
        rows.append(row)
    return rows


def run_prefill_points(dev, ops, with_flash=True):
    """Secondary line: the varlen-prefill kernel against the tensor roofline, flash-attn's
    flash_attn_varlen_func (the reference `flash` backend's call, attention.py:93-96) on the same tensors beside it.
    FLOPs = 4*Hq*128*sum_i[Lq_i*cached_i + Lq_i(Lq_i+1)/2]; peak = measured bf16 burst."""
    peak = 1590.0
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            peak = float(json.load(f)["bf16_tflops"])
    fa = None
    if with_flash:
        try:
            from flash_attn import flash_attn_varlen_func as fa
        except Exception:
            fa = None
    out = []

    def med_us(fn):
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
        return statistics.median(ts) * 1e3

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
        us = med_us(fn)
        row = {"workload": name, "kernel": ops.PREFILL_KERNEL_NAMES[ops.prefill_last_path()], "us": round(us, 1),
               "achieved": round(fl / us / 1e6, 1), "peak": peak, "unit": "TFLOP/s", "bound": "tensor",
               "frac": round(fl / us / 1e6 / peak, 4)}
        if fa is not None:
            try:
                ffn = lambda: fa(q, k, v, max_seqlen_q=lq, cu_seqlens_q=cu_q, max_seqlen_k=lk, cu_seqlens_k=cu_k,
                                 softmax_scale=D ** -0.5, causal=True, block_table=tab)
                fus = med_us(ffn)
                row["flash_attn_us"] = round(fus, 1)
                row["flash_attn_TFLOPs"] = round(fl / fus / 1e6, 1)
                row["max_abs_diff_vs_flash_attn"] = round((ffn().float() - o.float()).abs().max().item(), 5)
            except Exception as e:   # comparator only
                row["flash_attn_us"] = f"failed: {type(e).__name__}: {e}"[:200]
        out.append(row)
        del q, k, v, o
    out.append({"mbarrier_timeout_latch": ops.prefill_error()})
    return out


def run_vs_flash_headline(wl, ops, scale, splits, out, ws):
    """The headline decode launch (one layer of the bench workload) beside flash-attn's flash_attn_with_kvcache -- the
    exact call of the reference `flash` backend (attention.py:99-101) -- same tensors, same process, launches rotating
    over the per-layer caches (30 GB >> L2), back-to-back event timing."""
    fa = _flash_decode()
    if fa is None:
        return {"unavailable": "flash_attn not importable"}
    n = wl.layers

    def mine(i):
        ops.flash_attn_with_kvcache(wl.q[i].view(wl.bs, wl.hq, D), wl.kv[0, i], wl.kv[1, i], wl.lens, wl.tables,
                                    softmax_scale=scale, num_splits=splits, out=out, workspace=ws)

    def theirs(i):
        return fa(wl.q[i].view(wl.bs, 1, wl.hq, D), wl.kv[0, i], wl.kv[1, i], cache_seqlens=wl.lens,
                  block_table=wl.tables, softmax_scale=scale, causal=True)

    try:
        t_f = time_back_to_back(theirs, n, launches=56)
        t_m = time_back_to_back(mine, n, launches=56)
        mine(0)
        diff = (theirs(0).view(wl.bs, wl.hq, D).float() - out.float()).abs().max().item()
    except Exception as e:
        return {"failed": f"{type(e).__name__}: {e}"[:200]}
    return {"decode_us": t_m["median"], "decode_GBps": round(wl.bytes_decode / t_m["median"] / 1e3, 1),
            "flash_attn_decode_us": t_f["median"], "flash_attn_decode_GBps": round(wl.bytes_decode / t_f["median"] / 1e3, 1),
            "speedup": round(t_f["median"] / t_m["median"], 3), "max_abs_diff": round(diff, 5),
            "flash_attn_version": getattr(__import__("flash_attn"), "__version__", "?"),
            "timing": "median of 56 back-to-back launches each, rotating over the 28 per-layer caches"}


def run_fused_allreduce(args, dev, rank, world, ops, dist):
    """SURVEY §8f rank 2, N > 1 only: the fused RowParallelLinear (GEMM + all-reduce in one launch over NVLink peer
    memory) beside the reference's `F.linear` + `dist.all_reduce` (nanovllm/layers/linear.py:185-190) on the o_proj /
    down_proj shards of Qwen3-0.6B and the Qwen3-8B shape at tensor_parallel_size = N.
    Parity on the box first: every rank's y bit-identical, |y - exact| <= 2^-8 * max|exact| (exact = fp64 contraction
    all-reduced over the ranks), no timeout latched.  Then device time per call: both arms as CUDA-graph chains of 28
    calls, CUDA events, max over ranks."""
    import torch.nn.functional as F
    from nano_vllm_learn_b200 import rowparallel_cuda as rp
    shapes = [("qwen3-0.6b o_proj", 2048, 1024), ("qwen3-0.6b down_proj", 3072, 1024),
              ("qwen3-8b-shape o_proj", 4096, 4096), ("qwen3-8b-shape down_proj", 12288, 4096)]
    shapes = [sh for sh in shapes if sh[1] % world == 0 and (sh[1] // world) % 64 == 0]
    ops.rowparallel_set_plan(timeout_ms=5000)
    comm = rp.TPComm.from_process_group(rp.TPComm.required_bytes(world, [(k // world, h) for _, k, h in shapes]))
    chain, iters = 28, 5
    rows, all_ok = [], True

    def reference(x, w):
        y = F.linear(x, w)
        dist.all_reduce(y)
        return y

    def timed(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(chain):
                fn()
        g.replay()
        torch.cuda.synchronize()
        dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b) * 1e3 / (iters * chain)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        del g
        return round(float(t.item()), 2)

    for name, k, h in shapes:
        kr = k // world
        for n in (1, 16, 64, 128, 256, 512):
            gen = torch.Generator(device=dev).manual_seed(7000 + 13 * n + k + rank)
            x = torch.empty(n, kr, dtype=torch.bfloat16, device=dev).normal_(generator=gen)
            w = (torch.empty(h, kr, dtype=torch.float32, device=dev).normal_(generator=gen) * k ** -0.5).to(torch.bfloat16)
            y = ops.rowparallel_linear_allreduce(x, w, None, comm)
            exact = x.double() @ w.double().T
            dist.all_reduce(exact)
            torch.cuda.synchronize()
            scale = exact.abs().max().item()
            err = (y.double() - exact).abs().max().item()
            gathered = [torch.empty_like(y) for _ in range(world)]
            dist.all_gather(gathered, y)
            same = all(torch.equal(gathered[0].view(torch.int16), t.view(torch.int16)) for t in gathered)
            err_ref = (reference(x, w).double() - exact).abs().max().item()
            flag = torch.tensor([1 if (same and err <= 2.0 ** -8 * scale and rp.poll_error() == 0) else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok = bool(flag.item())
            all_ok = all_ok and ok
            row = {"shard": name, "in_features_per_rank": kr, "out_features": h, "tokens": n, "parity_ok": ok,
                   "max_abs_err_fused": float(f"{err:.3e}"), "max_abs_err_reference": float(f"{err_ref:.3e}"),
                   "bound_2^-8_scale": float(f"{2.0 ** -8 * scale:.3e}"), "ranks_bit_identical": same}
            if ok:
                row["us_fused"] = timed(lambda: ops.rowparallel_linear_allreduce(x, w, None, comm))
                row["us_nccl"] = timed(lambda: reference(x, w))
                row["speedup"] = round(row["us_nccl"] / row["us_fused"], 2)
                # rooflines of one call: W streamed once from HBM vs the fp32 partials a rank must push over NVLink
                wbytes, push = h * kr * 2, n * h * 4 * (world - 1)
                row["hbm_floor_us"] = round(wbytes / (peaks()[0] * 1e3), 2)
                row["nvlink_push_bytes"] = push
                row["nvlink_floor_us_at_770GBps"] = round(push / 770e3, 2)
                row["plan_feature_tile_ksplit_mode"] = list(ops.rowparallel_plan(n, kr, h, world))   # tile 128 = tcgen05
            rows.append(row)
    torch.cuda.synchronize()
    dist.barrier()
    rp.release() if False else None
    return {"parity_ok": all_ok, "world": world, "transport": comm.how, "timeout_latch": rp.poll_error(),
            "timing": f"CUDA-graph chains of {chain} calls x {iters} replays, max over ranks", "rows": rows}


def run_fused_embedding(args, dev, rank, world, ops, dist):
    """SURVEY §8f rank 2, third collective, N > 1 only: the fused VocabParallelEmbedding (lookup + row exchange in one
    launch over NVLink peer memory) beside the reference's masked F.embedding + dist.all_reduce
    (nanovllm/layers/embed_head.py:35-43), Qwen3-0.6B (hidden 1024) and Qwen3-8B (hidden 4096) tables at
    tensor_parallel_size = N.  Parity: bit-identical to the reference formula on every rank.  Then device time per call,
    CUDA-graph chains, max over ranks."""
    import torch.nn.functional as F
    from nano_vllm_learn_b200 import embedding_cuda
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    vocab = 151936
    if vocab % world:
        return {"skipped": f"vocab {vocab} not divisible by {world}"}
    per = vocab // world
    _lib = __import__("nano_vllm_learn_b200._lib", fromlist=["x"])
    _lib.check(_lib.load().nvllm_vocab_embedding_set_timeout_ms(5000), "set_timeout")
    rows, all_ok = [], True
    chain, iters = 28, 5
    for hidden in (1024, 4096):
        comm = TPComm.from_process_group(embedding_cuda.comm_bytes(world, hidden))
        gen = torch.Generator(device=dev).manual_seed(4000 + hidden + rank)
        w = torch.empty(per, hidden, dtype=torch.bfloat16, device=dev).normal_(generator=gen)
        start = rank * per

        def reference(x):
            mask = (x >= start) & (x < start + per)
            y = mask.unsqueeze(1) * F.embedding(mask * (x - start), w)
            dist.all_reduce(y)
            return y

        for n in (1, 16, 64, 256, 512):
            gx = torch.Generator(device=dev).manual_seed(77 + n)            # same ids on every rank
            x = torch.randint(0, vocab, (n,), generator=gx, device=dev)
            y = ops.vocab_parallel_embedding(x, w, start, comm)
            want = reference(x)
            torch.cuda.synchronize()
            flag = torch.tensor([1 if (torch.equal(y, want) and embedding_cuda.poll_error() == 0) else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            ok = bool(flag.item())
            all_ok = all_ok and ok
            row = {"hidden": hidden, "tokens": n, "parity_ok": ok, "nvlink_push_bytes_per_owner_row": hidden * 4 * (world - 1)}
            if ok:
                out = torch.empty(n, hidden, dtype=torch.bfloat16, device=dev)

                def timed(fn):
                    for _ in range(2):
                        fn()
                    torch.cuda.synchronize()
                    dist.barrier()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        for _ in range(chain):
                            fn()
                    g.replay()
                    torch.cuda.synchronize()
                    dist.barrier()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    for _ in range(iters):
                        g.replay()
                    b.record()
                    torch.cuda.synchronize()
                    t = torch.tensor([a.elapsed_time(b) * 1e3 / (iters * chain)], dtype=torch.float64, device=dev)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    del g
                    return round(float(t.item()), 2)
                row["us_fused"] = timed(lambda: ops.vocab_parallel_embedding(x, w, start, comm, out=out))
                row["us_nccl"] = timed(lambda: reference(x))
                row["speedup"] = round(row["us_nccl"] / row["us_fused"], 2)
            rows.append(row)
        torch.cuda.synchronize()
        dist.barrier()
    return {"parity_ok": all_ok, "world": world, "timeout_latch": embedding_cuda.poll_error(),
            "timing": f"CUDA-graph chains of {chain} calls x {iters} replays, max over ranks", "rows": rows}


def run_vocab_parallel_sampler(args, dev, rank, world, ops, dist):
    """SURVEY §8f rank 4 (second half), N > 1 only: vocabulary-parallel sampling -- every rank reduces ITS
    [bs, vocab/N] logits shard and swaps one 16-byte pair per row over NVLink peer memory inside the sampler kernel --
    beside the reference's path: dist.gather of the shards to rank 0 + torch.cat (nanovllm/layers/embed_head.py:66-70)
    + Sampler.forward on rank 0 (nanovllm/layers/sampler.py:10-18; the eager reference module, and this repo's fused
    one-launch sampler as a second comparator).  Parity on the box first (greedy + sampled rows with the same injected
    Exp(1) draw: tokens equal on EVERY rank to rank 0's reference tokens), then device time per call, CUDA events,
    max over ranks."""
    from nano_vllm_learn_b200 import sampler_cuda
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    vocab = 151936
    if vocab % (4 * world):
        return {"skipped": f"vocab {vocab} not divisible by 4*{world}"}
    shard = vocab // world
    _lib = __import__("nano_vllm_learn_b200._lib", fromlist=["x"])
    _lib.check(_lib.load().nvllm_sample_sharded_set_timeout_ms(5000), "set_timeout")
    comm = TPComm.from_process_group(sampler_cuda.comm_bytes(world))
    ref_cls = _reference_sampler_class()
    rows, all_ok = [], True
    call = 0
    for bs in (1, 16, 64, 256):
        gen = torch.Generator(device=dev).manual_seed(9000 + bs)          # same full tensors on every rank
        logits = (torch.empty(bs, vocab, dtype=torch.float32, device=dev).normal_(generator=gen) * 3).to(torch.bfloat16)
        temps = torch.tensor([0.0 if i % 2 == 0 else 0.8 for i in range(bs)], dtype=torch.float32, device=dev)
        noise = torch.empty(bs, vocab, dtype=torch.float32, device=dev).exponential_(1, generator=gen)
        mine = logits[:, rank * shard:(rank + 1) * shard].contiguous()
        nz = noise[:, rank * shard:(rank + 1) * shard].contiguous()
        tok = ops.sample_tokens_sharded(mine, temps if rank == 0 else None, rank * shard, comm, noise=nz)
        # the reference path on the same data: gather to rank 0, cat, sample there
        parts = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
        dist.gather(mine, parts, 0)
        want = torch.zeros(bs, dtype=torch.int64, device=dev)
        if rank == 0:
            full = torch.cat(parts, -1)
            # sampler.py:10-18 line by line in torch, the exponential_ draw replaced by the injected one
            lf = full.to(torch.float)
            probs = torch.softmax(lf / temps.unsqueeze(1), dim=-1, dtype=torch.float)
            want = torch.where(temps == 0, lf.argmax(-1), probs.div_(noise + 1e-10).argmax(-1))
            if ref_cls is not None:                              # greedy rows against the reference class itself
                rt = ref_cls()(full.clone(), temps)
                want = torch.where(temps == 0, rt, want)
        dist.broadcast(want, 0)
        torch.cuda.synchronize()
        flag = torch.tensor([1 if (torch.equal(tok, want) and sampler_cuda.poll_error() == 0) else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item())
        all_ok = all_ok and ok

        def timed(fn, iters=30):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(iters):
                fn()
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b) * 1e3 / iters], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return round(float(t.item()), 2)

        def sharded():
            ops.sample_tokens_sharded(mine, temps if rank == 0 else None, rank * shard, comm, seed=1, offset=bs)

        def gather_then(sample):
            def fn():
                dist.gather(mine, parts, 0)
                if rank == 0:
                    sample(torch.cat(parts, -1), temps)
            return fn
        row = {"bs": bs, "vocab": vocab, "parity_ok": ok, "gather_bytes_to_rank0": bs * shard * 2 * (world - 1),
               "exchange_bytes_per_rank": bs * 16 * (world - 1)}
        if ok:
            row["us_vocab_parallel"] = timed(sharded)
            row["us_gather_fused_sampler"] = timed(gather_then(lambda l, t: ops.sample_tokens(l, t, seed=1, offset=bs)))
            if ref_cls is not None:
                ref = ref_cls()
                row["us_gather_reference_sampler"] = timed(gather_then(lambda l, t: ref(l, t)), iters=10)
                row["speedup_vs_reference"] = round(row["us_gather_reference_sampler"] / row["us_vocab_parallel"], 2)
        rows.append(row)
    torch.cuda.synchronize()
    dist.barrier()
    return {"parity_ok": all_ok, "world": world, "transport": comm.how, "timeout_latch": sampler_cuda.poll_error(),
            "timing": "eager launches back to back (the sampler is outside the engine's CUDA graphs), CUDA events, "
                      "max over ranks", "rows": rows}


def run_tp_e2e(args, world):
    """The metric BASELINE.json names beside the op-level one: bench_my decode tok/s of the reference ENGINE at
    tensor_parallel_size = N (its own spawned ranks, NCCL, scheduler, CUDA graphs), `flash` backend vs `cuda` backend
    (and, when N > 1, `cuda` with the reference's NCCL collectives instead of the fused all-reduce / embedding exchange /
    vocabulary-parallel sampler), bounded: Qwen3-0.6B random-init, bs 64, input 512, output 128.
    Runs tools/bench_my_b200.py as a subprocess of rank 0 after the torchrun ranks have released their GPUs."""
    import signal
    script = os.path.join(ROOT, "tools", "bench_my_b200.py")
    env = {k: v for k, v in os.environ.items()
           if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "GROUP_RANK", "ROLE_RANK",
                        "LOCAL_WORLD_SIZE", "ROLE_WORLD_SIZE", "TORCHELASTIC_RUN_ID", "TORCHELASTIC_RESTART_COUNT",
                        "TORCHELASTIC_MAX_RESTARTS", "TORCHELASTIC_USE_AGENT_STORE", "TORCH_NCCL_ASYNC_ERROR_HANDLING",
                        "TORCHELASTIC_ERROR_FILE", "OMP_NUM_THREADS")}
    arms = [("flash", {}), ("cuda", {})]
    if world > 1:   # the cuda attention backend with every collective left to the reference's NCCL calls
        arms.append(("cuda", {"NVLLM_FUSED_ALLREDUCE": "0", "NVLLM_VOCAB_PARALLEL_SAMPLER": "0", "NVLLM_FUSED_EMBEDDING": "0"}))
    out = {"workload": f"reference engine, Qwen3-0.6B random-init, tensor_parallel_size={world}, bs 64, input 512, "
                       f"output 128, temperature 0, ignore_eos (bench_my protocol, tok/s = bs*out/wall)", "arms": []}
    deadline = time.time() + args.tp_e2e_timeout
    for i, (backend, extra) in enumerate(arms):
        left = deadline - time.time()
        label = backend + ("+nccl_collectives" if extra else ("+fused_collectives" if backend == "cuda" and world > 1 else ""))
        if left < 45:
            out["arms"].append({"arm": label, "skipped": "time budget of the tp_e2e leg spent"})
            continue
        try:   # a crashed engine leaves its control-plane segment behind (model_runner.py:52)
            from multiprocessing import shared_memory
            shm = shared_memory.SharedMemory(name="nanovllm")
            shm.close()
            shm.unlink()
        except Exception:
            pass
        res_path = os.path.join(tempfile.gettempdir(), f"nvllm_tp_e2e_{os.getpid()}_{i}.json")
        cmd = [sys.executable, script, "--attn-backend", backend, "--batch-sizes", "64", "--input-lens", "512",
               "--output-len", "128", "--tensor-parallel-size", str(world), "--port", str(2400 + 7 * i), "--out", res_path]
        e = dict(env)
        e.update(extra)
        t0 = time.time()
        proc = subprocess.Popen(cmd, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                                start_new_session=True)
        try:
            log, _ = proc.communicate(timeout=left)
        except subprocess.TimeoutExpired:
            try:
                os.killpg(proc.pid, signal.SIGKILL)
            except Exception:
                proc.kill()
            log, _ = proc.communicate()
            out["arms"].append({"arm": label, "failed": "timeout", "log_tail": (log or "")[-300:]})
            continue
        row = {"arm": label, "wall_s": round(time.time() - t0, 1)}
        try:
            with open(res_path) as f:
                r = json.load(f)["results"][backend]["64x512"]
            row.update(tok_s=r["tok_s"], ms=r["ms"])
            os.unlink(res_path)
        except Exception as ex:
            row.update(failed=f"rc={proc.returncode} {type(ex).__name__}", log_tail=(log or "")[-400:])
        out["arms"].append(row)
    good = {a["arm"]: a["tok_s"] for a in out["arms"] if "tok_s" in a}
    if "flash" in good:
        for k2, v in good.items():
            if k2 != "flash":
                out[f"{k2}_over_flash"] = round(v / good["flash"], 3)
    return out


def _reference_sampler_class():
    """The reference's UNMODIFIED Sampler class (nanovllm/layers/sampler.py), loaded from its source file in
    baseline/_ref so that the nanovllm package (engine, flash-attn, triton) is not imported; None if absent."""
    import importlib.util
    path = os.path.join(ROOT, "baseline", "_ref", "nanovllm", "layers", "sampler.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("_ref_sampler_bench", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.Sampler


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
