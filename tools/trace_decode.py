"""Where does a small paged-decode launch spend its time?  %globaltimer stamps from every CTA of the fixed-grid kernel
(tools build of the library: `python nano-vllm-learn_b200/build.py --tools`).

    python tools/trace_decode.py --bs 1 --ctx 32768 --hq 32 --hkv 8 --splits 18 [--variant 2]

Prints, relative to the earliest CTA entry of the launch: the spread of CTA entries (grid ramp), the time to the first
copies issued (context_lens + table lookups), to the first chunk landed (HBM latency under load), the end of the main
loops, the partial publish (merge + fence + arrival atomic) and the exit of the combining CTAs.  The launch sits in the
middle of a stream of identical launches on rotating inputs, as a decoder layer's launch does."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch  # noqa: E402

from nano_vllm_learn_b200 import _lib  # noqa: E402

lib = _lib.load_tools()   # python nano-vllm-learn_b200/build.py --tools
from nano_vllm_learn_b200 import ops  # noqa: E402
from sweep_decode import make, share_table  # noqa: E402

D, BLOCK = 128, 256


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bs", type=int, default=1)
    ap.add_argument("--ctx", type=int, default=32768)
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--width", type=int, default=160)
    ap.add_argument("--splits", type=int, default=18)
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--cluster", type=int, default=0)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    lib.nvllm_paged_decode_set_variant(a.variant)
    lib.nvllm_paged_decode_set_cluster(a.cluster)
    one = a.bs * a.ctx * a.hkv * D * 4
    ncopy = max(2, min(16, -(-(1024 << 20) // one)))
    sets = share_table([make(a.bs, a.ctx, a.hq, a.hkv, dev, a.width, seed=c) for c in range(ncopy)])
    lens = sets[0][3]
    out = torch.empty_like(sets[0][0])
    ws = torch.zeros(max(ops.decode_workspace_bytes(a.bs, a.hq, D, a.splits), 16), dtype=torch.uint8, device=dev)
    n_cta = a.bs * a.hkv * a.splits
    trace = torch.zeros(n_cta * 8, dtype=torch.int64, device=dev)

    def launch(i):
        q, kc, vc, _, tab, _ = sets[i % ncopy]
        ops.flash_attn_with_kvcache(q, kc, vc, lens, tab, softmax_scale=D ** -0.5, num_splits=a.splits, out=out, workspace=ws)

    for i in range(6):
        launch(i)
    torch.cuda.synchronize()
    rows = []
    for rep in range(5):
        trace.zero_()
        torch.cuda.synchronize()
        torch.cuda._sleep(1_000_000)
        for i in range(3):
            launch(i)
        assert lib.nvllm_paged_decode_set_trace(trace.data_ptr()) == 0     # synchronous symbol copy: stream is drained
        launch(3)
        torch.cuda.synchronize()
        assert lib.nvllm_paged_decode_set_trace(None) == 0
        t = trace.view(n_cta, 8).cpu().double()
        t0 = t[:, 0].min()
        rel = (t - t0) / 1e3      # microseconds
        fin = t[:, 5] > 0
        row = {"entry_last": rel[:, 0].max().item(), "issued_med": rel[:, 1].median().item(), "issued_last": rel[:, 1].max().item(),
               "first_chunk_med": rel[:, 2].median().item(), "first_chunk_last": rel[:, 2].max().item(),
               "loop_done_first": rel[:, 3].min().item(), "loop_done_med": rel[:, 3].median().item(),
               "loop_done_last": rel[:, 3].max().item(), "published_last": rel[:, 4].max().item() if a.splits > 1 else None,
               "exit_last": rel[fin, 5].max().item() if fin.any() else None}
        rows.append({k: (round(v, 2) if v is not None else None) for k, v in row.items()})
    # a launch behind its predecessor: entry times are spread because CTAs of the previous launch still hold SMs
    print(json.dumps(dict(shape=dict(bs=a.bs, ctx=a.ctx, hq=a.hq, hkv=a.hkv, splits=a.splits, variant=a.variant, cluster=a.cluster, ctas=n_cta),
                          us_since_first_cta_entry=rows)))


if __name__ == "__main__":
    main()
