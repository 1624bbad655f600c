"""HBM read-bandwidth probe: contiguous warp reads vs the decode kernel's access shape (see csrc/membench.cu)."""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from nano_vllm_learn_b200 import _lib  # noqa: E402

lib = _lib.load_tools()   # python nano-vllm-learn_b200/build.py --tools
dev = torch.device("cuda", 0)
nbytes = 2 << 30
bufs = [torch.empty(nbytes, dtype=torch.uint8, device=dev).random_(0, 255) for _ in range(2)]
sink = torch.zeros(4, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
for pattern, hkv in ((0, 1), (1, 8), (1, 4), (1, 2), (1, 1)):
    for blocks in (148 * 2, 148 * 4, 148 * 8, 148 * 16):
        ts = []
        for i in range(8):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(200_000)
            a.record()
            rc = lib.nvllm_debug_stream_read(bufs[i % 2].data_ptr(), nbytes, pattern, hkv, blocks, sink.data_ptr(), st)
            assert rc == 0
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ms = statistics.median(ts[2:])
        print(json.dumps(dict(pattern="contiguous" if pattern == 0 else f"rows256_stride{hkv * 256}", blocks=blocks,
                              ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 1))), flush=True)
