"""Item-level timeline of the tcgen05 prefill kernel (tools build): what CTA 0 does between consecutive work items
(sequence, head, pair of 128-query tiles).  clock64() stamps relative to the first item's start:

    python tools/trace_prefill_items.py [batch] [len] [hq] [hkv]      (default 16 1024 16 8: the bench prefill step)

softmax threads: start | first S ready | last P handed over | O tiles final | stores issued
MMA issuer:      start | first QK issued (Q and K landed) | last PV issued"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from nano_vllm_learn_b200 import _lib, ops  # noqa: E402

dev = "cuda:0"
lib = _lib.load_tools()   # python nano-vllm-learn_b200/build.py --tools
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
L = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
hq = int(sys.argv[3]) if len(sys.argv) > 3 else 16
hkv = int(sys.argv[4]) if len(sys.argv) > 4 else 8
d = 128
torch.manual_seed(0)
q = torch.randn(B * L, hq, d, device=dev).to(torch.bfloat16)
k = torch.randn(B * L, hkv, d, device=dev).to(torch.bfloat16)
v = torch.randn(B * L, hkv, d, device=dev).to(torch.bfloat16)
cu = torch.arange(0, B * L + 1, L, dtype=torch.int32, device=dev)
dbg = torch.zeros(32768 + 5 * 64 * 8 + 8, dtype=torch.float32, device=dev)
ops.set_prefill_variant(1)
for _ in range(2):
    ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=d ** -0.5)
lib.nvllm_varlen_prefill_set_debug_buffer(dbg.data_ptr())
ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=d ** -0.5)
torch.cuda.synchronize()
lib.nvllm_varlen_prefill_set_debug_buffer(None)
tr = dbg[32768 + 4 * 64 * 8:32768 + 5 * 64 * 8].view(torch.int32).view(64, 8).cpu().long() & 0xFFFFFFFF
t0 = int(tr[0, 0])
pairs = (L + 255) // 256
print(f"batch={B} len={L} heads={hq}/{hkv}: {pairs * hq * B} items over the CTAs; CTA 0, cycles since its first item began")
print("item | softmax: start  S ready  last P  O final  stored | issuer: start  first QK  last PV | item span")
prev = None
for i in range(64):
    if int(tr[i, 0]) == 0:
        break
    row = [(int(tr[i, kk]) - t0) & 0xFFFFFFFF for kk in range(8)]
    span = "" if prev is None else f"{row[0] - prev:7d}"
    print(f"{i:4d} | " + " ".join(f"{x:8d}" for x in row[:5]) + " | " + " ".join(f"{x:8d}" for x in row[5:]) + " | " + span)
    prev = row[0]
