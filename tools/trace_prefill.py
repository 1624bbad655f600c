"""Timeline of the tcgen05 prefill kernel's protocol (debug build): clock64() stamps of the MMA issuer and of
row 0 of both softmax warpgroups for CTA 0's first work item.  Prints cycles relative to the first stamp."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from nano_vllm_learn_b200 import _lib, ops  # noqa: E402

dev = "cuda:0"
lib = _lib.load_tools()   # python nano-vllm-learn_b200/build.py --tools
L = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
MODE = int(sys.argv[2]) if len(sys.argv) > 2 else 0  # bit0 no exp/P stores, bit1 no S loads, bit2 no K/V TMA
hq, hkv, d = 32, 8, 128
torch.manual_seed(0)
q = torch.randn(L, hq, d, device=dev).to(torch.bfloat16)
k = torch.randn(L, hkv, d, device=dev).to(torch.bfloat16)
v = torch.randn(L, hkv, d, device=dev).to(torch.bfloat16)
cu = torch.tensor([0, L], dtype=torch.int32, device=dev)
dbg = torch.zeros(32768 + 3 * 64 * 8 + 8, dtype=torch.float32, device=dev)
dbg.view(torch.int32)[32768 + 3 * 64 * 8] = MODE
print(f"L={L} debug mode={MODE}")
ops.set_prefill_variant(1)
for _ in range(2):
    ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=d ** -0.5)
lib.nvllm_varlen_prefill_set_debug_buffer(dbg.data_ptr())
ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=d ** -0.5)
torch.cuda.synchronize()
lib.nvllm_varlen_prefill_set_debug_buffer(None)
tr = dbg[32768:32768 + 3 * 64 * 8].view(torch.int32).view(3, 64, 8).cpu().long() & 0xFFFFFFFF
t0 = int(tr[0, 0, 0])
names = ["MMA issuer  [top, PV_A issued, QK_A(next) issued, PV_B issued, QK_B(next) issued]",
         "softmax A   [top, S ready, S in regs, max/rescale done, exp+P stores issued, P arrived]",
         "softmax B   [same]"]
n = {0: 5, 1: 6, 2: 6}
for role in range(3):
    print(names[role])
    prev = None
    for j in range(4, 14):
        row = [int(tr[role, j, kk]) - t0 for kk in range(n[role])]
        delta = "" if prev is None else f"   (+{row[0] - prev} since previous tile)"
        print(f"  tile {j:2d}: " + " ".join(f"{x:7d}" for x in row) + delta)
        prev = row[0]
