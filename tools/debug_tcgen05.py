"""Blind-debug helper for the tcgen05 prefill kernel: dumps the first raw S tile and the first raw O
partial of CTA (q-tile 0, head 0, seq 0) and compares them, and the final output, with torch."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from nano_vllm_learn_b200 import _lib, ops  # noqa: E402


def stats(name, got, want):
    err = (got - want).abs()
    print(f"{name}: max|err|={err.max().item():.4e} mean|err|={err.mean().item():.4e} "
          f"max|want|={want.abs().max().item():.3f} nan={torch.isnan(got).sum().item()}")
    return err.max().item()


def main():
    dev = "cuda:0"
    lib = _lib.load_tools()   # python nano-vllm-learn_b200/build.py --tools
    torch.manual_seed(0)
    L = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    hq, hkv, d = 4, 2, 128
    q = torch.randn(L, hq, d, device=dev).to(torch.bfloat16)
    k = torch.randn(L, hkv, d, device=dev).to(torch.bfloat16)
    v = torch.randn(L, hkv, d, device=dev).to(torch.bfloat16)
    cu = torch.tensor([0, L], dtype=torch.int32, device=dev)
    dbg = torch.zeros(2 * 128 * 128, dtype=torch.float32, device=dev)
    lib.nvllm_varlen_prefill_set_debug_buffer(dbg.data_ptr())
    scale = d ** -0.5
    ops.set_prefill_variant(0)
    ref = ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=scale)
    ops.set_prefill_variant(1)
    out = ops.flash_attn_varlen_func(q, k, v, L, cu, L, cu, softmax_scale=scale)
    torch.cuda.synchronize()
    lib.nvllm_varlen_prefill_set_debug_buffer(None)
    ops.set_prefill_variant(-1)
    n = min(L, 128)
    S = dbg[:128 * 128].view(128, 128)
    O = dbg[128 * 128:].view(128, 128)
    qf, kf, vf = q[:, 0].float(), k[:, 0].float(), v[:, 0].float()
    S_want = torch.zeros(128, 128, device=dev)
    S_want[:n, :n] = qf[:n] @ kf[:n].T
    e = stats("S tile0 (raw QK^T)", S[:n, :n], S_want[:n, :n])
    if e > 0.5:
        print("  S[0,:8]   =", S[0, :8].tolist())
        print("  want[0,:8]=", S_want[0, :8].tolist())
        print("  S[:8,0]   =", S[:8, 0].tolist())
        print("  want[:8,0]=", S_want[:8, 0].tolist())
        stats("  hypothesis S^T", S[:n, :n].T, S_want[:n, :n])
    # expected O partial of tile 0: causal mask inside the tile (shift = 0), P relative to the tile max
    mask = torch.ones(n, n, dtype=torch.bool, device=dev).tril()
    s = (S_want[:n, :n] * scale).masked_fill(~mask, float("-inf"))
    p = torch.exp(s - s.max(dim=-1, keepdim=True).values).to(torch.bfloat16).float()
    O_want = p @ vf[:n]
    e = stats("O partial tile0 (P.V)", O[:n], O_want)
    if e > 0.1:
        print("  O[0,:8]   =", O[0, :8].tolist())
        print("  want[0,:8]=", O_want[0, :8].tolist())
        print("  O[5,:8]   =", O[5, :8].tolist())
        print("  want[5,:8]=", O_want[5, :8].tolist())
        print("  O[5,64:72]   =", O[5, 64:72].tolist())
        print("  want[5,64:72]=", O_want[5, 64:72].tolist())
    stats("final out vs mma.sync kernel", out.float(), ref.float())


if __name__ == "__main__":
    main()
