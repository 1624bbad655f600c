"""Seeded synthetic inputs shared by the golden generator, the CPU tests and the GPU parity tests.

Inputs are regenerated from seeds (fixtures stay small); every fixture stores a sha256 of the
regenerated inputs so an RNG change shows up as "fixture inputs differ", not as a parity failure.
"""
from __future__ import annotations

import hashlib

import torch


def sha(*tensors) -> str:
    h = hashlib.sha256()
    for t in tensors:
        if t is None:
            continue
        t = t.contiguous()
        if t.dtype == torch.bfloat16:
            t = t.view(torch.int16)
        h.update(t.numpy().tobytes())
    return h.hexdigest()


def bf16_bits(t: torch.Tensor):
    return t.contiguous().view(torch.int16).numpy()


def from_bits(a) -> torch.Tensor:
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


# name -> (seed, ctx lens, Hq, Hkv, num_blocks, block_size)
DECODE_CASES = {
    # BASELINE.json configs[0]: bs=1, ctx=512, 16 q / 8 kv heads, head_dim 128, block 256
    "c1_bs1_ctx512": (0, [512], 16, 8, 4, 256),
    "ragged_g4": (1, [1, 255, 700, 256, 257], 8, 2, 12, 256),
    "ragged_g2_tail": (2, [17, 513, 31, 300], 4, 2, 10, 256),
    "g1_single_head": (3, [129, 64], 2, 2, 4, 256),
    "g8": (4, [333, 16], 8, 1, 6, 256),
}

# name -> (seed, lens, Hq, Hkv)   (no prefix: Lq == Lk)
PREFILL_CASES = {
    "p_small": (10, [5, 64, 130], 4, 2),
    "p_tiles": (11, [63, 65, 128, 1], 4, 1),
    "p_qwen06_heads": (12, [200], 16, 8),
}

# name -> (seed, [(cached, new)], Hq, Hkv, num_blocks)   prefix-cache prefill from the paged cache
PREFIX_CASES = {
    "px_shared512": (20, [(0, 300), (512, 300), (256, 70)], 4, 2, 16),
    "px_one_token": (21, [(256, 1), (512, 64)], 8, 2, 12),
    "px_g4_long": (22, [(768, 129)], 8, 2, 8),
}

D = 128


def make_decode_case(name: str):
    seed, lens, hq, hkv, nblk, bsz = DECODE_CASES[name]
    g = torch.Generator().manual_seed(seed)
    bs = len(lens)
    q = torch.randn(bs, hq, D, generator=g).to(torch.bfloat16)
    k_cache = torch.randn(nblk, bsz, hkv, D, generator=g).to(torch.bfloat16)
    v_cache = torch.randn(nblk, bsz, hkv, D, generator=g).to(torch.bfloat16)
    need = [(c + bsz - 1) // bsz for c in lens]
    width = max(need)
    perm = torch.randperm(nblk, generator=g).tolist()
    assert sum(need) <= nblk
    tables = torch.full((bs, width), -1, dtype=torch.int32)
    pos = 0
    for b, n in enumerate(need):
        for j in range(n):
            tables[b, j] = perm[pos]
            pos += 1
    ctx = torch.tensor(lens, dtype=torch.int32)
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, context_lens=ctx, block_tables=tables,
                scale=D ** -0.5, hq=hq, hkv=hkv, block_size=bsz)


def make_prefill_case(name: str):
    seed, lens, hq, hkv = PREFILL_CASES[name]
    g = torch.Generator().manual_seed(seed)
    total = sum(lens)
    q = torch.randn(total, hq, D, generator=g).to(torch.bfloat16)
    k = torch.randn(total, hkv, D, generator=g).to(torch.bfloat16)
    v = torch.randn(total, hkv, D, generator=g).to(torch.bfloat16)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    return dict(q=q, k=k, v=v, cu_seqlens_q=cu, cu_seqlens_k=cu.clone(), max_seqlen_q=max(lens),
                max_seqlen_k=max(lens), scale=D ** -0.5, hq=hq, hkv=hkv)


def make_prefix_case(name: str, block_size: int = 256):
    """Prefill with cached prefixes: all keys (cached + new) already sit in the paged cache."""
    seed, pairs, hq, hkv, nblk = PREFIX_CASES[name]
    g = torch.Generator().manual_seed(seed)
    lq = [n for _, n in pairs]
    lk = [c + n for c, n in pairs]
    q = torch.randn(sum(lq), hq, D, generator=g).to(torch.bfloat16)
    k_cache = torch.randn(nblk, block_size, hkv, D, generator=g).to(torch.bfloat16)
    v_cache = torch.randn(nblk, block_size, hkv, D, generator=g).to(torch.bfloat16)
    need = [(c + block_size - 1) // block_size for c in lk]
    assert sum(need) <= nblk
    perm = torch.randperm(nblk, generator=g).tolist()
    tables = torch.full((len(pairs), max(need)), -1, dtype=torch.int32)
    pos = 0
    for b, n in enumerate(need):
        for j in range(n):
            tables[b, j] = perm[pos]
            pos += 1
    cu_q = torch.tensor([0] + list(torch.tensor(lq).cumsum(0)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lk).cumsum(0)), dtype=torch.int32)
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, cu_seqlens_q=cu_q, cu_seqlens_k=cu_k,
                max_seqlen_q=max(lq), max_seqlen_k=max(lk), block_tables=tables, scale=D ** -0.5,
                hq=hq, hkv=hkv, block_size=block_size)


def make_store_case(seed: int, n: int, hkv: int, nblk: int, block_size: int = 256, strided_v: bool = True,
                    dup_slot0: int = 0, neg: int = 0):
    """Store inputs shaped like the model's: K contiguous, V a strided view of a fused qkv row."""
    g = torch.Generator().manual_seed(seed)
    row = hkv * D
    key = torch.randn(n, hkv, D, generator=g).to(torch.bfloat16)
    if strided_v:
        fused = torch.randn(n, 3 * row, generator=g).to(torch.bfloat16)
        value = fused[:, 2 * row:].view(n, hkv, D)
    else:
        value = torch.randn(n, hkv, D, generator=g).to(torch.bfloat16)
    slots = torch.randperm(nblk * block_size, generator=g)[:n].to(torch.int32)
    if dup_slot0:  # zero-padded graph rows: slot 0 repeated with identical data (model_runner.py:291-293)
        slots[-dup_slot0:] = 0
        key[-dup_slot0:] = key[-1]
        value[-dup_slot0:] = value[-1]
    if neg:
        slots[:neg] = -1
    k_cache = torch.randn(nblk, block_size, hkv, D, generator=g).to(torch.bfloat16)
    v_cache = torch.randn(nblk, block_size, hkv, D, generator=g).to(torch.bfloat16)
    return dict(key=key, value=value, slot_mapping=slots, k_cache=k_cache, v_cache=v_cache)


def make_pre_attention_case(seed: int = 30, n: int = 37, hq: int = 16, hkv: int = 8, nblk: int = 3, block_size: int = 256):
    """Inputs of the fused pre-attention step: fused qkv rows, norm weights, positions, slots."""
    g = torch.Generator().manual_seed(seed)
    qkv = (torch.randn(n, (hq + 2 * hkv) * D, generator=g) * 1.5).to(torch.bfloat16)
    positions = torch.randint(0, 4000, (n,), generator=g)
    q_weight = (1 + 0.1 * torch.randn(D, generator=g)).to(torch.bfloat16)
    k_weight = (1 + 0.1 * torch.randn(D, generator=g)).to(torch.bfloat16)
    slots = torch.randperm(nblk * block_size, generator=g)[:n].to(torch.int32)
    return dict(qkv=qkv, positions=positions, q_weight=q_weight, k_weight=k_weight, eps=1e-6, slots=slots,
                hq=hq, hkv=hkv, nblk=nblk, block_size=block_size, rope_base=1000000.0, max_position=4096)


def rope_cache(base: float, max_position: int):
    """cos_sin_cache exactly as RotaryEmbedding.__init__ builds it (rotary_embedding.py:30-37)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, D, 2, dtype=torch.float) / D))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


# name -> (seed, bs, vocab, dtype, temperatures)   token sampler (reference sampler.py)
SAMPLER_CASES = {
    "s_greedy_bf16": (40, 5, 151936, torch.bfloat16, [0.0, 0.0, 0.0, 0.0, 0.0]),
    "s_mixed_bf16": (41, 6, 4099, torch.bfloat16, [0.0, 0.7, 1.0, 0.0, 1.3, 0.5]),
    "s_mixed_fp32": (42, 4, 1000, torch.float32, [1.0, 0.0, 0.25, 2.0]),
    "s_peaked": (43, 8, 2048, torch.bfloat16, [0.6] * 8),
}


def make_sampler_case(name: str):
    """Logits with deliberate exact ties at the maximum (bf16 rounding makes them common anyway); `s_peaked`
    scales the logits so the sampled token is not always the noise's favourite."""
    seed, bs, vocab, dtype, temps = SAMPLER_CASES[name]
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(bs, vocab, generator=g) * (6.0 if name == "s_peaked" else 2.0)
    logits = logits.to(dtype)
    # rows 0 and 1: the row maximum appears three times; the lowest index must win
    for r in range(min(2, bs)):
        top = logits[r].float().max().to(dtype)
        idx = torch.randint(0, vocab, (3,), generator=g)
        logits[r, idx] = top
    return dict(logits=logits, temperatures=torch.tensor(temps, dtype=torch.float32), noise_seed=1000 + seed)


# ---- vocabulary-parallel lm_head + sampler (second half) -----------------------------------
# name -> (seed, tokens in x, hidden, vocab, world, temperatures per sampled row, prefill cu_seqlens_q or None)
LM_HEAD_CASES = {
    "vp_decode_tp2": (50, 6, 64, 4096, 2, [0.0, 0.0, 0.8, 1.0, 0.0, 0.5], None),
    "vp_decode_tp4": (51, 5, 128, 2048, 4, [0.0, 0.7, 0.0, 1.2, 0.0], None),
    "vp_prefill_tp2": (52, 11, 64, 1024, 2, [0.0, 0.9, 0.0], [0, 4, 5, 11]),     # last token of 3 sequences
    "vp_greedy_tp8": (53, 4, 32, 1024, 8, [0.0, 0.0, 0.0, 0.0], None),
}


def make_lm_head_case(name: str):
    """Hidden states + an lm_head weight; rows 0 and 1 of the weight matrix get duplicates in OTHER shards so that
    exact logit ties across ranks occur (the lowest global index must win)."""
    seed, n, hidden, vocab, world, temps, cu = LM_HEAD_CASES[name]
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, hidden, generator=g).to(torch.bfloat16)
    w = (torch.randn(vocab, hidden, generator=g) * (hidden ** -0.5)).to(torch.bfloat16)
    # make the greedy winner of row 0 appear again in the last shard (same weight row -> identical logit)
    logits0 = (x[0 if cu is None else cu[1] - 1].float() @ w.float().T)
    top = int(logits0.argmax())
    w[vocab - 3] = w[top]
    return dict(x=x, weight=w, world=world, temperatures=torch.tensor(temps, dtype=torch.float32),
                cu_seqlens_q=None if cu is None else torch.tensor(cu, dtype=torch.int32), noise_seed=2000 + seed)


# ---- fused VocabParallelEmbedding (third collective) ----------------------------------------
# name -> (seed, n_tokens, vocab, hidden, world)
EMBEDDING_CASES = {
    "emb_tp2": (60, 16, 4096, 256, 2),
    "emb_tp4_ragged": (61, 5, 2048, 128, 4),
    "emb_tp8": (62, 33, 1024, 64, 8),
    "emb_one_owner_tp2": (63, 8, 1024, 1024, 2),      # every token in rank 1's shard: rank 1 waits for nobody
}


def make_embedding_case(name: str):
    seed, n, vocab, hidden, world = EMBEDDING_CASES[name]
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(vocab, hidden, generator=g).to(torch.bfloat16)
    if name == "emb_one_owner_tp2":
        x = torch.randint(vocab // 2, vocab, (n,), generator=g)
    else:
        x = torch.randint(0, vocab, (n,), generator=g)
        x[0], x[-1] = 0, vocab - 1                     # shard edges
    return dict(x=x.to(torch.int64), weight=w, world=world)


# ---- fused RowParallelLinear -------------------------------------------------------------
# name -> (seed, n_tokens, in_features (all ranks), out_features, world, bias)
ROWPARALLEL_CASES = {
    "rp_qwen06_oproj_tp2": (40, 16, 2048, 1024, 2, False),     # Qwen3-0.6B o_proj, tp=2, decode bs 16
    "rp_qwen06_down_tp2": (41, 5, 3072, 1024, 2, False),       # down_proj, ragged token count
    "rp_bias_tp2": (42, 33, 512, 128, 2, True),                # bias added once (rank 0 in the reference)
    "rp_tp4": (43, 7, 1024, 256, 4, False),
}


def make_rowparallel_case(name: str):
    seed, n, k, h, world, with_bias = ROWPARALLEL_CASES[name]
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, k, generator=g).to(torch.bfloat16)
    w = (torch.randn(h, k, generator=g) * (k ** -0.5)).to(torch.bfloat16)
    b = torch.randn(h, generator=g).to(torch.bfloat16) if with_bias else None
    return dict(x=x, weight=w, bias=b, world=world, n=n, k=k, h=h)
