"""Synthetic Qwen3 model directory: published shape, seeded random-init weights, stub tokenizer.

    python tools/make_synthetic_model.py --shape 0.6b --out /tmp/qwen3-0.6b-synth

There is no network for checkpoints (BASELINE.json: "random-init weights").  Shapes are the
published Qwen3-0.6B / Qwen3-8B configs.  Weights ~ N(0, 0.02) in bf16, norm weights 1.
"""
from __future__ import annotations

import argparse
import json
import os

import torch

SHAPES = {
    "0.6b": dict(hidden_size=1024, num_hidden_layers=28, num_attention_heads=16, num_key_value_heads=8, head_dim=128,
                 intermediate_size=3072, vocab_size=151936, tie_word_embeddings=True),
    "8b": dict(hidden_size=4096, num_hidden_layers=36, num_attention_heads=32, num_key_value_heads=8, head_dim=128,
               intermediate_size=12288, vocab_size=151936, tie_word_embeddings=False),
}


def make_tokenizer(out: str):
    from tokenizers import Tokenizer
    from tokenizers.models import WordLevel
    from tokenizers.pre_tokenizers import Whitespace
    from transformers import PreTrainedTokenizerFast
    words = ["<unk>", "<eos>"] + [f"w{i}" for i in range(14)] + ["<pad>", "warmup", "Benchmark", ":", "hello"]
    vocab = {w: i for i, w in enumerate(words)}
    tok = Tokenizer(WordLevel(vocab, unk_token="<unk>"))
    tok.pre_tokenizer = Whitespace()
    fast = PreTrainedTokenizerFast(tokenizer_object=tok, unk_token="<unk>", eos_token="<eos>", pad_token="<pad>")
    fast.save_pretrained(out)


def make(shape: str, out: str, layers: int | None = None, seed: int = 0) -> str:
    from safetensors.torch import save_file
    cfg = dict(SHAPES[shape])
    if layers:
        cfg["num_hidden_layers"] = layers
    os.makedirs(out, exist_ok=True)
    config = dict(architectures=["Qwen3ForCausalLM"], model_type="qwen3", max_position_embeddings=40960,
                  rms_norm_eps=1e-6, rope_theta=1000000, torch_dtype="bfloat16", hidden_act="silu",
                  attention_bias=False, **cfg)
    with open(os.path.join(out, "config.json"), "w") as f:
        json.dump(config, f, indent=1)
    make_tokenizer(out)
    g = torch.Generator().manual_seed(seed)
    h, inter, hq, hkv, d = (cfg["hidden_size"], cfg["intermediate_size"], cfg["num_attention_heads"],
                            cfg["num_key_value_heads"], cfg["head_dim"])

    def w(*shape_):
        return (torch.randn(*shape_, generator=g) * 0.02).to(torch.bfloat16)

    tensors = {"model.embed_tokens.weight": w(cfg["vocab_size"], h), "model.norm.weight": torch.ones(h, dtype=torch.bfloat16)}
    if not cfg["tie_word_embeddings"]:
        tensors["lm_head.weight"] = w(cfg["vocab_size"], h)
    for i in range(cfg["num_hidden_layers"]):
        p = f"model.layers.{i}."
        tensors[p + "self_attn.q_proj.weight"] = w(hq * d, h)
        tensors[p + "self_attn.k_proj.weight"] = w(hkv * d, h)
        tensors[p + "self_attn.v_proj.weight"] = w(hkv * d, h)
        tensors[p + "self_attn.o_proj.weight"] = w(h, hq * d)
        tensors[p + "self_attn.q_norm.weight"] = torch.ones(d, dtype=torch.bfloat16)
        tensors[p + "self_attn.k_norm.weight"] = torch.ones(d, dtype=torch.bfloat16)
        tensors[p + "mlp.gate_proj.weight"] = w(inter, h)
        tensors[p + "mlp.up_proj.weight"] = w(inter, h)
        tensors[p + "mlp.down_proj.weight"] = w(h, inter)
        tensors[p + "input_layernorm.weight"] = torch.ones(h, dtype=torch.bfloat16)
        tensors[p + "post_attention_layernorm.weight"] = torch.ones(h, dtype=torch.bfloat16)
    save_file(tensors, os.path.join(out, "model.safetensors"))
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="0.6b", choices=list(SHAPES))
    ap.add_argument("--out", required=True)
    ap.add_argument("--layers", type=int, default=None, help="override layer count (tests)")
    a = ap.parse_args()
    print(make(a.shape, a.out, a.layers))
