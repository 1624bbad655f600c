"""Materialise the reference-side binding of INTEGRATION.md on a COPY of the installed reference.

    python tools/make_integration.py [--src baseline/_ref] [--dst /tmp/nvllm_integration/nanovllm_cuda]

Copies `<src>/nanovllm` to `<dst>/nanovllm` (under the system temp directory by default, never inside the repository) and applies exactly the
edits a maintainer would make to admit `--attn-backend cuda` (A1-A4, and the optional fusions A5-A9), plus two environment-compat
edits (C1-C2) that have nothing to do with the backend but are needed for the unmodified reference
to construct under this image's transformers 5.5 and network setup.  Every edit is an
anchored string replacement that must match exactly once; the flash/triton/sdpa code paths are not
touched, so running the copy with `attn_backend="flash"` is the reference's stock path.
"""
from __future__ import annotations

import argparse
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# the generated copy lives OUTSIDE the repository (it is the reference plus a dozen edited lines)
DEFAULT_DST = os.path.join(__import__("tempfile").gettempdir(), "nvllm_integration", "nanovllm_cuda")

EDITS = [
    # ---- A1: admit the backend name (nanovllm/config.py:6)
    ("nanovllm/config.py",
     'VALID_ATTN_BACKENDS = ("flash", "sdpa", "sdpa.math", "triton")',
     'VALID_ATTN_BACKENDS = ("flash", "sdpa", "sdpa.math", "triton", "cuda")'),
    # ---- A2: string -> class dispatch (nanovllm/models/qwen3.py:52-56)
    ("nanovllm/models/qwen3.py",
     '        else:\n            raise ValueError(f"Unknown attention backend: {attn_backend}")',
     '        elif attn_backend == "cuda":\n'
     '            from nanovllm.layers.attention_cuda import Attention\n'
     '            attn_kwargs = {"block_size": block_size}\n'
     '        else:\n            raise ValueError(f"Unknown attention backend: {attn_backend}")'),
    # ---- A4: runner glue (prepare_* / capture / replay) for the cuda backend, appended to model_runner.py
    ("nanovllm/engine/model_runner.py", None,
     '\n\n# --- cuda attention backend: B200-native prepare/capture glue (other backends keep the methods above)\n'
     'from nano_vllm_learn_b200.runner_glue import install as _install_cuda_glue\n'
     '_install_cuda_glue(ModelRunner, set_context, get_context, reset_context, sequence_cls=Sequence)\n'),
    # ---- A5 (optional: NVLLM_FUSED_PRE_ATTN=0 keeps the reference's separate kernels):
    #      q/k RMSNorm + RoPE + store_kvcache fused into one launch before attention (nanovllm/models/qwen3.py:105-118)
    ("nanovllm/models/qwen3.py",
     '        qkv = self.qkv_proj(hidden_states)\n',
     '        qkv = self.qkv_proj(hidden_states)\n'
     '        if self.attn_backend == "cuda" and __import__("os").environ.get("NVLLM_FUSED_PRE_ATTN", "1") != "0":\n'
     '            return self.o_proj(self.attn.forward_fused(qkv, positions, self.q_norm, self.k_norm, self.rotary_emb))\n'),
    # ---- A6 (optional: NVLLM_FUSED_SAMPLER=0 keeps the reference's eager sampler):
    #      one fused kernel for Sampler.forward (nanovllm/engine/model_runner.py:38, nanovllm/layers/sampler.py:10-18)
    ("nanovllm/engine/model_runner.py",
     '        self.sampler = Sampler()\n',
     '        self.sampler = Sampler()\n'
     '        if attn_backend == "cuda" and __import__("os").environ.get("NVLLM_FUSED_SAMPLER", "1") != "0":\n'
     '            from nano_vllm_learn_b200.sampler_cuda import Sampler as _CudaSampler\n'
     '            self.sampler = _CudaSampler()\n'),
    # ---- A7 (optional: NVLLM_FUSED_ALLREDUCE=0 keeps F.linear + NCCL all_reduce):
    #      RowParallelLinear GEMM + all-reduce in one launch over NVLink peer memory when tensor_parallel_size > 1
    #      (nanovllm/layers/linear.py:185-190; o_proj qwen3.py:118, down_proj qwen3.py:147), appended to linear.py
    ("nanovllm/layers/linear.py", None,
     '\n\n# --- cuda backend: fused RowParallelLinear (GEMM + tensor-parallel all-reduce, one sm_100a kernel)\n'
     'from nano_vllm_learn_b200.rowparallel_cuda import install as _install_fused_rowparallel\n'
     '_install_fused_rowparallel(RowParallelLinear)\n'),
    ("nanovllm/engine/model_runner.py",
     '        self.model = Qwen3ForCausalLM(hf_config)\n',
     '        from nano_vllm_learn_b200 import rowparallel_cuda as _fused_rowparallel   # A7: only with the cuda backend\n'
     '        _fused_rowparallel.configure(attn_backend == "cuda")\n'
     '        self.model = Qwen3ForCausalLM(hf_config)\n'),
    # ---- A8 (second half; optional: NVLLM_VOCAB_PARALLEL_SAMPLER=0 keeps dist.gather + rank-0 sampling):
    #      vocabulary-parallel sampling -- ParallelLMHead keeps its logits shard (nanovllm/layers/embed_head.py:66-70),
    #      ModelRunner.run samples on every rank and swaps 16 bytes per row over NVLink (nanovllm/engine/model_runner.py:306-313)
    ("nanovllm/layers/embed_head.py", None,
     '\n\n# --- cuda backend: vocabulary-parallel sampling (no logits gather)\n'
     'from nano_vllm_learn_b200.sampler_cuda import install_lm_head as _install_vocab_parallel_head\n'
     '_install_vocab_parallel_head(ParallelLMHead)\n'),
    ("nanovllm/engine/model_runner.py", None,
     'from nano_vllm_learn_b200.sampler_cuda import install_runner as _install_vocab_parallel_run\n'
     '_install_vocab_parallel_run(ModelRunner, reset_context)\n'),
    ("nanovllm/engine/model_runner.py",
     '        load_model(self.model, config.model)\n',
     '        load_model(self.model, config.model)\n'
     '        from nano_vllm_learn_b200 import sampler_cuda as _vocab_parallel   # A8: only with the cuda backend\n'
     '        _vocab_parallel.configure(attn_backend == "cuda")\n'),
    # ---- A9 (third collective; optional: NVLLM_FUSED_EMBEDDING=0 keeps masked lookup + dist.all_reduce):
    #      VocabParallelEmbedding lookup + row exchange in one launch over NVLink peer memory for <= 512 tokens
    #      (nanovllm/layers/embed_head.py:35-43)
    ("nanovllm/layers/embed_head.py", None,
     'from nano_vllm_learn_b200.embedding_cuda import install as _install_fused_embedding\n'
     '_install_fused_embedding(VocabParallelEmbedding)\n'),
    ("nanovllm/engine/model_runner.py",
     '        torch.set_default_dtype(hf_config.torch_dtype)\n',
     '        torch.set_default_dtype(hf_config.torch_dtype)\n'
     '        from nano_vllm_learn_b200 import embedding_cuda as _fused_embedding   # A9: only with the cuda backend\n'
     '        _fused_embedding.configure(attn_backend == "cuda")\n'),
    # ---- C1: transformers>=5 exposes rope parameters as a dict even for the default rope
    ("nanovllm/config.py",
     '        self.hf_config = AutoConfig.from_pretrained(self.model)\n',
     '        self.hf_config = AutoConfig.from_pretrained(self.model)\n'
     '        _rs = getattr(self.hf_config, "rope_scaling", None)\n'
     '        if isinstance(_rs, dict) and _rs.get("rope_type", _rs.get("type", "default")) == "default":\n'
     '            self.hf_config.rope_scaling = None   # compat: get_rope() asserts None and lru_caches its args\n'),
    # ---- C2: "localhost" may not resolve inside the GPU box; allow a port override for back-to-back engines
    ("nanovllm/engine/model_runner.py",
     'dist.init_process_group("nccl", "tcp://localhost:2333", world_size=self.world_size, rank=rank)',
     'dist.init_process_group("nccl", "tcp://127.0.0.1:" + __import__("os").environ.get("NVLLM_PORT", "2333"), '
     'world_size=self.world_size, rank=rank)'),
]

# ---- A3: the backend module the dispatch imports (new file nanovllm/layers/attention_cuda.py)
ATTENTION_CUDA = '''"""cuda attention backend: hand-written sm_100a kernels behind the C ABI of nvllm_attn.h."""
from nano_vllm_learn_b200.attention_cuda import Attention, bind_context
from nanovllm.utils.context import get_context

bind_context(get_context)

__all__ = ["Attention"]
'''


def make(src: str, dst: str) -> str:
    src_pkg = os.path.join(src, "nanovllm")
    if not os.path.isdir(src_pkg):
        raise SystemExit(f"{src_pkg} not found (pip --target install of the reference, see DESIGN.md)")
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    os.makedirs(dst)
    shutil.copytree(src_pkg, os.path.join(dst, "nanovllm"), ignore=shutil.ignore_patterns("__pycache__"))
    for rel, old, new in EDITS:
        path = os.path.join(dst, rel)
        text = open(path).read()
        if old is None:
            text = text + new
        else:
            if text.count(old) != 1:
                raise SystemExit(f"anchor for {rel} matched {text.count(old)} times: {old[:60]!r}")
            text = text.replace(old, new)
        open(path, "w").write(text)
    with open(os.path.join(dst, "nanovllm", "layers", "attention_cuda.py"), "w") as f:
        f.write(ATTENTION_CUDA)
    return dst


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--src", default=os.path.join(ROOT, "baseline", "_ref"))
    ap.add_argument("--dst", default=DEFAULT_DST)
    a = ap.parse_args()
    print(make(a.src, a.dst))
