"""B200-native paged-attention hot path for nano-vllm (`--attn-backend cuda`).

The product is the sm_100a CUDA library behind `include/nvllm_attn.h`; this package is the thin
Python host side that mirrors the reference's operator interface
(`nanovllm/layers/attention.py`, `nanovllm/utils/context.py`, the prepare/capture glue of
`nanovllm/engine/model_runner.py`).  There is no CPU or eager fallback: importing `ops` without a
built library raises.
"""
from .context import Context, get_context, set_context, reset_context  # noqa: F401

__all__ = ["Context", "get_context", "set_context", "reset_context"]
