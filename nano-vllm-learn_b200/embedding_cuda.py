"""Fused VocabParallelEmbedding for tensor parallelism: lookup + row exchange in one sm_100a kernel over NVLink peer
memory (csrc/vocab_embedding.cu), instead of a masked lookup followed by an NCCL all-reduce of mostly-zero rows.

Drop-in for `VocabParallelEmbedding.forward` (nanovllm/layers/embed_head.py:35-43; `embed_tokens` of the model,
qwen3.py) -- the third collective of a tensor-parallel decode step, after the two RowParallelLinear all-reduces per
layer (rowparallel_cuda.py) -- SURVEY §8f rank 2.  Batches above 512 tokens (long prefill: bandwidth-, not latency-bound)
keep the reference forward.  No CPU path: CUDA tensors and the built library are required.
"""
from __future__ import annotations

import ctypes
import os

import torch

from . import _lib, ops
from .rowparallel_cuda import TPComm

_ENABLED = False
_COMMS: dict[int, TPComm] = {}      # hidden size -> communicator (one embedding per model)


def comm_bytes(world: int, hidden: int) -> int:
    n = int(_lib.load().nvllm_vocab_embedding_comm_bytes(world, ops.EMBEDDING_MAX_TOKENS, hidden))
    if n == 0:
        raise ValueError(f"fused vocab-parallel embedding: unsupported world={world} / hidden={hidden}")
    return n


def poll_error(clear: bool = False) -> int:
    err = ctypes.c_uint32()
    _lib.check(_lib.load().nvllm_vocab_embedding_poll_error(ctypes.byref(err), 1 if clear else 0), "vocab_embedding_poll_error")
    return err.value


def configure(cuda_backend: bool) -> None:
    """Called by ModelRunner.__init__ next to rowparallel_cuda.configure (INTEGRATION.md edit A9).
    NVLLM_FUSED_EMBEDDING: "0" keeps the reference's masked lookup + dist.all_reduce; default on with the cuda backend."""
    global _ENABLED
    _ENABLED = bool(cuda_backend) and os.environ.get("NVLLM_FUSED_EMBEDDING", "1") != "0"
    _COMMS.clear()
    tmo = os.environ.get("NVLLM_FUSED_ALLREDUCE_TIMEOUT_MS")
    if tmo is not None and int(tmo) > 0:
        _lib.check(_lib.load().nvllm_vocab_embedding_set_timeout_ms(int(tmo)), "vocab_embedding_set_timeout_ms")


def enabled() -> bool:
    return _ENABLED


def _comm(hidden: int) -> TPComm:
    c = _COMMS.get(hidden)
    if c is None:
        import torch.distributed as dist
        c = _COMMS[hidden] = TPComm.from_process_group(comm_bytes(dist.get_world_size(), hidden))
    return c


def install(VocabParallelEmbedding) -> None:
    """Patch the reference class in place.  ParallelLMHead overrides forward and is not affected."""
    ref_forward = VocabParallelEmbedding.forward

    def forward(self, x: torch.Tensor):
        if not (_ENABLED and self.tp_size > 1 and x.is_cuda and x.dim() == 1 and x.dtype == torch.int64 and
                0 < x.numel() <= ops.EMBEDDING_MAX_TOKENS and self.weight.dtype == torch.bfloat16 and
                self.weight.shape[1] % 4 == 0):
            return ref_forward(self, x)
        y = ops.vocab_parallel_embedding(x.contiguous(), self.weight, self.vocab_start_idx, _comm(self.weight.shape[1]))
        if not torch.cuda.is_current_stream_capturing():
            err = poll_error()          # a timeout of an EARLIER launch: one host memory read
            if err:
                poll_error(clear=True)
                raise _lib.NvllmError(f"fused vocab-parallel embedding: a tensor-parallel peer never arrived (token row "
                                      f"{err - 1}); NVLLM_FUSED_EMBEDDING=0 restores the NCCL all-reduce")
        return y

    VocabParallelEmbedding.forward = forward
