"""Drop-in for the reference `Sampler` module (nanovllm/layers/sampler.py:5-18) on the fused sm_100a kernel.

Same constructor, same `forward(logits, temperatures) -> int64 token ids`; one launch instead of eight eager
kernels and no [bs, vocab] fp32 temporaries.  Greedy rows (temperature == 0, what bench_my uses) are bit-exact
with the reference; sampled rows draw the same distribution (Gumbel-max over `logits / T`) from a Philox stream
keyed by torch's seed and a per-call offset instead of `Tensor.exponential_`.

Vocabulary-parallel mode (tensor_parallel_size > 1, INTEGRATION.md edit A8): `ParallelLMHead.forward` keeps its
`[bs, vocab/tp]` logits shard on every rank instead of gathering them to rank 0 over NCCL
(nanovllm/layers/embed_head.py:66-70), and `ModelRunner.run` (nanovllm/engine/model_runner.py:306-313) samples with
`Sampler.forward_sharded`: each rank reduces its shard to one (key, global index) pair per row and the pairs are
swapped through NVLink peer memory inside the same kernel (csrc/sampler.cu).  No NCCL kernel runs after the last
decoder layer, and every rank holds the sampled token ids.  No CPU path.
"""
from __future__ import annotations

import ctypes
import os

import torch
from torch import nn

from . import _lib, ops
from .rowparallel_cuda import TPComm

MAX_BS = 512      # the engine's max_num_seqs ceiling for graphs (model_runner.py:325); comm sized for it


def comm_bytes(world: int, max_bs: int = MAX_BS) -> int:
    n = int(_lib.load().nvllm_sample_comm_bytes(world, max_bs))
    if n == 0:
        raise ValueError(f"vocabulary-parallel sampler: unsupported world={world} / max_bs={max_bs}")
    return n


def poll_error(clear: bool = False) -> int:
    """Timeout latch of the vocabulary-parallel exchange (0 = none, else 1 + row): a mapped host word, no CUDA call."""
    err = ctypes.c_uint32()
    _lib.check(_lib.load().nvllm_sample_sharded_poll_error(ctypes.byref(err), 1 if clear else 0), "sample_sharded_poll_error")
    return err.value


class Sampler(nn.Module):

    def __init__(self):
        super().__init__()
        self._calls = 0
        self._comm: TPComm | None = None
        self._host_tokens: torch.Tensor | None = None
        self._dev_tokens: torch.Tensor | None = None
        # int64 ids of the most recent call, still on the device: the runner glue feeds them to the next decode step
        # as input_ids without a host round trip (runner_glue.RunnerGlue.prepare_decode)
        self.last_tokens: torch.Tensor | None = None

    def _out(self, bs: int, device) -> torch.Tensor:
        if self._dev_tokens is None or self._dev_tokens.numel() < bs or self._dev_tokens.device != device:
            self._dev_tokens = torch.empty(max(bs, MAX_BS), dtype=torch.int64, device=device)
        return self._dev_tokens[:bs]

    @torch.inference_mode()
    def forward(self, logits: torch.Tensor, temperatures: torch.Tensor, noise: torch.Tensor | None = None):
        self._calls += 1
        self.last_tokens = ops.sample_tokens(logits, temperatures, noise=noise, seed=torch.initial_seed(),
                                             offset=self._calls, out=self._out(logits.shape[0], logits.device))
        return self.last_tokens

    # ---- vocabulary-parallel mode ---------------------------------------------------------------------------------
    def bind_comm(self, comm: TPComm) -> None:
        self._comm = comm

    def _get_comm(self) -> TPComm:
        if self._comm is None:
            import torch.distributed as dist
            self._comm = TPComm.from_process_group(comm_bytes(dist.get_world_size()))
        return self._comm

    @torch.inference_mode()
    def forward_sharded(self, logits_shard: torch.Tensor, temperatures: torch.Tensor | None,
                        noise_shard: torch.Tensor | None = None, to_host: bool = False):
        """logits_shard [bs, vocab/tp] = this rank's columns of the full logits (rank r owns [r*V/tp, (r+1)*V/tp),
        embed_head.py:21-23).  Every rank must call with the same bs in the same order; rank 0 supplies the
        temperatures.  Returns int64 token ids [bs] on every rank; `to_host=True` writes them straight into pinned host
        memory through its device mapping and returns that host tensor after a stream synchronise (what replaces the
        `.tolist()` device read at model_runner.py:311)."""
        comm = self._get_comm()
        self._calls += 1
        bs, shard = logits_shard.shape
        host = None
        if to_host:
            if self._host_tokens is None or self._host_tokens.numel() < bs:
                self._host_tokens = torch.empty(max(bs, MAX_BS), dtype=torch.int64, device="cpu", pin_memory=True)
            host = self._host_tokens
        tokens = ops.sample_tokens_sharded(logits_shard, temperatures, comm.rank * shard, comm, noise=noise_shard,
                                           seed=torch.initial_seed(), offset=self._calls,
                                           out=self._out(bs, logits_shard.device), host_out=host)
        self.last_tokens = tokens
        if to_host:
            tokens = host
            torch.cuda.current_stream().synchronize()
            err = poll_error()
            if err:
                poll_error(clear=True)
                raise _lib.NvllmError(f"vocabulary-parallel sampler: a tensor-parallel peer never arrived (row {err - 1}); "
                                      f"NVLLM_VOCAB_PARALLEL_SAMPLER=0 restores the NCCL gather")
            return tokens[:bs]
        return tokens


# ---------------------------------------------------------------------- engine integration (edit A8)
_VP_ENABLED = False


def configure(cuda_backend: bool) -> None:
    """Called by ModelRunner.__init__ before the model is built (next to rowparallel_cuda.configure).
    NVLLM_VOCAB_PARALLEL_SAMPLER: "1" on, "0" the reference's gather + rank-0 sampler; unset: on with the cuda
    backend's fused sampler (NVLLM_FUSED_SAMPLER != 0) whenever tensor_parallel_size > 1."""
    global _VP_ENABLED
    env = os.environ.get("NVLLM_VOCAB_PARALLEL_SAMPLER")
    fused = os.environ.get("NVLLM_FUSED_SAMPLER", "1") != "0"
    _VP_ENABLED = bool(cuda_backend) and fused and (env != "0")
    tmo = os.environ.get("NVLLM_FUSED_ALLREDUCE_TIMEOUT_MS")
    if tmo is not None and int(tmo) > 0:
        _lib.check(_lib.load().nvllm_sample_sharded_set_timeout_ms(int(tmo)), "sample_sharded_set_timeout_ms")


def vocab_parallel_enabled() -> bool:
    return _VP_ENABLED


def install_lm_head(ParallelLMHead) -> None:
    """Patch the reference class in place: with vocabulary-parallel sampling on and tp_size > 1, forward() returns
    this rank's logits shard on EVERY rank and launches no collective (embed_head.py:61-71 otherwise unchanged)."""
    ref_forward = ParallelLMHead.forward
    get_context = ref_forward.__globals__["get_context"]
    F = torch.nn.functional

    def forward(self, x: torch.Tensor):
        if not (_VP_ENABLED and self.tp_size > 1):
            return ref_forward(self, x)
        context = get_context()
        if context.is_prefill:
            last_indices = context.cu_seqlens_q[1:] - 1
            x = x[last_indices].contiguous()
        return F.linear(x, self.weight, self.bias)

    ParallelLMHead.forward = forward


def install_runner(model_runner_cls, reset_context) -> None:
    """Patch ModelRunner.run (model_runner.py:306-313): every rank samples on its own shard, rank 0 alone knows the
    temperatures and alone returns token ids, as in the reference."""
    ref_run = model_runner_cls.run

    def run(self, seqs, is_prefill: bool):
        if not (_VP_ENABLED and self.world_size > 1 and isinstance(self.sampler, Sampler)):
            return ref_run(self, seqs, is_prefill)
        input_ids, positions = self.prepare_prefill(seqs) if is_prefill else self.prepare_decode(seqs)
        temperatures = self.prepare_sample(seqs) if self.rank == 0 else None
        logits = self.run_model(input_ids, positions, is_prefill)
        tokens = self.sampler.forward_sharded(logits, temperatures, to_host=self.rank == 0)
        token_ids = tokens.tolist() if self.rank == 0 else None
        reset_context()
        return token_ids

    model_runner_cls.run = run
