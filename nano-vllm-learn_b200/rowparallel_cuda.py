"""Fused RowParallelLinear for tensor parallelism: GEMM + all-reduce in one sm_100a kernel over NVLink peer memory.

Drop-in for `RowParallelLinear.forward` (nanovllm/layers/linear.py:185-190; o_proj at qwen3.py:118, down_proj at
qwen3.py:147) — SURVEY §8f rank 2.  The reference launches cuBLAS, then an NCCL all-reduce on the [N, hidden]
result (two latency-bound collectives per layer in decode); here every rank's CTAs push their fp32 partial tiles
straight into the peers' inboxes and reduce what arrives, so one launch replaces both and no NCCL kernel runs.

`TPComm` owns the peer-mapped buffers: one cudaMalloc per rank, exchanged between the rank PROCESSES (the reference
spawns one process per GPU, llm_engine.py:29-36) as CUDA IPC handles through the existing process group.  Batches
above `engine_max_tokens()` (prefill: GEMM-bound, the collective is bandwidth- not latency-bound) keep the reference's
own F.linear + dist.all_reduce.  No CPU path: CUDA tensors and the built library are required.
"""
from __future__ import annotations

import ctypes
import os

import torch

from . import _lib, ops

MAX_TOKENS = 512          # nvllm_rowparallel_linear_allreduce's n_tokens limit (include/nvllm_attn.h (6))
# Batches the engine hands to the fused kernel, by weight-shard size and world size: the largest token count at which the
# fused kernel still beat cuBLAS + NCCL on B200s (bench.py fused_allreduce leg, profiles/r02_bench_n{2,4,8}*.json).
# Larger batches keep the reference's own forward.
_ENV_MAX = os.environ.get("NVLLM_FUSED_ALLREDUCE_MAX_TOKENS")


def engine_max_tokens(in_features_per_rank: int, out_features: int, world: int = 2) -> int:
    if _ENV_MAX is not None:
        return min(MAX_TOKENS, int(_ENV_MAX))
    wbytes = in_features_per_rank * out_features * 2
    if world >= 4:
        # round 2, 4 B200s (profiles/r02_bench_n4.json): every shard 1.1-1.5x at 64 tokens, 0.64-0.96x at 256 -- the
        # one-shot exchange pushes world-1 fp32 copies, so the break-even moves down as the world grows
        # 8 B200s (profiles/r02_bench_n8.json): 1.3-1.65x at 64 tokens, 0.72-1.07x at 256 on every shard
        return 64 if world >= 8 else (128 if wbytes <= (4 << 20) else 64)
    # 2 B200s, round 2 with the tcgen05 tile from 64-128 tokens up on big shards (profiles/r02_bench_n2_tcgen05_tile.json):
    # 0.6B shards 1.04-1.15x at 512 tokens; 8B-shape o_proj / down_proj 1.05x / 1.14x at 128 and 0.97x / 0.92x at 256
    return 512 if wbytes <= (4 << 20) else 128


_IPC_HANDLE_BYTES = 64


class TPComm:
    """Peer-mapped communication buffers of all tensor-parallel ranks, as seen from this rank's process."""

    def __init__(self, rank: int, world: int, nbytes: int, local_ptr: int, peer_ptrs: list[int], how: str,
                 keepalive=None):
        assert len(peer_ptrs) == world and peer_ptrs[rank] == local_ptr
        self.rank, self.world, self.nbytes, self.how = rank, world, nbytes, how
        self.local_ptr = local_ptr
        self.peer_ptrs = list(peer_ptrs)
        self.ptr_array = (ctypes.c_void_p * world)(*peer_ptrs)
        self._keepalive = keepalive

    # -- construction -------------------------------------------------------------------------------------------
    @staticmethod
    def required_bytes(world: int, shapes: list[tuple[int, int]], max_tokens: int = MAX_TOKENS) -> int:
        """Buffer size covering every (in_features_per_rank, out_features) in `shapes` (pure host arithmetic)."""
        lib = _lib.load()
        hdr, reg, ll = ctypes.c_size_t(), ctypes.c_size_t(), ctypes.c_size_t()
        max_reg = max_ll = 0
        for k, h in shapes:
            if lib.nvllm_rowparallel_comm_layout(world, max_tokens, k, h, ctypes.byref(hdr), ctypes.byref(reg),
                                                 ctypes.byref(ll)) != 0:
                raise ValueError(f"rowparallel: unsupported shape in_features_per_rank={k} out_features={h} world={world}")
            max_reg, max_ll = max(max_reg, reg.value), max(max_ll, ll.value)
        # the flag-in-data inbox sits at the END of each epoch parity, everything else at its start: sized so the
        # largest of each, taken over all shapes sharing the buffer, never overlap
        return hdr.value + 2 * (max_reg + max_ll) if shapes else 0

    @classmethod
    def from_process_group(cls, nbytes: int, group=None) -> "TPComm":
        """One process per GPU (the reference's layout): cudaMalloc here, CUDA IPC handles through `group`.
        Falls back to torch symmetric memory (also peer-mapped device memory) when legacy CUDA IPC is refused
        by the environment; raises if neither works."""
        import torch.distributed as dist
        lib = _lib.load()
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        ptr = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(_IPC_HANDLE_BYTES)
        st = lib.nvllm_tp_comm_alloc(nbytes, ctypes.byref(ptr), handle)
        mine = (bytes(handle.raw), torch.cuda.current_device(), os.getpid()) if st == 0 else None
        gathered = [None] * world
        dist.all_gather_object(gathered, mine, group=group)
        ok = all(g is not None for g in gathered)
        ptrs = [0] * world
        if ok:
            for r, (h, _dev, _pid) in enumerate(gathered):
                if r == rank:
                    ptrs[r] = ptr.value
                    continue
                peer = ctypes.c_void_p()
                if lib.nvllm_tp_comm_open(ctypes.create_string_buffer(h, _IPC_HANDLE_BYTES), ctypes.byref(peer)) != 0:
                    ok = False
                    break
                ptrs[r] = peer.value
        flags = [None] * world
        dist.all_gather_object(flags, bool(ok), group=group)
        if all(flags):
            torch.cuda.synchronize()
            dist.barrier(group=group)
            return cls(rank, world, nbytes, ptr.value, ptrs, "cuda-ipc")
        # some rank could not export/map: release and try torch symmetric memory on every rank
        for r, p in enumerate(ptrs):
            if p and r != rank:
                lib.nvllm_tp_comm_close(ctypes.c_void_p(p))
        if st == 0:
            lib.nvllm_tp_comm_free(ptr)
        return cls._from_symmetric_memory(nbytes, group)

    @classmethod
    def _from_symmetric_memory(cls, nbytes: int, group=None) -> "TPComm":
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
        hdl = symm_mem.rendezvous(buf, group)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        buf.zero_()   # the flag-in-data inboxes must not hold anything that looks like a call epoch
        _lib.check(_lib.load().nvllm_tp_comm_reset(ctypes.c_void_p(ptrs[rank]), None), "tp_comm_reset")
        torch.cuda.synchronize()
        dist.barrier(group=group)
        return cls(rank, world, nbytes, ptrs[rank], ptrs, "symmetric-memory", keepalive=(buf, hdl))

    @classmethod
    def single_process(cls, devices: list[int], nbytes: int) -> list["TPComm"]:
        """One process driving several GPUs (tests, microbenchmarks): direct peer access, no IPC."""
        lib = _lib.load()
        world = len(devices)
        ptrs = []
        for d in devices:
            with torch.cuda.device(d):
                for e in devices:
                    _lib.check(lib.nvllm_tp_comm_enable_peer(e), f"enable peer access {d}->{e}")
                p = ctypes.c_void_p()
                _lib.check(lib.nvllm_tp_comm_alloc(nbytes, ctypes.byref(p), None), "tp_comm_alloc")
                ptrs.append(p.value)
        return [cls(r, world, nbytes, ptrs[r], ptrs, "peer-access") for r in range(world)]

    # -- inspection ---------------------------------------------------------------------------------------------
    def status(self) -> tuple[int, int]:
        """(call epoch, latched error word) — synchronous; error != 0 means a peer never arrived."""
        ep, err = ctypes.c_uint32(), ctypes.c_uint32()
        _lib.check(_lib.load().nvllm_tp_comm_status(ctypes.c_void_p(self.local_ptr), ctypes.byref(ep), ctypes.byref(err)),
                   "tp_comm_status")
        return ep.value, err.value


def poll_error(clear: bool = False) -> int:
    """Timeout latch of the fused kernels launched by this process (0 = none): a mapped host word, no CUDA call."""
    err = ctypes.c_uint32()
    _lib.check(_lib.load().nvllm_rowparallel_poll_error(ctypes.byref(err), 1 if clear else 0), "rowparallel_poll_error")
    return err.value


def raise_on_timeout() -> None:
    """Turn a latched peer timeout into an exception, as NCCL would: called after every eager fused
    forward and after every graph replay by the runner glue.  Costs one ctypes call and one host memory read."""
    err = poll_error()
    if err & 0x80000000:    # the tcgen05 tile's mbarrier-timeout latch (csrc/tcgen05.cuh), not a peer timeout
        poll_error(clear=True)
        raise _lib.NvllmError(f"fused RowParallelLinear tcgen05 tile: an mbarrier wait timed out (code {err:#x}); results "
                              f"of the launches queued since are undefined.  The latch has been cleared")
    if err:
        poll_error(clear=True)
        raise _lib.NvllmError(f"fused RowParallelLinear all-reduce: a tensor-parallel peer never arrived (tile {err - 1}); "
                              f"hidden states of that step are undefined.  NVLLM_FUSED_ALLREDUCE_TIMEOUT_MS sets the "
                              f"wait (default 20000), NVLLM_FUSED_ALLREDUCE=0 restores the NCCL all-reduce")


_COMM: TPComm | None = None
_SHAPES: set[tuple[int, int]] = set()
_ENABLED = False
# NVLLM_FUSED_ALLREDUCE: "1" fused kernel, "0" the reference's F.linear + NCCL all_reduce.  Unset: fused at the world
# sizes where the FINAL kernel (flag-in-data exchange) passed tests/test_gpu_rowparallel*.py on B200s and bench.py's
# fused_allreduce leg reported parity_ok: 2 GPUs (profiles/r02_tp2_tests.log, r02_bench_n2.json) and 4 GPUs
# (profiles/r02_tp4_tests.log, r02_bench_n4.json) and 8 GPUs (profiles/r02_tp8_tests.log: 118 tests, r02_bench_n8.json).
# Other world sizes opt in with "1".
VALIDATED_WORLDS = (2, 4, 8)


def configure(cuda_backend: bool) -> None:
    """Called by ModelRunner.__init__ (INTEGRATION.md edit A7) after the process group exists and before the model
    is built: the fused op is part of the `cuda` backend only, so the reference's other backends keep their stock
    collective.  Also drops the communicator of a previous engine in this process (its peers are gone)."""
    global _ENABLED
    env = os.environ.get("NVLLM_FUSED_ALLREDUCE")
    if env is None:
        import torch.distributed as dist
        world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        want = world in VALIDATED_WORLDS
    else:
        want = env != "0"
    _ENABLED = bool(cuda_backend) and want
    tmo = os.environ.get("NVLLM_FUSED_ALLREDUCE_TIMEOUT_MS")
    if tmo is not None and int(tmo) > 0:
        ops.rowparallel_set_plan(timeout_ms=int(tmo))
    release()


def enabled() -> bool:
    return _ENABLED


def release() -> None:
    global _COMM
    _SHAPES.clear()
    if _COMM is None:
        return
    comm, _COMM = _COMM, None
    if comm.how == "cuda-ipc":
        lib = _lib.load()
        torch.cuda.synchronize()
        for r, p in enumerate(comm.peer_ptrs):
            if r != comm.rank:
                lib.nvllm_tp_comm_close(ctypes.c_void_p(p))
        lib.nvllm_tp_comm_free(ctypes.c_void_p(comm.local_ptr))


def get_comm() -> TPComm:
    """The process-wide communicator, created on first use from the default process group and sized for every
    RowParallelLinear registered so far (all are constructed before the first forward)."""
    global _COMM
    if _COMM is None:
        import torch.distributed as dist
        world = dist.get_world_size()
        _COMM = TPComm.from_process_group(TPComm.required_bytes(world, sorted(_SHAPES)))
    return _COMM


def supported(in_features_per_rank: int, out_features: int, world: int) -> bool:
    return 2 <= world <= 8 and in_features_per_rank % 64 == 0 and out_features % 64 == 0 and \
        int(_lib.load().nvllm_rowparallel_comm_bytes(world, MAX_TOKENS, in_features_per_rank, out_features)) > 0


def install(RowParallelLinear) -> None:
    """Patch the reference class in place (INTEGRATION.md edit A7).  tp_size == 1, unsupported shapes, batches
    above MAX_TOKENS and NVLLM_FUSED_ALLREDUCE=0 run the reference's own forward unchanged."""
    ref_forward = RowParallelLinear.forward
    ref_init = RowParallelLinear.__init__

    def __init__(self, *args, **kwargs):
        ref_init(self, *args, **kwargs)
        self._nvllm_fused = bool(enabled() and self.tp_size > 1 and
                                 supported(self.input_size_per_partition, self.output_size, self.tp_size))
        if self._nvllm_fused:
            _SHAPES.add((self.input_size_per_partition, self.output_size))
        self._nvllm_fused_max = engine_max_tokens(self.input_size_per_partition, self.output_size, self.tp_size)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not getattr(self, "_nvllm_fused", False) or x.dim() != 2 or x.shape[0] > self._nvllm_fused_max or \
                x.dtype != torch.bfloat16 or not x.is_cuda:
            return ref_forward(self, x)
        y = ops.rowparallel_linear_allreduce(x, self.weight, self.bias, get_comm())
        if not torch.cuda.is_current_stream_capturing():
            raise_on_timeout()   # a timeout of an EARLIER launch (this one is still running): one host memory read
        return y

    RowParallelLinear.__init__ = __init__
    RowParallelLinear.forward = forward
