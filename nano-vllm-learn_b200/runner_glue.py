"""Runner glue for the cuda attention backend: prepare_prefill / prepare_decode / CUDA-graph capture.

Re-implements, bit-identically in every tensor it publishes, the glue of the reference
`nanovllm/engine/model_runner.py`:
    prepare_block_tables :160-169     prepare_prefill :171-242     prepare_decode :244-269
    run_model (graph replay) :278-303   capture_cudagraph :316-370
B200-first differences (host side only, results unchanged):
  * persistent pinned staging + ONE packed host->device copy per step instead of five pinned
    allocations and five copies; index math for slots runs in C (nvllm_prepare_*_slots);
  * in graph mode the metadata is staged directly in the captured buffers' layout (zero-padded
    rows, -1-padded table columns inside the live width, exactly what `zero_()` + row copies
    produce in the reference), so replay needs no extra device ops;
  * a small ring of staging slots guarded by CUDA events keeps worker ranks (which never sync with
    the host) from overwriting a slot whose copy is still in flight;
  * device-side decode state: while the running batch is unchanged from one decode
    step to the next, nothing is staged on the host at all -- one tiny kernel (nvllm_advance_decode_state)
    takes input_ids from the sampler's device output, bumps positions / context_lens, recomputes the slots
    from the table already on the device and patches in the few block ids BlockManager.may_append opened
    for this step (passed as kernel arguments).  Any change of batch composition, layout or an intervening
    prefill falls back to the staged path, which re-publishes everything.
The engine-facing quirks are kept on purpose: decode `positions = len(seq)` (model_runner.py:251),
zero padding of graph rows (:291-293), `block_tables` only when a prefix hit exists (:225-232).
"""
from __future__ import annotations

import ctypes
from typing import Callable, Optional, Sequence

import numpy as np
import torch

from . import _lib
from . import rowparallel_cuda as _rp

_I32P = ctypes.POINTER(ctypes.c_int32)


def _i32ptr(a: np.ndarray):
    return a.ctypes.data_as(_I32P)


class _Slot:
    """One pinned staging slot + its typed numpy view and the event that guards reuse."""

    def __init__(self, nbytes: int, pinned: bool):
        self.host = torch.empty(nbytes, dtype=torch.uint8, device="cpu", pin_memory=pinned)
        self.np = self.host.numpy()
        self.event = torch.cuda.Event() if pinned else None

    def wait(self):
        if self.event is not None:
            self.event.synchronize()

    def mark(self):
        if self.event is not None:
            self.event.record()


def _align(n: int, a: int = 16) -> int:
    return (n + a - 1) // a * a


class RunnerGlue:
    """Host-side metadata staging for one rank.  `set_context` is the engine's context setter."""

    RING = 3

    def __init__(self, block_size: int, max_num_seqs: int, max_model_len: int, max_num_batched_tokens: int,
                 device, set_context: Callable, pad_slot: int = 0):
        self.block_size = block_size
        self.max_bs = max_num_seqs
        self.max_blocks = (max_model_len + block_size - 1) // block_size
        self.max_tokens = max_num_batched_tokens
        self.device = torch.device(device)
        self.pinned = self.device.type == "cuda"
        self.set_context = set_context
        self.pad_slot = pad_slot  # 0 = reference behaviour (model_runner.py:291-293); -1 = skip padded rows
        self.lib = _lib.load()
        # decode: ids i64 | pos i64 | slots i32 | lens i32 | tables i32[bs, W]
        self._dec_bytes = self._decode_layout(self.max_bs, self.max_blocks)[-1]
        # prefill: ids i64[T] | pos i64[T] | slots i32[T] | cu_q i32[B+1] | cu_k i32[B+1] | tables i32[B, W]
        t, b = self.max_tokens, self.max_bs
        self._pre_bytes = _align(8 * t) * 2 + _align(4 * t) + 2 * _align(4 * (b + 1)) + _align(4 * b * self.max_blocks)
        nbytes = max(self._dec_bytes, self._pre_bytes)
        self._ring = [_Slot(nbytes, self.pinned) for _ in range(self.RING)]
        self._turn = 0
        self.dev = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
        self._scratch_len = np.zeros(self.max_bs, dtype=np.int32)
        self._scratch_last = np.zeros(self.max_bs, dtype=np.int32)
        self._scratch_last_ptr = _i32ptr(self._scratch_last)
        self._slot_sig: list = [None] * self.RING   # what each ring slot's decode table currently holds
        self._host_view_cache: dict = {}
        self._dev_view_cache: dict = {}
        self.graph_bs: Optional[list[int]] = None   # set by capture; enables the bucketed decode layout
        # device-side decode state: what the device buffers hold after the last decode step (None = unknown)
        import os
        self.device_state = self.pinned and os.environ.get("NVLLM_DEVICE_DECODE_STATE", "1") != "0"
        self._fast = None
        self._max_patches = int(self.lib.nvllm_advance_decode_state_max_patches())
        self._patch = np.zeros((3, self._max_patches), dtype=np.int32)
        self._patch_ptrs = tuple(_i32ptr(self._patch[i]) for i in range(3))
        self._adv_err = ctypes.c_uint32(0)
        self.fast_steps = 0      # decode steps served by the device-side path (tests, profiling)
        self.staged_steps = 0

    # ------------------------------------------------------------------ layouts
    @staticmethod
    def _decode_layout(gbs: int, width: int):
        o_ids = 0
        o_pos = o_ids + _align(8 * gbs)
        o_slot = o_pos + _align(8 * gbs)
        o_len = o_slot + _align(4 * gbs)
        o_tab = o_len + _align(4 * gbs)
        end = o_tab + _align(4 * gbs * width)
        return o_ids, o_pos, o_slot, o_len, o_tab, end

    def _typed(self, buf, off: int, count: int, dtype):
        """Typed view [count] at byte offset `off` of a uint8 numpy array or torch tensor."""
        if isinstance(buf, np.ndarray):
            return buf[off:off + count * np.dtype(dtype).itemsize].view(dtype)
        tdt = {np.int64: torch.int64, np.int32: torch.int32}[dtype]
        return buf[off:off + count * np.dtype(dtype).itemsize].view(tdt)

    def decode_views(self, buf, gbs: int, width: int):
        o_ids, o_pos, o_slot, o_len, o_tab, _ = self._decode_layout(gbs, width)
        ids = self._typed(buf, o_ids, gbs, np.int64)
        pos = self._typed(buf, o_pos, gbs, np.int64)
        slots = self._typed(buf, o_slot, gbs, np.int32)
        lens = self._typed(buf, o_len, gbs, np.int32)
        tab = self._typed(buf, o_tab, gbs * width, np.int32)
        tab = tab.reshape(gbs, width) if isinstance(tab, np.ndarray) else tab.view(gbs, width)
        return ids, pos, slots, lens, tab

    def _next_slot(self) -> _Slot:
        slot = self._ring[self._turn]
        self._turn = (self._turn + 1) % self.RING
        slot.wait()
        return slot

    def _upload(self, slot: _Slot, nbytes: int):
        self.dev[:nbytes].copy_(slot.host[:nbytes], non_blocking=True)
        slot.mark()

    # ------------------------------------------------------------------ reference: model_runner.py:160-169
    def prepare_block_tables(self, seqs: Sequence) -> torch.Tensor:
        width = max(len(s.block_table) for s in seqs)
        n = len(seqs)
        self._slot_sig = [None] * self.RING
        self._fast = None
        slot = self._next_slot()
        tab = slot.np[:4 * n * width].view(np.int32).reshape(n, width)
        tab.fill(-1)
        for i, s in enumerate(seqs):
            bt = s.block_table
            tab[i, :len(bt)] = bt
        # standalone use only (prepare_prefill/prepare_decode stage their tables inline): fresh device tensor
        out = torch.empty((n, width), dtype=torch.int32, device=self.device)
        out.copy_(slot.host[:4 * n * width].view(torch.int32).view(n, width), non_blocking=True)
        slot.mark()
        return out

    # ------------------------------------------------------------------ reference: model_runner.py:171-242
    def prepare_prefill(self, seqs: Sequence):
        bsz = self.block_size
        n = len(seqs)
        total_q = sum(len(s) - s.num_cached_tokens for s in seqs)
        width = max(len(s.block_table) for s in seqs) if n else 0
        o_ids = 0
        o_pos = o_ids + _align(8 * total_q)
        o_slot = o_pos + _align(8 * total_q)
        # slots of a sequence cover its uncached BLOCKS (>= its uncached tokens only if cached is block aligned,
        # which BlockManager guarantees), so total_q bounds the count
        o_cuq = o_slot + _align(4 * total_q)
        o_cuk = o_cuq + _align(4 * (n + 1))
        o_tab = o_cuk + _align(4 * (n + 1))
        end = o_tab + _align(4 * n * width)
        if end > self.dev.numel():
            raise ValueError("prefill batch exceeds max_num_batched_tokens/max_num_seqs staging capacity")
        self._slot_sig = [None] * self.RING   # new sequences may reuse freed blocks: no decode table survives a prefill
        self._fast = None                     # and the prefill staging overwrites the device buffer
        slot = self._next_slot()
        h = slot.np
        ids = h[o_ids:o_ids + 8 * total_q].view(np.int64)
        pos = h[o_pos:o_pos + 8 * total_q].view(np.int64)
        slots = h[o_slot:o_slot + 4 * total_q].view(np.int32)
        cu_q = h[o_cuq:o_cuq + 4 * (n + 1)].view(np.int32)
        cu_k = h[o_cuk:o_cuk + 4 * (n + 1)].view(np.int32)
        cu_q[0] = cu_k[0] = 0
        max_q = max_k = 0
        tq = 0
        n_slots = 0
        for i, s in enumerate(seqs):
            seqlen = len(s)
            cached = s.num_cached_tokens
            lq = seqlen - cached
            ids[tq:tq + lq] = s[cached:]
            pos[tq:tq + lq] = np.arange(cached, seqlen, dtype=np.int64)
            tq += lq
            cu_q[i + 1] = cu_q[i] + lq
            cu_k[i + 1] = cu_k[i] + seqlen
            max_q = max(max_q, lq)
            max_k = max(max_k, seqlen)
            bt = s.block_table
            if not bt:
                continue  # warmup sequences have no blocks (model_runner.py:209-210)
            bt_np = np.asarray(bt, dtype=np.int32)
            wrote = self.lib.nvllm_prepare_prefill_slots(_i32ptr(bt_np), len(bt), seqlen, cached, bsz,
                                                         _i32ptr(slots[n_slots:]) if total_q > n_slots else None,
                                                         total_q - n_slots)
            if wrote < 0:
                raise AssertionError(f"prepare_prefill slot math failed with status {-wrote}")
            n_slots += int(wrote)
        has_prefix = int(cu_k[n]) > int(cu_q[n])
        if has_prefix:
            tab = h[o_tab:o_tab + 4 * n * width].view(np.int32).reshape(n, width)
            tab.fill(-1)
            for i, s in enumerate(seqs):
                bt = s.block_table
                tab[i, :len(bt)] = bt
        self._upload(slot, end if has_prefix else o_tab)
        d = self.dev
        d_ids = d[o_ids:o_ids + 8 * total_q].view(torch.int64)
        d_pos = d[o_pos:o_pos + 8 * total_q].view(torch.int64)
        d_slots = d[o_slot:o_slot + 4 * n_slots].view(torch.int32)
        d_cuq = d[o_cuq:o_cuq + 4 * (n + 1)].view(torch.int32)
        d_cuk = d[o_cuk:o_cuk + 4 * (n + 1)].view(torch.int32)
        d_tab = d[o_tab:o_tab + 4 * n * width].view(torch.int32).view(n, width) if has_prefix else None
        self.set_context(True, d_cuq, d_cuk, max_q, max_k, d_slots, None, d_tab)
        return d_ids, d_pos

    # ------------------------------------------------------------------ reference: model_runner.py:244-269
    def decode_bucket(self, bs: int) -> Optional[int]:
        """Captured graph batch size that serves `bs` (model_runner.py:288), None -> eager."""
        if not self.graph_bs or bs > self.graph_bs[-1]:
            return None
        return next(x for x in self.graph_bs if x >= bs)

    def _host_decode_views(self, slot_idx: int, rows: int, cols: int):
        """Typed numpy views (and ctypes pointers) of ring slot `slot_idx` for a [rows, cols] decode layout."""
        key = (slot_idx, rows, cols)
        v = self._host_view_cache.get(key)
        if v is None:
            ids, pos, slots, lens, tab = self.decode_views(self._ring[slot_idx].np, rows, cols)
            v = (ids, pos, slots, lens, tab, _i32ptr(lens), _i32ptr(slots))
            if len(self._host_view_cache) > 512:
                self._host_view_cache.clear()
            self._host_view_cache[key] = v
        return v

    def _dev_decode_views(self, rows: int, cols: int, bs: int, width: int):
        """Device views published through the Context for this step's geometry (cached: slicing costs host time)."""
        key = (rows, cols, bs, width)
        v = self._dev_view_cache.get(key)
        if v is None:
            d_ids, d_pos, d_slots, d_lens, d_tab = self.decode_views(self.dev, rows, cols)
            nbytes = self._decode_layout(rows, cols)[-1]
            v = (d_ids[:bs], d_pos[:bs], d_slots[:bs], d_lens[:bs], d_tab[:bs, :width], self.dev[:nbytes],
                 [slot.host[:nbytes] for slot in self._ring])
            if len(self._dev_view_cache) > 512:
                self._dev_view_cache.clear()
            self._dev_view_cache[key] = v
        return v

    def _advance_on_device(self, seqs: Sequence, ids_key: tuple, device_tokens: torch.Tensor):
        """One decode step of an unchanged batch without host staging; None -> the caller takes the staged path."""
        st = self._fast
        if st is None or st["ids"] != ids_key or device_tokens is None:
            return None
        # a row the kernel had to park at an EARLIER step (no block behind its new token): one host memory read
        self.lib.nvllm_advance_decode_state_poll_error(ctypes.byref(self._adv_err), 0)
        if self._adv_err.value:
            row = self._adv_err.value - 1
            self.lib.nvllm_advance_decode_state_poll_error(ctypes.byref(self._adv_err), 1)
            self._fast = None
            raise _lib.NvllmError(f"device-side prepare_decode: row {row} had no block for its new token (a block-table "
                                  f"patch was missing); that row's output of the affected step is zeros")
        bs, bsz = st["bs"], self.block_size
        # every row must have grown by exactly one token, else the device state is stale: restage.  All rows are
        # compared (bs <= 512: a few microseconds), and before the host copy of the lengths is touched.
        now = np.fromiter(map(len, seqs), dtype=np.int64, count=bs) if len(seqs) == bs else None
        if now is None or not np.array_equal(now, st["lens"] + 1):
            self._fast = None
            return None
        lens = st["lens"] = now
        opened = np.flatnonzero((lens - 1) % bsz == 0)                 # rows whose new token starts a block
        n_patch = int(opened.size)
        width = old_width = st["width"]
        if n_patch:
            if n_patch > self._max_patches:
                self._fast = None
                return None
            cols = (lens[opened] - 1) // bsz
            new_width = max(width, int(cols.max()) + 1)
            if new_width != width and st["gbs"] is None:               # eager layout is [bs, width]: it changes
                self._fast = None
                return None
            width = st["width"] = new_width
            self._patch[0, :n_patch] = opened
            self._patch[1, :n_patch] = cols
            self._patch[2, :n_patch] = [seqs[i].block_table[-1] for i in opened]
        rows, cols_l = st["rows"], st["cols"]
        d_ids, d_pos, d_slots, d_lens, d_tab, _, _ = self._dev_decode_views(rows, cols_l, bs, width)
        ptrs = st.get("ptrs")
        if ptrs is None:
            f_ids, f_pos, f_slots, f_lens, f_tab = self.decode_views(self.dev, rows, cols_l)
            ptrs = st["ptrs"] = (f_ids.data_ptr(), f_pos.data_ptr(), f_slots.data_ptr(), f_lens.data_ptr(),
                                 f_tab.data_ptr(), cols_l)
        rc = self.lib.nvllm_advance_decode_state(ptrs[0], ptrs[1], ptrs[2], ptrs[3], ptrs[4], ptrs[5],
                                                 device_tokens.data_ptr(), bs, bsz, *self._patch_ptrs, n_patch,
                                                 old_width, width,
                                                 torch.cuda.current_stream().cuda_stream if self.pinned else None)
        if rc != 0:
            self._fast = None
            _lib.check(rc, "advance_decode_state")
        self.fast_steps += 1
        self.set_context(False, slot_mapping=d_slots, context_lens=d_lens, block_tables=d_tab)
        return d_ids, d_pos

    def prepare_decode(self, seqs: Sequence, device_tokens: Optional[torch.Tensor] = None):
        """`device_tokens`: the sampler's int64 output of the PREVIOUS step, still on the device, row i = seqs[i]
        (sampler_cuda.Sampler.last_tokens); enables the device-side path for an unchanged batch."""
        bsz = self.block_size
        bs = len(seqs)
        ids_key = None
        if self.device_state:
            ids_key = tuple(map(seq_identity, seqs))
            if ids_key is not None and device_tokens is not None and self._fast is not None:
                out = self._advance_on_device(seqs, ids_key, device_tokens)
                if out is not None:
                    return out
        self.staged_steps += 1
        tables = [s.block_table for s in seqs]
        width = max(map(len, tables))
        gbs = self.decode_bucket(bs)
        if gbs is None:          # eager: tight [bs, width] table, -1 padded (prepare_block_tables)
            rows, cols = bs, width
        else:                    # graph: captured [gbs, max_blocks] layout, zero padded outside [:bs, :width]
            rows, cols = gbs, self.max_blocks
        slot_idx = self._turn
        slot = self._next_slot()
        ids, pos, slots, lens, tab, lens_ptr, slots_ptr = self._host_decode_views(slot_idx, rows, cols)
        last = self._scratch_last
        n_tok = [len(s) for s in seqs]
        last_blocks = [bt[-1] for bt in tables]
        # The table of a running batch changes only when a block is appended or the batch changes; between two
        # prefills a (row count, layout, per-row length, per-row last block) signature pins the content of this
        # ring slot, so the common step rewrites nothing but ids / positions / lens / slots.
        # Sequence identity is part of it: a batch where one running sequence was swapped for another of equal length
        # and equal last block must not inherit the staged table (tests/test_runner_glue.py::test_decode_swap...).
        sig = (rows, cols, bs, width, ids_key if ids_key is not None else tuple(map(seq_identity, seqs)),
               tuple(map(len, tables)), tuple(last_blocks))
        if self._slot_sig[slot_idx] != sig:
            self._slot_sig[slot_idx] = sig
            if gbs is not None:
                tab.fill(0)
                ids[bs:] = 0
                pos[bs:] = 0
                lens[bs:] = 0
                slots[bs:] = self.pad_slot
            else:
                tab.fill(-1)
            if min(map(len, tables)) == width:      # rectangular (the common case): one C-level conversion
                tab[:bs, :width] = tables
            else:
                tab[:bs, :width] = -1
                for i, bt in enumerate(tables):
                    tab[i, :len(bt)] = bt
        ids[:bs] = [s.last_token for s in seqs]
        pos[:bs] = n_tok        # sic: the reference feeds len(seq), not len(seq)-1 (model_runner.py:251)
        lens[:bs] = n_tok
        last[:bs] = last_blocks
        st = self.lib.nvllm_prepare_decode_slots(lens_ptr, self._scratch_last_ptr, bs, bsz, slots_ptr)
        if st != 0:
            raise AssertionError("prepare_decode slot math failed")
        d_ids, d_pos, d_slots, d_lens, d_tab, d_all, h_all = self._dev_decode_views(rows, cols, bs, width)
        d_all.copy_(h_all[slot_idx], non_blocking=True)
        slot.mark()
        if ids_key is not None:
            self._fast = dict(ids=ids_key, rows=rows, cols=cols, bs=bs, width=width, gbs=gbs,
                              lens=np.asarray(n_tok, dtype=np.int64))
        self.set_context(False, slot_mapping=d_slots, context_lens=d_lens, block_tables=d_tab)
        return d_ids, d_pos

    # ------------------------------------------------------------------ reference: model_runner.py:316-370
    def graph_buffers(self, gbs: int):
        """Device views a graph of batch size `gbs` is captured on (and later replayed from)."""
        return self.decode_views(self.dev, gbs, self.max_blocks)


def seq_identity(seq):
    """A key that names the same running sequence at every step and on every tensor-parallel rank.

    `Sequence.seq_id` where the object has one.  The reference does not pickle it (engine/sequence.py:77-90), so the
    copies a worker rank unpickles each step have none unless `install_sequence_identity` patched the class; for those
    the key is (prompt length, first block id) -- stable over the life of a sequence, and distinct between running
    sequences except for two that share their first prefix-cache block AND have equal prompt lengths, which the
    per-row length / last-block terms of the staging signature still tell apart.  Never `id(seq)`: workers get
    fresh objects every step, so an address is neither stable nor unique over time.
    """
    sid = getattr(seq, "seq_id", None)
    if sid is not None:
        return sid
    bt = seq.block_table
    return ("anon", getattr(seq, "num_prompt_tokens", -1), bt[0] if bt else -1)


def install_sequence_identity(sequence_cls) -> None:
    """Make `seq_id` survive the rank-0 -> worker pickle (reference: engine/sequence.py:77-90 drops it, so on worker
    ranks every decode step took the staged path and the step was gated by the slowest rank).  The id travels as one
    tagged trailing element that the patched `__setstate__` strips before the reference's own code sees the state,
    so the reference's five fields keep their positions; both ends of the pipe import the same patched class
    (model_runner.py is imported by every rank process)."""
    if getattr(sequence_cls, "_nvllm_identity", False):
        return
    get0, set0 = sequence_cls.__getstate__, sequence_cls.__setstate__

    def __getstate__(self):
        return (*get0(self), ("nvllm_seq_id", getattr(self, "seq_id", None)))

    def __setstate__(self, state):
        tail = state[-1] if len(state) else None
        if isinstance(tail, tuple) and len(tail) == 2 and tail[0] == "nvllm_seq_id":
            state = state[:-1]
            if tail[1] is not None:
                self.seq_id = tail[1]
        set0(self, state)

    sequence_cls.__getstate__ = __getstate__
    sequence_cls.__setstate__ = __setstate__
    sequence_cls._nvllm_identity = True


# ---------------------------------------------------------------------- engine integration
def install(model_runner_cls, set_context: Callable, get_context: Callable, reset_context: Callable,
            sequence_cls=None) -> None:
    """Swap the glue methods of the host engine's ModelRunner for the cuda backend.

    Called from the reference-side binding: the original methods stay in place for
    every other backend; dispatch is on `hf_config.attn_backend == "cuda"` per instance.
    `sequence_cls` (the engine's Sequence): its pickled state additionally carries `seq_id`, which the device-side
    decode path needs on worker ranks (install_sequence_identity).
    """
    if sequence_cls is not None:
        install_sequence_identity(sequence_cls)
    orig = {name: getattr(model_runner_cls, name) for name in
            ("prepare_block_tables", "prepare_prefill", "prepare_decode", "run_model", "capture_cudagraph")}

    def _is_cuda(self) -> bool:
        return getattr(self.config.hf_config, "attn_backend", "flash") == "cuda"

    def _glue(self) -> RunnerGlue:
        g = getattr(self, "_nvllm_glue", None)
        if g is None:
            cfg = self.config
            import os
            g = RunnerGlue(self.block_size, cfg.max_num_seqs, cfg.max_model_len, cfg.max_num_batched_tokens,
                           torch.device("cuda", torch.cuda.current_device()), set_context,
                           pad_slot=int(os.environ.get("NVLLM_PAD_SLOT", "0")))
            self._nvllm_glue = g
        return g

    def prepare_block_tables(self, seqs):
        return _glue(self).prepare_block_tables(seqs) if _is_cuda(self) else orig["prepare_block_tables"](self, seqs)

    def prepare_prefill(self, seqs):
        return _glue(self).prepare_prefill(seqs) if _is_cuda(self) else orig["prepare_prefill"](self, seqs)

    def prepare_decode(self, seqs):
        if not _is_cuda(self):
            return orig["prepare_decode"](self, seqs)
        # the fused sampler keeps its output on the device (sampler_cuda.Sampler.last_tokens): an unchanged batch is
        # advanced there instead of being re-staged from the host
        return _glue(self).prepare_decode(seqs, getattr(self.sampler, "last_tokens", None))

    @torch.inference_mode()
    def run_model(self, input_ids, positions, is_prefill):
        if not _is_cuda(self):
            return orig["run_model"](self, input_ids, positions, is_prefill)
        bs = input_ids.size(0)
        g = _glue(self)
        gbs = None if (is_prefill or self.enforce_eager) else g.decode_bucket(bs)
        if gbs is None:
            return self.model.compute_logits(self.model(input_ids, positions))
        # prepare_decode already staged ids/positions/context in the captured buffers of bucket `gbs`
        self.graphs[gbs].replay()
        if _rp.enabled():
            _rp.raise_on_timeout()   # fused all-reduce peer timeout of an earlier replay -> exception (one host read)
        return self.model.compute_logits(self.graph_vars["outputs"][:bs])

    @torch.inference_mode()
    def capture_cudagraph(self):
        if not _is_cuda(self):
            return orig["capture_cudagraph"](self)
        config = self.config
        hf_config = config.hf_config
        g = _glue(self)
        max_bs = min(config.max_num_seqs, 512)
        outputs = torch.zeros(max_bs, hf_config.hidden_size)
        self.graph_bs = [1, 2, 4, 8] + list(range(16, max_bs + 1, 16))
        self.graphs = {}
        self.graph_pool = None
        g.dev.zero_()
        g._fast = None          # the staging buffer was just cleared: no device-side decode state survives
        for bs in reversed(self.graph_bs):
            ids, pos, slots, lens, tab = g.graph_buffers(bs)
            graph = torch.cuda.CUDAGraph()
            set_context(False, slot_mapping=slots, context_lens=lens, block_tables=tab)
            outputs[:bs] = self.model(ids, pos)            # warm-up outside capture
            with torch.cuda.graph(graph, self.graph_pool):
                outputs[:bs] = self.model(ids, pos)
            if self.graph_pool is None:
                self.graph_pool = graph.pool()
            self.graphs[bs] = graph
            torch.cuda.synchronize()
            reset_context()
        g.graph_bs = list(self.graph_bs)
        self.graph_vars = dict(outputs=outputs, staging=g.dev)

    orig_exit = getattr(model_runner_cls, "exit", None)

    def exit(self):
        # NVLLM_GLUE_STATS_DIR: every rank leaves its decode-path counters behind (tests assert that worker ranks take
        # the device-side path too; spawned workers skip atexit handlers, so this hangs on the engine's own exit call)
        import json
        import os
        d = os.environ.get("NVLLM_GLUE_STATS_DIR")
        g = getattr(self, "_nvllm_glue", None)
        if d and g is not None:
            with open(os.path.join(d, f"glue_rank{getattr(self, 'rank', 0)}.json"), "w") as f:
                json.dump({"rank": getattr(self, "rank", 0), "fast_steps": g.fast_steps, "staged_steps": g.staged_steps}, f)
        return orig_exit(self)

    if orig_exit is not None:
        model_runner_cls.exit = exit
    for name, fn in (("prepare_block_tables", prepare_block_tables), ("prepare_prefill", prepare_prefill),
                     ("prepare_decode", prepare_decode), ("run_model", run_model),
                     ("capture_cudagraph", capture_cudagraph)):
        setattr(model_runner_cls, name, fn)
