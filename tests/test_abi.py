"""CPU: the C-ABI library loads, exports every symbol include/nvllm_attn.h declares, validates
arguments without touching a GPU, and its host-side index math matches the reference fixtures."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "nvllm_attn.h")
I32P = ctypes.POINTER(ctypes.c_int32)


def _declared():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"NVLLM_API[^;(]*?\b(nvllm_\w+)\s*\(", src)))


def test_header_symbols_exported(lib):
    names = _declared()
    assert len(names) >= 15
    from nano_vllm_learn_b200 import _lib
    assert sorted(_lib.SIGNATURES) == names, "ctypes binding and header disagree"
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/nvllm_attn.h but not exported"


def test_abi_version_and_strings(lib):
    assert lib.nvllm_abi_version() == 2
    assert lib.nvllm_status_string(0) == b"ok"
    for s in range(1, 6):
        assert lib.nvllm_status_string(s) not in (b"ok", b"unknown status")
    assert lib.nvllm_status_string(99) == b"unknown status"


def test_argument_validation_without_gpu(lib):
    # all of these return before any CUDA call
    assert lib.nvllm_store_kvcache(None, 0, None, 0, None, None, None, 0, 8, 128, 2, None) == 0   # n == 0 is a no-op
    assert lib.nvllm_store_kvcache(None, 1024, None, 1024, None, None, None, 4, 8, 128, 2, None) == 1
    assert lib.nvllm_store_kvcache(16, 100, 16, 1024, 16, 16, 16, 4, 8, 128, 2, None) == 1          # stride < row
    assert lib.nvllm_store_kvcache(8, 1024, 16, 1024, 16, 16, 16, 4, 8, 128, 2, None) == 3           # misaligned
    assert lib.nvllm_store_kvcache(16, 1028, 16, 1024, 16, 16, 16, 4, 8, 128, 2, None) == 3          # stride*2 % 16
    # decode: unsupported head_dim / group, missing workspace
    args = dict(out=16, q=16, qs=2048, k=16, v=16, blk=256 * 1024, tok=1024, bt=16, bts=4, cl=16)

    def dec(bs=1, hq=16, hkv=8, d=128, bsz=256, splits=1, ws=None, wsb=0):
        return lib.nvllm_paged_decode(args["out"], args["q"], args["qs"], args["k"], args["v"], args["blk"],
                                      args["tok"], args["bt"], args["bts"], args["cl"], bs, hq, hkv, d, bsz, 0.1,
                                      splits, ws, wsb, None)
    assert dec(bs=0) == 0
    assert dec(d=64) == 2
    assert dec(hq=24, hkv=1) == 2            # G = 24 > 8
    assert dec(hq=10, hkv=4) == 2            # not divisible
    assert dec(bsz=100) == 2
    assert dec(splits=-513) == 1             # balanced grid: at most 512 CTAs
    assert dec(splits=0) == 5                # balanced grid: workspace required
    assert dec(splits=4) == 5                # workspace required
    assert lib.nvllm_paged_decode_workspace_bytes(2, 16, 128, 1) == 0
    assert lib.nvllm_paged_decode_workspace_bytes(2, 16, 128, 4) == 65536 + 2 * 16 * 4 * (128 * 4 + 4)
    # prefill
    def pre(batch=1, mq=8, d=128, hq=4, hkv=2, bt=None, bsz=0):
        return lib.nvllm_varlen_prefill(16, 16, 512, 16, 256, 16, 256, 0, 16, 16, batch, 8, 8, mq, mq, bt, 0, hq, hkv, d,
                                        bsz, 0.1, None)
    assert pre(batch=0) == 0
    assert pre(mq=0) == 0
    assert pre(d=64) == 2
    assert pre(hq=3) == 2
    assert pre(bt=16, bsz=100) == 2
    assert lib.nvllm_paged_decode_set_variant(999) == 1
    assert lib.nvllm_paged_decode_set_variant(-1) == 0
    assert lib.nvllm_paged_decode_num_variants() >= 1


def test_plan_is_sane(lib, monkeypatch):
    monkeypatch.delenv("NVLLM_DECODE_PLAN", raising=False)
    for bs, hkv, ctx in [(1, 8, 512), (16, 8, 4096), (64, 8, 4096), (128, 8, 4096), (128, 1, 4096), (1, 1, 40960),
                         (4, 8, 40960), (5, 1, 32768), (3, 8, 8192)]:
        s = lib.nvllm_paged_decode_plan(bs, 2 * hkv, hkv, ctx, 148)
        assert 1 <= s <= 64
        assert ctx // s >= 256 or s == 1
        assert bs * hkv * s <= 148 or s == 1          # one wave
    # tiny grids must be split, saturated grids must not be split at all
    assert lib.nvllm_paged_decode_plan(1, 4, 1, 40960, 148) == 64
    assert lib.nvllm_paged_decode_plan(1, 32, 8, 32768, 148) == 16          # power of two: equal chunks per warp
    assert lib.nvllm_paged_decode_plan(2, 32, 8, 32768, 148) == 8
    assert lib.nvllm_paged_decode_plan(4, 32, 8, 32768, 148) == 4
    assert lib.nvllm_paged_decode_plan(5, 4, 1, 32768, 148) == 29           # 16 would idle half the SMs
    assert lib.nvllm_paged_decode_plan(512, 16, 8, 4096, 148) == 1
    assert lib.nvllm_paged_decode_plan(128, 2, 1, 4096, 148) == 1            # the tp=8 shard of the 8B shape
    # opt-in: 0 = the balanced persistent grid, wherever its arrival counters cover the batch
    monkeypatch.setenv("NVLLM_DECODE_PLAN", "balanced")
    for bs, hkv, ctx in [(1, 8, 512), (64, 8, 4096), (128, 1, 4096), (1, 1, 40960), (1024, 8, 4096)]:
        assert lib.nvllm_paged_decode_plan(bs, 4 * hkv, hkv, ctx, 148) == 0
    assert lib.nvllm_paged_decode_plan(4096, 16, 8, 4096, 148) >= 1          # beyond the arrival-counter area
    # batch-independent workspace: arrival counters + one head and one tail partial for each of at most 512 CTAs
    ws0 = lib.nvllm_paged_decode_workspace_bytes(64, 16, 128, 0)
    assert ws0 == 65536 + 512 * 2 * 8 * 129 * 4
    assert lib.nvllm_paged_decode_workspace_bytes(1, 4, 128, -10) == ws0


def _glue_fixture():
    with open(os.path.join(os.path.dirname(__file__), "golden", "glue.json")) as f:
        return json.load(f)["traces"]


def test_prepare_decode_slots_matches_reference_fixture(lib):
    for tr in _glue_fixture():
        steps = tr.get("steps") or [dict(decode=tr["decode"], state=[tr["state"]])]
        for st in steps:
            lens = np.array([s["num_tokens"] for s in st["state"]], dtype=np.int32)
            last = np.array([s["block_table"][-1] for s in st["state"]], dtype=np.int32)
            out = np.zeros_like(lens)
            rc = lib.nvllm_prepare_decode_slots(lens.ctypes.data_as(I32P), last.ctypes.data_as(I32P), len(lens), 256,
                                                out.ctypes.data_as(I32P))
            assert rc == 0
            assert out.tolist() == st["decode"]["slot_mapping"]


def test_prepare_prefill_slots_matches_reference_fixture(lib):
    for tr in _glue_fixture():
        if "prefill" not in tr:
            continue
        want = tr["prefill"]["slot_mapping"]
        got = []
        for prompt, state in zip(tr["prompts"], tr["prefill_state"]):
            bt = np.array(state["block_table"], dtype=np.int32)
            buf = np.zeros(len(prompt) + 8, dtype=np.int32)
            n = lib.nvllm_prepare_prefill_slots(bt.ctypes.data_as(I32P), len(bt), len(prompt),
                                                state["num_cached_tokens"], 256, buf.ctypes.data_as(I32P), len(buf))
            assert n >= 0
            got.extend(buf[:n].tolist())
        assert got == want
    # capacity and argument errors are negative statuses, never a crash
    bt = np.array([3], dtype=np.int32)
    buf = np.zeros(4, dtype=np.int32)
    assert lib.nvllm_prepare_prefill_slots(bt.ctypes.data_as(I32P), 1, 10, 0, 256, buf.ctypes.data_as(I32P), 4) == -5
    assert lib.nvllm_prepare_prefill_slots(bt.ctypes.data_as(I32P), 1, 600, 0, 256, buf.ctypes.data_as(I32P), 4) == -1
    assert lib.nvllm_prepare_prefill_slots(None, 0, 600, 0, 256, None, 0) == 0


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from nano_vllm_learn_b200 import _lib
    monkeypatch.setattr(_lib, "_LIB", None)
    monkeypatch.setenv("NVLLM_ATTN_LIB", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.NvllmError):
        _lib.load()


def test_ops_refuse_cpu_tensors(lib):
    import torch
    from nano_vllm_learn_b200 import ops
    k = torch.zeros(2, 2, 128, dtype=torch.bfloat16)
    cache = torch.zeros(1, 256, 2, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        ops.store_kvcache(k, k, cache, cache, torch.zeros(2, dtype=torch.int32))


def test_sharded_sampler_host_arithmetic_without_gpu(lib):
    # exchange buffer: header + (2 parities x world x 16 B pairs + 2 x 8 B temperature words) per row, never below
    # nvllm_tp_comm_alloc's minimum
    assert lib.nvllm_sample_comm_bytes(1, 64) == 0 and lib.nvllm_sample_comm_bytes(9, 64) == 0
    assert lib.nvllm_sample_comm_bytes(2, 0) == 0 and lib.nvllm_sample_comm_bytes(2, 4097) == 0
    assert lib.nvllm_sample_comm_bytes(2, 512) == 299008
    assert lib.nvllm_sample_comm_bytes(8, 4096) == 4096 + (2 * 8 * 16 + 16) * 4096
    ptrs = (ctypes.c_void_p * 2)(16, 32)
    # world < 2, bs == 0 no-op, rank out of range, rank 0 without temperatures, offset not a multiple of 4
    base = dict(out=16, logits=16, eb=2, stride=1024, temps=16, noise=None, ns=0, bs=2, vocab=1024, off=0, ws=16, wsb=1 << 20,
                comm=ptrs, cb=299008, rank=0, world=2)

    def call(**kw):
        a = dict(base, **kw)
        return lib.nvllm_sample_tokens_sharded(a["out"], None, a["logits"], a["eb"], a["stride"], a["temps"], a["noise"], a["ns"],
                                               a["bs"], a["vocab"], a["off"], 0, 0, a["ws"], a["wsb"], a["comm"], a["cb"],
                                               a["rank"], a["world"], None)
    assert call(world=1) == 1
    assert call(bs=0) == 0
    assert call(rank=2) == 1
    assert call(temps=None) == 1
    assert call(off=1022) == 2
    assert call(cb=1024) == 5
    assert call(bs=4096, cb=299008) == 5          # buffer holds (299008 - 4096) / 80 = 3686 rows
    err = ctypes.c_uint32(7)
    assert lib.nvllm_sample_sharded_poll_error(ctypes.byref(err), 0) in (0, 4)   # no device: pinned alloc may be refused
    assert lib.nvllm_sample_sharded_set_timeout_ms(0) == 1
