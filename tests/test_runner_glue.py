"""CPU: the rewritten runner glue publishes tensors bit-identical to the reference's
prepare_prefill / prepare_decode (fixtures frozen from the reference + live reference when present)."""
import json
import os
from types import SimpleNamespace

import pytest
import torch

from oracle import runner_glue_oracle as gorc

GOLD = os.path.join(os.path.dirname(__file__), "golden", "glue.json")
BS = 256


class Seq:
    """Duck-typed stand-in for nanovllm.engine.sequence.Sequence (sequence.py:14-69) built from fixture state."""

    def __init__(self, token_ids, block_table, num_cached_tokens=0, num_tokens=None, last_token=None):
        self.token_ids = list(token_ids)
        self.block_table = list(block_table)
        self.num_cached_tokens = num_cached_tokens
        self.num_tokens = num_tokens if num_tokens is not None else len(self.token_ids)
        self.last_token = last_token if last_token is not None else self.token_ids[-1]

    def __len__(self):
        return self.num_tokens

    def __getitem__(self, key):
        return self.token_ids[key]


class Ctx:
    def __init__(self):
        self.args = None

    def set(self, is_prefill, cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=0, max_seqlen_k=0,
            slot_mapping=None, context_lens=None, block_tables=None):
        self.args = SimpleNamespace(is_prefill=is_prefill, cu_seqlens_q=cu_seqlens_q, cu_seqlens_k=cu_seqlens_k,
                                    max_seqlen_q=max_seqlen_q, max_seqlen_k=max_seqlen_k, slot_mapping=slot_mapping,
                                    context_lens=context_lens, block_tables=block_tables)


def _glue(lib, ctx, **kw):
    from nano_vllm_learn_b200.runner_glue import RunnerGlue
    args = dict(block_size=BS, max_num_seqs=64, max_model_len=4096, max_num_batched_tokens=16384, device="cpu",
                set_context=ctx.set)
    args.update(kw)
    return RunnerGlue(**args)


def _traces():
    with open(GOLD) as f:
        return json.load(f)["traces"]


def _eq(t: torch.Tensor, want, dtype: str):
    assert str(t.dtype) == dtype
    assert t.tolist() == want


def test_prefill_matches_reference_fixture(lib):
    for tr in _traces():
        if "prefill" not in tr:
            continue
        ctx = Ctx()
        g = _glue(lib, ctx)
        seqs = [Seq(p, st["block_table"], st["num_cached_tokens"]) for p, st in zip(tr["prompts"], tr["prefill_state"])]
        ids, pos = g.prepare_prefill(seqs)
        want, dt = tr["prefill"], tr["prefill_dtypes"]
        _eq(ids, want["input_ids"], dt["input_ids"])
        _eq(pos, want["positions"], dt["positions"])
        a = ctx.args
        assert a.is_prefill is True and a.context_lens is None
        _eq(a.cu_seqlens_q, want["cu_seqlens_q"], dt["cu_seqlens_q"])
        _eq(a.cu_seqlens_k, want["cu_seqlens_k"], dt["cu_seqlens_k"])
        _eq(a.slot_mapping, want["slot_mapping"], dt["slot_mapping"])
        assert (a.max_seqlen_q, a.max_seqlen_k) == (want["max_seqlen_q"], want["max_seqlen_k"])
        if want["block_tables"] is None:
            assert a.block_tables is None
        else:
            _eq(a.block_tables, want["block_tables"], dt["block_tables"])


@pytest.mark.parametrize("graphs", [False, True])
def test_decode_matches_reference_fixture(lib, graphs):
    for tr in _traces():
        steps = tr.get("steps") or [dict(decode=tr["decode"], dtypes=tr["dtypes"], state=[tr["state"]])]
        ctx = Ctx()
        g = _glue(lib, ctx)
        if graphs:
            g.graph_bs = [1, 2, 4, 8] + list(range(16, 65, 16))
        for st in steps:
            seqs = [Seq([s["last_token"]], s["block_table"], 0, s["num_tokens"], s["last_token"]) for s in st["state"]]
            ids, pos = g.prepare_decode(seqs)
            want, dt = st["decode"], st["dtypes"]
            _eq(ids, want["input_ids"], dt["input_ids"])
            _eq(pos, want["positions"], dt["positions"])
            a = ctx.args
            assert a.is_prefill is False
            _eq(a.slot_mapping, want["slot_mapping"], dt["slot_mapping"])
            _eq(a.context_lens, want["context_lens"], dt["context_lens"])
            _eq(a.block_tables, want["block_tables"], dt["block_tables"])
            if graphs:
                # the captured buffers hold exactly what the reference's zero_() + row copies leave
                # (model_runner.py:291-299): zeros outside [:bs, :width]
                bs, width = len(seqs), len(want["block_tables"][0])
                gbs = g.decode_bucket(bs)
                gi, gp, gs, gl, gt = g.graph_buffers(gbs)
                assert gt.shape == (gbs, g.max_blocks)
                ref_tab = torch.zeros(gbs, g.max_blocks, dtype=torch.int32)
                ref_tab[:bs, :width] = torch.tensor(want["block_tables"], dtype=torch.int32)
                assert torch.equal(gt, ref_tab)
                for buf, key in ((gi, "input_ids"), (gp, "positions"), (gs, "slot_mapping"), (gl, "context_lens")):
                    ref = torch.zeros(gbs, dtype=buf.dtype)
                    ref[:bs] = torch.tensor(want[key], dtype=buf.dtype)
                    assert torch.equal(buf, ref)


def test_pad_slot_flag(lib):
    ctx = Ctx()
    g = _glue(lib, ctx, pad_slot=-1)
    g.graph_bs = [1, 2, 4, 8]
    seqs = [Seq([5], [3], 0, 10, 5) for _ in range(3)]
    g.prepare_decode(seqs)
    _, _, gs, _, _ = g.graph_buffers(4)
    assert gs.tolist() == [3 * 256 + 9] * 3 + [-1]


def test_ring_reuse_does_not_alias_previous_step(lib):
    ctx = Ctx()
    g = _glue(lib, ctx)
    a = [Seq([1], [2], 0, 5, 1)]
    b = [Seq([9], [4, 7], 0, 300, 9)]
    g.prepare_decode(a)
    first = ctx.args.slot_mapping.clone()
    g.prepare_decode(b)
    assert first.tolist() == [2 * 256 + 4]
    assert ctx.args.slot_mapping.tolist() == [7 * 256 + 43]


@pytest.mark.parametrize("graphs", [False, True])
def test_decode_swap_of_equal_shaped_sequence_restages_table(lib, graphs):
    """two batches whose per-row lengths and last blocks agree but whose sequences (and so the
    earlier table entries) differ must not share a staged block table.  Every ring slot is visited."""
    ctx = Ctx()
    g = _glue(lib, ctx)
    if graphs:
        g.graph_bs = [1, 2, 4, 8]

    def seq(seq_id, table):
        s = Seq([7], table, 0, 300, 7)
        s.seq_id = seq_id
        return s

    a = [seq(0, [2, 9]), seq(1, [3, 8])]
    b = [seq(0, [2, 9]), seq(5, [6, 8])]          # row 1 swapped: same length, same LAST block, different first block
    for step in range(2 * g.RING + 1):
        batch = a if step % 2 == 0 else b
        g.prepare_decode(batch)
        want = gorc.prepare_decode_ref(batch)
        assert torch.equal(ctx.args.block_tables, want["block_tables"]), step
        assert torch.equal(ctx.args.slot_mapping, want["slot_mapping"])
        assert torch.equal(ctx.args.context_lens, want["context_lens"])


def test_live_reference_trace(lib):
    """Drive the REAL BlockManager/Sequence/ModelRunner.prepare_* (unmodified) side by side with the glue."""
    root = gorc.find_reference_root()
    if root is None:
        pytest.skip("reference not present")
    try:
        ref = gorc.ReferenceGlue(root)
    except Exception as e:
        pytest.skip(f"reference not importable: {e}")
    import random
    random.seed(7)
    bm = ref.block_manager.BlockManager(128, BS)
    shared = [random.randint(0, 10000) for _ in range(768)]
    prompts = [shared[:512] + [random.randint(0, 10000) for _ in range(n)] for n in (1, 255, 256, 600)]
    prompts.append(shared + [3, 4, 5])
    seqs = [ref.sequence.Sequence(p) for p in prompts]
    for s in seqs:
        bm.allocate(s)
    ctx = Ctx()
    g = _glue(lib, ctx)
    want = ref.prepare_prefill(seqs)
    ids, pos = g.prepare_prefill(seqs)
    assert torch.equal(ids, want["input_ids"]) and torch.equal(pos, want["positions"])
    for k in ("cu_seqlens_q", "cu_seqlens_k", "slot_mapping", "block_tables"):
        got = getattr(ctx.args, k)
        assert got.dtype == want[k].dtype and torch.equal(got, want[k]), k
    for step in range(1, 520):
        for i, s in enumerate(seqs):
            s.append_token((step * 13 + i) % 10000)
            bm.may_append(s)
        if step in (1, 2, 255, 256, 257, 258, 511, 512, 513, 519):
            want = ref.prepare_decode(seqs)
            ids, pos = g.prepare_decode(seqs)
            assert torch.equal(ids, want["input_ids"]) and torch.equal(pos, want["positions"])
            for k in ("slot_mapping", "context_lens", "block_tables"):
                got = getattr(ctx.args, k)
                assert got.dtype == want[k].dtype and torch.equal(got, want[k]), (step, k)
            # and the plain-Python restatement agrees with both
            m = gorc.prepare_decode_ref(seqs)
            for k in ("slot_mapping", "context_lens", "block_tables", "positions", "input_ids"):
                assert torch.equal(m[k], want[k])


def test_device_side_advance_restatement_tracks_prepare_decode():
    """CPU statement of the algorithm behind nvllm_advance_decode_state (the kernel itself is checked on the box,
    tests/test_gpu_decode_state.py): starting from one staged step, 600 increments with the may_append patches reproduce
    what prepare_decode builds from the sequences at every step -- including rows crossing block boundaries at the same
    step and a new column joining the table width."""
    free = list(range(900, 100, -1))

    def mk(n):
        s = Seq([1], [free.pop() for _ in range((n + BS - 1) // BS)], 0, n, 1)
        return s
    seqs = [mk(n) for n in (250, 3, 511, 250, 700)]
    state = gorc.prepare_decode_ref(seqs)
    for step in range(1, 600):
        toks, patches = [], []
        width = state["block_tables"].shape[1]
        for i, s in enumerate(seqs):
            t = (step * 131 + i * 7) % 10000
            s.last_token = t
            s.num_tokens += 1
            if s.num_tokens % BS == 1:
                s.block_table.append(free.pop())
                patches.append((i, len(s.block_table) - 1, s.block_table[-1]))
            toks.append(t)
        new_width = max(len(s.block_table) for s in seqs)
        state = gorc.advance_decode_state_ref(state, toks, patches, (width, new_width), BS)
        want = gorc.prepare_decode_ref(seqs)
        for k in ("input_ids", "positions", "slot_mapping", "context_lens", "block_tables"):
            assert state[k].dtype == want[k].dtype and torch.equal(state[k], want[k]), (step, k)


class _KernelShim:
    """Stands in for the C library in the host-logic test below: forwards everything to the real library except
    nvllm_advance_decode_state, which it executes on the (CPU) staging buffer exactly as csrc/decode_state.cu is
    specified in include/nvllm_attn.h (4b) -- so the HOST side of the device path (what it decides to patch, pad and
    fall back on) runs in the CPU suite; the kernel itself is tested on the box (tests/test_gpu_decode_state.py)."""

    def __init__(self, real):
        self._real = real
        self.calls = 0
        self.max_patches_seen = 0

    def __getattr__(self, name):
        return getattr(self._real, name)

    def nvllm_advance_decode_state(self, ids, pos, slots, lens, tab, stride, sampled, bs, bsz, prow, pcol, pval, n_patch,
                                   pad_lo, pad_hi, stream):
        import ctypes
        import numpy as np

        def arr(ptr, n, ct, dt):
            return np.frombuffer((ct * n).from_address(ptr), dtype=dt)
        a_ids, a_pos = arr(ids, bs, ctypes.c_int64, np.int64), arr(pos, bs, ctypes.c_int64, np.int64)
        a_slots, a_lens = arr(slots, bs, ctypes.c_int32, np.int32), arr(lens, bs, ctypes.c_int32, np.int32)
        a_tab = arr(tab, bs * stride, ctypes.c_int32, np.int32).reshape(bs, stride)
        a_sampled = arr(sampled, bs, ctypes.c_int64, np.int64)
        assert 0 <= n_patch <= self._real.nvllm_advance_decode_state_max_patches() and stream is None
        self.calls += 1
        self.max_patches_seen = max(self.max_patches_seen, n_patch)
        a_tab[:, pad_lo:pad_hi] = -1
        for i in range(n_patch):
            a_tab[prow[i], pcol[i]] = pval[i]
        a_lens += 1
        last = a_lens - 1
        a_ids[:] = a_sampled
        a_pos[:] = a_lens
        a_slots[:] = a_tab[np.arange(bs), last // bsz] * bsz + last % bsz
        return 0


@pytest.mark.parametrize("graphs", [False, True])
def test_device_side_path_host_logic_on_cpu(lib, graphs):
    """The glue's decisions on the device-side decode path (patch lists, -1 padding of new columns, fall-backs on layout
    change / too many patches / changed batch) with the kernel emulated on the CPU staging buffer: every step must
    publish what the reference's prepare_decode builds."""
    import itertools
    ctx = Ctx()
    g = _glue(lib, ctx, max_num_seqs=64)
    g.device_state = True                       # normally CUDA-only
    shim = g.lib = _KernelShim(g.lib)
    if graphs:
        g.graph_bs = [1, 2, 4, 8] + list(range(16, 65, 16))
    free = itertools.count(100)
    ids = itertools.count()

    def mk(n):
        s = Seq([1], [next(free) for _ in range((n + BS - 1) // BS)], 0, n, 1)
        s.seq_id = next(ids)
        return s
    # 20 sequences of EQUAL length cross a block boundary at the same step (more patches than one launch carries)
    seqs = [mk(n) for n in (250, 3, 511, 700)] + [mk(200) for _ in range(20)]
    g.prepare_decode(seqs, None)
    for step in range(1, 400):
        toks = []
        for i, s in enumerate(seqs):
            s.last_token = (step * 31 + i) % 10000
            s.num_tokens += 1
            if s.num_tokens % BS == 1:
                s.block_table.append(next(free))
            toks.append(s.last_token)
        if step == 300:
            seqs = seqs[:3] + seqs[5:]         # two sequences finish: rows shift
            toks = toks[:3] + toks[5:]
        ids_t, pos_t = g.prepare_decode(seqs, torch.tensor(toks, dtype=torch.int64))
        want = gorc.prepare_decode_ref(seqs)
        assert torch.equal(ids_t, want["input_ids"]) and torch.equal(pos_t, want["positions"]), step
        for k in ("slot_mapping", "context_lens", "block_tables"):
            got = getattr(ctx.args, k)
            assert got.dtype == want[k].dtype and torch.equal(got, want[k]), (step, k)
    assert shim.calls == g.fast_steps and g.fast_steps + g.staged_steps == 400
    assert g.fast_steps >= 380 and 1 <= shim.max_patches_seen <= 16


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _reference_sequence_cls():
    import importlib
    import sys
    for cand in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if os.path.isdir(os.path.join(cand, "nanovllm")):
            if cand not in sys.path:
                sys.path.insert(0, cand)
            return importlib.import_module("nanovllm.engine.sequence")
    return None


def test_worker_rank_copies_keep_their_identity_and_take_the_device_path(lib):
    """worker ranks unpickle fresh Sequence objects every step and the reference's __getstate__ drops
    seq_id, so rank > 0 never took the device-side path.  With install_sequence_identity the id travels in the pickle;
    driving the glue with a pickled copy of the batch per step (what ModelRunner.loop hands a worker) must advance on
    the device at every step after the first, and publish the reference's tensors."""
    import copy
    import itertools
    import pickle
    from nano_vllm_learn_b200 import runner_glue as rg
    mod = _reference_sequence_cls()
    if mod is None:
        pytest.skip("reference tree not on this machine")
    Sequence = type("Sequence", (mod.Sequence,), {})       # patch a subclass: the imported reference class stays stock
    Sequence.__module__ = __name__
    globals()["Sequence"] = Sequence                        # picklable by reference from this module
    try:
        plain = pickle.loads(pickle.dumps(Sequence([1, 2, 3])))
        assert not hasattr(plain, "seq_id")                 # the reference behaviour the patch is for
        rg.install_sequence_identity(Sequence)
        rg.install_sequence_identity(Sequence)              # idempotent
        free = itertools.count(10)
        seqs = []
        for n in (250, 3, 511, 700, 256):
            s = Sequence(list(range(n)))
            s.block_table = [next(free) for _ in range((n + BS - 1) // BS)]
            seqs.append(s)
        for s in seqs:                                      # prompt state and decode state both round-trip
            c = pickle.loads(pickle.dumps(s))
            assert c.seq_id == s.seq_id and c.token_ids == s.token_ids and c.block_table == s.block_table
            assert (c.num_tokens, c.num_prompt_tokens, c.num_cached_tokens) == (s.num_tokens, s.num_prompt_tokens, 0)
        ctx = Ctx()
        g = _glue(lib, ctx, max_num_seqs=16)
        g.device_state = True
        shim = g.lib = _KernelShim(g.lib)
        toks = None
        for step in range(300):
            toks = []
            for i, s in enumerate(seqs):
                s.append_token((step * 31 + i) % 10000)
                if s.num_tokens % BS == 1:
                    s.block_table.append(next(free))
                toks.append(s.last_token)
            worker_seqs = pickle.loads(pickle.dumps(seqs))   # fresh objects, as on rank > 0
            assert worker_seqs[0] is not seqs[0] and not hasattr(worker_seqs[0], "token_ids")
            assert [rg.seq_identity(w) for w in worker_seqs] == [s.seq_id for s in seqs]
            ids_t, pos_t = g.prepare_decode(worker_seqs, torch.tensor(toks, dtype=torch.int64))
            want = gorc.prepare_decode_ref(seqs)
            assert torch.equal(ids_t, want["input_ids"]) and torch.equal(pos_t, want["positions"]), step
            for k in ("slot_mapping", "context_lens", "block_tables"):
                assert torch.equal(getattr(ctx.args, k), want[k]), (step, k)
        # staged: the first step, and the one where the longest row opens a 4th block (the eager table gets wider)
        assert g.staged_steps == 2 and g.fast_steps == 298 == shim.calls
        # objects without any id (stock reference pickles): a stable surrogate, never the object's address
        anon = [copy.copy(s) for s in seqs]
        for a in anon:
            del a.seq_id
        assert [rg.seq_identity(a) for a in anon] == [("anon", s.num_prompt_tokens, s.block_table[0]) for s in seqs]
    finally:
        globals().pop("Sequence", None)


def test_device_path_restages_when_a_middle_row_did_not_grow_by_one(lib):
    """only the first and last rows were checked for 'one more token each'."""
    import itertools
    ctx = Ctx()
    g = _glue(lib, ctx, max_num_seqs=16)
    g.device_state = True
    g.lib = _KernelShim(g.lib)
    free = itertools.count(100)
    seqs = []
    for i, n in enumerate((250, 300, 40)):
        s = Seq([1], [next(free) for _ in range((n + BS - 1) // BS)], 0, n, 1)
        s.seq_id = i
        seqs.append(s)
    g.prepare_decode(seqs, None)
    for s in seqs:
        s.num_tokens += 1
    g.prepare_decode(seqs, torch.tensor([1, 2, 3]))
    assert (g.fast_steps, g.staged_steps) == (1, 1)
    seqs[0].num_tokens += 1
    seqs[1].num_tokens += 2                                  # the middle row jumps by two
    seqs[2].num_tokens += 1
    g.prepare_decode(seqs, torch.tensor([4, 5, 6]))
    assert (g.fast_steps, g.staged_steps) == (1, 2)          # restaged from the host ...
    want = gorc.prepare_decode_ref(seqs)
    for k in ("slot_mapping", "context_lens", "block_tables"):
        assert torch.equal(getattr(ctx.args, k), want[k]), k
    for s in seqs:
        s.num_tokens += 1
    g.prepare_decode(seqs, torch.tensor([7, 8, 9]))
    assert (g.fast_steps, g.staged_steps) == (2, 2)          # ... and the device path resumes from the new state
    want = gorc.prepare_decode_ref(seqs)
    for k in ("slot_mapping", "context_lens", "block_tables"):
        assert torch.equal(getattr(ctx.args, k), want[k]), k
