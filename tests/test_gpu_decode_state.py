"""GPU: device-side prepare_decode (csrc/decode_state.cu + RunnerGlue._advance_on_device).

While the running batch is unchanged the glue stages nothing on the host: one kernel advances input_ids / positions /
context_lens / slot_mapping / block_tables in the buffers the captured graph reads.  Integer work: every published
tensor must be bit-identical, at EVERY step, to what the reference's prepare_decode builds from the sequences
(nanovllm/engine/model_runner.py:244-269) -- checked against the plain-Python restatement (oracle) and, where the
reference is importable, against the unmodified ModelRunner.prepare_decode driven by the real BlockManager/Sequence."""
import itertools
from types import SimpleNamespace

import pytest
import torch

from oracle import runner_glue_oracle as gorc

pytestmark = pytest.mark.gpu
BS = 256


class Seq:
    """Duck-typed nanovllm.engine.sequence.Sequence: what prepare_decode reads (sequence.py:14-69)."""
    _ids = itertools.count()

    def __init__(self, num_tokens, block_table):
        self.seq_id = next(Seq._ids)
        self.num_tokens = num_tokens
        self.block_table = list(block_table)
        self.last_token = 1
        self.num_cached_tokens = 0

    def __len__(self):
        return self.num_tokens

    @property
    def last_block_num_tokens(self):
        return self.num_tokens - (len(self.block_table) - 1) * BS


class Ctx:
    def set(self, is_prefill, cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=0, max_seqlen_k=0,
            slot_mapping=None, context_lens=None, block_tables=None):
        self.args = SimpleNamespace(is_prefill=is_prefill, slot_mapping=slot_mapping, context_lens=context_lens,
                                    block_tables=block_tables)


def _glue(ctx, max_model_len=4096):
    from nano_vllm_learn_b200.runner_glue import RunnerGlue
    return RunnerGlue(block_size=BS, max_num_seqs=64, max_model_len=max_model_len, max_num_batched_tokens=16384,
                      device="cuda", set_context=ctx.set)


def _check(ctx, ids, pos, want, step):
    torch.cuda.synchronize()
    assert ids.dtype == torch.int64 and torch.equal(ids.cpu(), want["input_ids"]), step
    assert pos.dtype == torch.int64 and torch.equal(pos.cpu(), want["positions"]), step
    for k in ("slot_mapping", "context_lens", "block_tables"):
        got = getattr(ctx.args, k)
        assert got.dtype == torch.int32 and torch.equal(got.cpu(), want[k]), (step, k)


def _step(seqs, free_blocks, step):
    """What Scheduler.postprocess + BlockManager.may_append do to a running sequence between two decode steps."""
    toks = []
    for i, s in enumerate(seqs):
        t = (step * 131 + i * 7) % 10000
        s.last_token = t
        s.num_tokens += 1
        if s.num_tokens % BS == 1:                 # block_manager.py:134-159: a new block for the token just appended
            s.block_table.append(free_blocks.pop())
        toks.append(t)
    return torch.tensor(toks, dtype=torch.int64, device="cuda")


@pytest.mark.parametrize("graphs", [False, True], ids=["eager_layout", "graph_layout"])
def test_unchanged_batch_advances_on_the_device(lib, graphs):
    ctx = Ctx()
    g = _glue(ctx)
    if graphs:
        g.graph_bs = [1, 2, 4, 8]
    free = list(range(500, 100, -1))
    # lengths chosen so rows cross block boundaries at different steps, two of them at the SAME step
    seqs = [Seq(n, [free.pop() for _ in range((n + BS - 1) // BS)]) for n in (250, 3, 511, 250, 700)]
    ids, pos = g.prepare_decode(seqs)                      # first decode step after a prefill: staged from the host
    _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), 0)
    for step in range(1, 600):
        toks = _step(seqs, free, step)
        ids, pos = g.prepare_decode(seqs, toks)
        _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), step)
    if graphs:
        assert g.fast_steps == 599 and g.staged_steps == 1
        # zero padding of the captured layout outside [:bs, :width] is untouched (model_runner.py:291-293)
        _, _, slots, lens, tab = g.graph_buffers(8)
        assert int(lens[5:].abs().sum()) == 0 and int(tab[5:].abs().sum()) == 0
    else:
        # the tight [bs, width] layout changes whenever the widest table grows: those steps are re-staged
        assert g.fast_steps >= 590 and g.fast_steps + g.staged_steps == 600


def test_batch_change_prefill_and_missing_tokens_fall_back(lib):
    ctx = Ctx()
    g = _glue(ctx)
    g.graph_bs = [1, 2, 4, 8]
    free = list(range(500, 100, -1))
    seqs = [Seq(n, [free.pop() for _ in range((n + BS - 1) // BS)]) for n in (10, 300, 255)]
    g.prepare_decode(seqs)
    toks = _step(seqs, free, 1)
    ids, pos = g.prepare_decode(seqs, toks)
    assert g.fast_steps == 1
    _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), 1)
    # a sequence finishes: the rows shift, everything is re-staged
    seqs = [seqs[0], seqs[2]]
    toks = _step(seqs, free, 2)
    ids, pos = g.prepare_decode(seqs, toks)
    assert g.fast_steps == 1
    _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), 2)
    toks = _step(seqs, free, 3)
    ids, pos = g.prepare_decode(seqs, toks)
    assert g.fast_steps == 2
    _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), 3)
    # no device tokens (reference sampler / worker rank without vocabulary-parallel sampling): staged
    _step(seqs, free, 4)
    ids, pos = g.prepare_decode(seqs, None)
    assert g.fast_steps == 2
    _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), 4)
    # a prefill in between overwrites the staging buffer: the next decode step must not trust the device state
    toks = _step(seqs, free, 5)
    new = Seq(40, [free.pop()])
    new.token_ids = list(range(40))
    new.__class__ = type("SeqP", (Seq,), {"__getitem__": lambda self, k: self.token_ids[k]})
    g.prepare_prefill([new])
    ids, pos = g.prepare_decode(seqs, toks)
    assert g.fast_steps == 2
    _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), 5)
    # a sequence that did NOT grow by exactly one token is detected (first / last row probes) and re-staged
    toks = _step(seqs, free, 6)
    seqs[0].num_tokens += 1
    ids, pos = g.prepare_decode(seqs, toks)
    assert g.fast_steps == 2
    _check(ctx, ids, pos, gorc.prepare_decode_ref(seqs), 6)


@pytest.fixture
def unmodified_reference():
    """The UNMODIFIED reference package, imported for this test only: later tests import the edited integration copy
    under the same package name (tests/test_gpu_engine_e2e.py), so every trace of it is removed again."""
    import sys
    root = gorc.find_reference_root()
    if root is None:
        pytest.skip("reference not present")
    before = set(sys.modules)
    path_before = list(sys.path)
    try:
        try:
            yield gorc.ReferenceGlue(root)
        except ImportError as e:
            pytest.skip(f"reference not importable: {e}")
    finally:
        for name in [m for m in sys.modules if m not in before and (m == "nanovllm" or m.startswith("nanovllm."))]:
            del sys.modules[name]
        sys.path[:] = path_before


def test_live_reference_trace_on_device(lib, unmodified_reference):
    """The REAL BlockManager / Sequence / ModelRunner.prepare_decode (unmodified), 520 decode steps, every step."""
    ref = unmodified_reference
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
    g = _glue(ctx)
    g.graph_bs = [1, 2, 4, 8]
    g.prepare_prefill(seqs)
    toks = None
    for step in range(1, 521):
        new = [(step * 13 + i) % 10000 for i in range(len(seqs))]
        for s, t in zip(seqs, new):
            s.append_token(t)
            bm.may_append(s)
        want = ref.prepare_decode(seqs)
        ids, pos = g.prepare_decode(seqs, toks)
        toks = None
        _check(ctx, ids, pos, want, step)
        toks = torch.tensor([(( step + 1) * 13 + i) % 10000 for i in range(len(seqs))], dtype=torch.int64, device="cuda")
    assert g.fast_steps == 519 and g.staged_steps == 1


def test_missing_block_parks_the_row_and_reaches_the_host(lib):
    """a new token that lands in a table column holding -1 (its patch never arrived) used to become a
    negative slot and an out-of-range page.  The kernel now parks that row (slot -1, context length 0), writes
    1 + row to a mapped host word, and leaves the other rows exactly as the reference computes them."""
    import ctypes
    dev = "cuda"
    bs, width = 3, 2
    lens = torch.tensor([10, BS, 300], dtype=torch.int32, device=dev)          # row 1's next token opens column 1
    tab = torch.tensor([[7, -1], [8, -1], [9, 11]], dtype=torch.int32, device=dev)
    ids = torch.zeros(bs, dtype=torch.int64, device=dev)
    pos = torch.zeros(bs, dtype=torch.int64, device=dev)
    slots = torch.zeros(bs, dtype=torch.int32, device=dev)
    sampled = torch.tensor([101, 102, 103], dtype=torch.int64, device=dev)
    err = ctypes.c_uint32(7)
    assert lib.nvllm_advance_decode_state_poll_error(ctypes.byref(err), 1) == 0
    empty = (ctypes.c_int32 * 1)()
    rc = lib.nvllm_advance_decode_state(ids.data_ptr(), pos.data_ptr(), slots.data_ptr(), lens.data_ptr(), tab.data_ptr(),
                                        width, sampled.data_ptr(), bs, BS, empty, empty, empty, 0, 0, 0,
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    assert ids.tolist() == [101, 102, 103] and pos.tolist() == [11, BS + 1, 301]
    assert lens.tolist() == [11, 0, 301]                       # parked: the decode kernel reads nothing for row 1
    assert slots.tolist() == [7 * BS + 10, -1, 11 * BS + 300 - BS]   # -1: store_kvcache skips it
    assert lib.nvllm_advance_decode_state_poll_error(ctypes.byref(err), 0) == 0 and err.value == 1 + 1
    assert lib.nvllm_advance_decode_state_poll_error(ctypes.byref(err), 1) == 0
    assert lib.nvllm_advance_decode_state_poll_error(ctypes.byref(err), 0) == 0 and err.value == 0
    # the glue reads the word before its next device-side step, raises, and re-stages from the host afterwards
    from nano_vllm_learn_b200 import _lib
    ctx = Ctx()
    g = _glue(ctx)
    free = list(range(500, 100, -1))
    seqs = [Seq(n, [free.pop() for _ in range((n + BS - 1) // BS)]) for n in (10, 300)]
    g.prepare_decode(seqs)
    g.prepare_decode(seqs, _step(seqs, free, 1))
    assert g.fast_steps == 1
    lens.copy_(torch.tensor([10, BS, 300], dtype=torch.int32))                # some launch parks a row ...
    rc = lib.nvllm_advance_decode_state(ids.data_ptr(), pos.data_ptr(), slots.data_ptr(), lens.data_ptr(), tab.data_ptr(),
                                        width, sampled.data_ptr(), bs, BS, empty, empty, empty, 0, 0, 0,
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    toks = _step(seqs, free, 2)
    with pytest.raises(_lib.NvllmError, match="no block for its new token"):  # ... and the next step says so
        g.prepare_decode(seqs, toks)
    ids2, pos2 = g.prepare_decode(seqs, toks)                                  # word cleared, state dropped: staged
    assert g.fast_steps == 1
    _check(ctx, ids2, pos2, gorc.prepare_decode_ref(seqs), 2)
