"""GPU: fused VocabParallelEmbedding (third collective; csrc/vocab_embedding.cu) vs the reference
class (nanovllm/layers/embed_head.py:35-43: masked lookup + dist.all_reduce), frozen under gloo in
tests/golden/vocab_embedding.npz.  A pure copy: the bar is bit equality on every rank.  "Loop-back" runs all ranks on one
GPU (own exchange buffer + stream per rank); the peer tests need >= 2 GPUs."""
import os

import numpy as np
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "vocab_embedding.npz"))


def _devices(world, spread):
    if not spread:
        return [0] * world
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    return [r % n for r in range(world)]


def _world(devices, weight, hidden):
    from nano_vllm_learn_b200 import embedding_cuda
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    world = len(devices)
    comms = TPComm.single_process(devices, embedding_cuda.comm_bytes(world, hidden))
    shards = [w.contiguous().to(torch.device("cuda", devices[r])) for r, w in enumerate(weight.chunk(world, dim=0))]
    streams = [torch.cuda.Stream(device=torch.device("cuda", devices[r])) for r in range(world)]
    return comms, shards, streams


def _call(devices, comms, shards, streams, x_cpu, order=None):
    from nano_vllm_learn_b200 import ops
    world = len(devices)
    per = shards[0].shape[0]
    outs = [None] * world
    for r in (order or range(world)):
        dev = torch.device("cuda", devices[r])
        with torch.cuda.device(dev), torch.cuda.stream(streams[r]):
            outs[r] = ops.vocab_parallel_embedding(x_cpu.to(dev), shards[r], r * per, comms[r])
    return outs


@pytest.mark.parametrize("spread", [False, True], ids=["loopback", "peer"])
@pytest.mark.parametrize("name", list(cases.EMBEDDING_CASES))
def test_matches_reference_fixture(lib, name, spread):
    from nano_vllm_learn_b200 import embedding_cuda
    c = cases.make_embedding_case(name)
    devices = _devices(c["world"], spread)
    comms, shards, streams = _world(devices, c["weight"], c["weight"].shape[1])
    outs = _call(devices, comms, shards, streams, c["x"])
    for d in set(devices):
        torch.cuda.synchronize(d)
    want = cases.from_bits(GOLD[f"{name}_y"])
    for r, y in enumerate(outs):
        assert y.dtype == torch.bfloat16 and torch.equal(y.cpu().view(torch.int16), want.view(torch.int16)), f"rank {r}"
    assert embedding_cuda.poll_error() == 0 and all(cm.status()[1] == 0 for cm in comms)


@pytest.mark.parametrize("spread", [False, True], ids=["loopback", "peer"])
def test_many_calls_without_host_sync_and_a_rank_that_owns_everything(lib, spread):
    """Back-to-back calls with changing batch sizes; in half of them every token belongs to ONE rank, which therefore
    waits for no data and could run ahead -- the 'entered' handshake must keep it from overwriting unread rows."""
    from nano_vllm_learn_b200 import embedding_cuda
    world, vocab, hidden = 2, 2048, 512
    devices = _devices(world, spread)
    g = torch.Generator().manual_seed(17)
    weight = torch.randn(vocab, hidden, generator=g).to(torch.bfloat16)
    comms, shards, streams = _world(devices, weight, hidden)
    calls = []
    for i in range(40):
        n = [1, 7, 64, 300, 512][i % 5]
        lo, hi = (0, vocab) if i % 2 == 0 else ((vocab // 2, vocab) if i % 4 == 1 else (0, vocab // 2))
        x = torch.randint(lo, hi, (n,), generator=g)
        # alternate which rank is launched first
        calls.append((x, _call(devices, comms, shards, streams, x, order=[1, 0] if i % 3 == 0 else [0, 1])))
    for d in set(devices):
        torch.cuda.synchronize(d)
    for i, (x, outs) in enumerate(calls):
        want = weight[x]
        for r, y in enumerate(outs):
            assert torch.equal(y.cpu().view(torch.int16), want.view(torch.int16)), (i, r)
    assert embedding_cuda.poll_error() == 0
    assert all(cm.status() == (40, 0) for cm in comms)


def test_ids_no_rank_owns_give_zero_rows_at_once(lib):
    """an id outside [0, world * vocab_per_rank) has no owner; every rank used to wait the full 20 s timeout
    for a row nobody sends.  The reference answers zeros (no mask matches, embed_head.py:35-43) and so does the kernel,
    without any wait and without latching an error; owned ids in the same call are unaffected."""
    import time
    from nano_vllm_learn_b200 import embedding_cuda
    world, vocab, hidden = 2, 1024, 256
    g = torch.Generator().manual_seed(23)
    weight = torch.randn(vocab, hidden, generator=g).to(torch.bfloat16)
    comms, shards, streams = _world([0, 0], weight, hidden)
    x = torch.tensor([5, vocab, 700, -1, vocab + 12345, 0, vocab - 1], dtype=torch.int64)
    want = torch.zeros(len(x), hidden, dtype=torch.bfloat16)
    owned = (x >= 0) & (x < vocab)
    want[owned] = weight[x[owned]]
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(3):
        outs = _call([0, 0], comms, shards, streams, x)
    torch.cuda.synchronize()
    assert time.time() - t0 < 5.0
    for r, y in enumerate(outs):
        assert torch.equal(y.cpu().view(torch.int16), want.view(torch.int16)), f"rank {r}"
    assert embedding_cuda.poll_error() == 0 and all(cm.status() == (3, 0) for cm in comms)


def test_graph_replay_loopback(lib):
    """The decode graph captures the embedding: the call epoch lives on the device, so replays need no host help."""
    from nano_vllm_learn_b200 import ops
    world, vocab, hidden, n = 2, 1024, 256, 16
    g = torch.Generator().manual_seed(5)
    weight = torch.randn(vocab, hidden, generator=g).to(torch.bfloat16)
    comms, shards, streams = _world([0, 0], weight, hidden)
    xs = [torch.zeros(n, dtype=torch.int64, device="cuda") for _ in range(world)]
    ys = [torch.empty(n, hidden, dtype=torch.bfloat16, device="cuda") for _ in range(world)]
    graphs = []
    for r in range(world):
        with torch.cuda.stream(streams[r]):
            ops.vocab_parallel_embedding(xs[r], shards[r], r * (vocab // 2), comms[r], out=ys[r])   # warm-up, epoch 1
    torch.cuda.synchronize()
    for r in range(world):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=streams[r]):
            ops.vocab_parallel_embedding(xs[r], shards[r], r * (vocab // 2), comms[r], out=ys[r])
        graphs.append(gr)
    for it in range(5):
        x = torch.randint(0, vocab, (n,), generator=g)
        for r in range(world):
            xs[r].copy_(x)
        torch.cuda.synchronize()
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                graphs[r].replay()
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(ys[r].cpu().view(torch.int16), weight[x].view(torch.int16)), (it, r)


def test_argument_validation(lib):
    from nano_vllm_learn_b200 import ops
    weight = torch.zeros(512, 256, dtype=torch.bfloat16)
    comms, shards, _ = _world([0, 0], weight, 256)
    with pytest.raises(ValueError):
        ops.vocab_parallel_embedding(torch.zeros(513, dtype=torch.int64, device="cuda"), shards[0], 0, comms[0])
    with pytest.raises(AssertionError):
        ops.vocab_parallel_embedding(torch.zeros(4, dtype=torch.int32, device="cuda"), shards[0], 0, comms[0])
