"""GPU: vocabulary-parallel sampling (second half; csrc/sampler.cu, nvllm_sample_tokens_sharded) vs
the reference's gather path: ParallelLMHead.forward (nanovllm/layers/embed_head.py:61-71: shard GEMM + dist.gather +
torch.cat) followed by Sampler.forward on rank 0 (nanovllm/layers/sampler.py:10-18, engine/model_runner.py:311).

Token ids are index work: the bar is exact equality -- with the fixture frozen from the reference classes under gloo
(tests/golden/lm_head_sampler.npz) on the reference's own noise draw, and with the unsharded kernel on the concatenated
logits in Philox mode.  "Loop-back" tests run all `world` ranks on ONE GPU (each rank = its own exchange buffer, stream
and workspace; the kernels are co-resident and swap pairs through ordinary device memory), so the exchange protocol is
covered on the 1-GPU box; the peer-memory tests need >= 2 GPUs.
"""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import sampler_oracle as so

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "lm_head_sampler.npz"))


def _comms(devices):
    from nano_vllm_learn_b200 import sampler_cuda
    from nano_vllm_learn_b200.rowparallel_cuda import TPComm
    world = len(devices)
    return TPComm.single_process(devices, sampler_cuda.comm_bytes(world))


def _run_world(comms, devices, shards, temps, noise=None, seed=0, offset=0, temps_on_workers=True, streams=None):
    """One call of every rank (rank 0 first); returns the list of per-rank token tensors."""
    from nano_vllm_learn_b200 import ops
    world = len(devices)
    outs = [None] * world
    off = 0
    for r in range(world):
        dev = torch.device("cuda", devices[r])
        v = shards[r].shape[1]
        with torch.cuda.device(dev), torch.cuda.stream(streams[r]):
            t = temps[r] if (r == 0 or temps_on_workers) else None
            n = None if noise is None else noise[r]
            outs[r] = ops.sample_tokens_sharded(shards[r], t, off, comms[r], noise=n, seed=seed, offset=offset)
        off += v
    return outs


def _setup(devices, logits_cpu, temps_cpu, noise_cpu=None):
    world = len(devices)
    shards, temps, noise, streams = [], [], [], []
    for r, chunk in enumerate(logits_cpu.chunk(world, dim=1)):
        dev = torch.device("cuda", devices[r])
        shards.append(chunk.contiguous().to(dev))
        temps.append(temps_cpu.to(dev))
        streams.append(torch.cuda.Stream(device=dev))
    if noise_cpu is not None:
        noise = [c.contiguous().to(torch.device("cuda", devices[r])) for r, c in enumerate(noise_cpu.chunk(world, dim=1))]
    else:
        noise = None
    return shards, temps, noise, streams


def _sync(devices):
    for d in set(devices):
        torch.cuda.synchronize(d)


def _case_logits(c):
    last = None if c["cu_seqlens_q"] is None else (c["cu_seqlens_q"][1:] - 1).long()
    return so.lm_head_gather_ref(so.lm_head_shards_ref(c["x"], c["weight"], c["world"], last))


def _devices(world, spread):
    if not spread:
        return [0] * world
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    return [r % n for r in range(world)]


@pytest.mark.parametrize("spread", [False, True], ids=["loopback", "peer"])
@pytest.mark.parametrize("name", list(cases.LM_HEAD_CASES))
def test_reference_noise_matches_gather_fixture(lib, name, spread):
    c = cases.make_lm_head_case(name)
    logits = _case_logits(c)
    devices = _devices(c["world"], spread)
    noise = so.reference_noise(logits.shape, c["noise_seed"])
    shards, temps, nz, streams = _setup(devices, logits, c["temperatures"], noise)
    comms = _comms(devices)
    outs = _run_world(comms, devices, shards, temps, noise=nz, streams=streams)
    _sync(devices)
    want = GOLD[f"{name}_tokens"].tolist()
    for r, o in enumerate(outs):
        assert o.dtype == torch.int64 and o.cpu().tolist() == want, f"rank {r}"
    from nano_vllm_learn_b200 import sampler_cuda
    assert sampler_cuda.poll_error() == 0
    assert all(cm.status()[1] == 0 for cm in comms)


@pytest.mark.parametrize("spread", [False, True], ids=["loopback", "peer"])
@pytest.mark.parametrize("world,bs,vocab,dtype", [(2, 16, 151936, torch.bfloat16), (8, 64, 151936, torch.bfloat16),
                                                  (4, 7, 4096, torch.float32)])
def test_philox_mode_equals_unsharded_kernel(lib, world, bs, vocab, dtype, spread):
    """Same (seed, offset): the sharded kernels draw the very noise the unsharded kernel draws (global counters), so
    the tokens are identical on every rank -- greedy and sampled rows, many calls back to back, no host sync between."""
    from nano_vllm_learn_b200 import ops
    devices = _devices(world, spread)
    g = torch.Generator().manual_seed(world * 100 + bs)
    logits = (torch.randn(bs, vocab, generator=g) * 3).to(dtype)
    temps = torch.tensor([0.0 if i % 3 == 0 else 0.5 + 0.1 * (i % 7) for i in range(bs)], dtype=torch.float32)
    shards, tl, _, streams = _setup(devices, logits, temps)
    comms = _comms(devices)
    full = logits.to("cuda:0")
    runs = []
    for call in range(6):
        outs = _run_world(comms, devices, shards, tl, seed=77, offset=call + 1, streams=streams)
        runs.append(outs)
    _sync(devices)
    for call, outs in enumerate(runs):
        want = ops.sample_tokens(full, tl[0].to("cuda:0"), seed=77, offset=call + 1).cpu().tolist()
        for r, o in enumerate(outs):
            assert o.cpu().tolist() == want, f"call {call} rank {r}"
    greedy = logits.float().argmax(-1).tolist()
    for i in range(bs):
        if temps[i] == 0:
            assert runs[0][0][i].item() == greedy[i]
    assert all(cm.status() == (6, 0) for cm in comms)


@pytest.mark.parametrize("spread", [False, True], ids=["loopback", "peer"])
def test_workers_take_the_temperatures_from_rank0(lib, spread):
    """The reference's worker ranks never learn the temperatures (sequence.py:77-83 does not pickle them,
    model_runner.py:309 prepares them on rank 0 only): ranks > 0 call with temperatures=None."""
    from nano_vllm_learn_b200 import ops
    world, bs, vocab = 2, 4, 8192
    devices = _devices(world, spread)
    g = torch.Generator().manual_seed(31)
    logits = (torch.randn(bs, vocab, generator=g) * 2).to(torch.bfloat16)
    temps = torch.tensor([0.0, 0.9, 1.1, 0.0])
    shards, tl, _, streams = _setup(devices, logits, temps)
    comms = _comms(devices)
    for call in range(4):
        outs = _run_world(comms, devices, shards, tl, seed=5, offset=call, temps_on_workers=False, streams=streams)
        _sync(devices)
        want = ops.sample_tokens(logits.to("cuda:0"), temps.to("cuda:0"), seed=5, offset=call).cpu().tolist()
        assert [o.cpu().tolist() for o in outs] == [want] * world


def test_varying_batch_sizes_and_pinned_host_output(lib):
    """Decode batches shrink as sequences finish; rank 0 reads the ids from pinned host memory (Sampler.forward_sharded)."""
    from nano_vllm_learn_b200 import ops
    from nano_vllm_learn_b200.sampler_cuda import Sampler
    world, vocab = 2, 4096
    devices = [0, 0]
    comms = _comms(devices)
    samplers = [Sampler() for _ in range(world)]
    for s, cm in zip(samplers, comms):
        s.bind_comm(cm)
    streams = [torch.cuda.Stream() for _ in range(world)]
    g = torch.Generator().manual_seed(3)
    for bs in (9, 4, 4, 1, 12):
        logits = (torch.randn(bs, vocab, generator=g) * 2).to(torch.bfloat16).cuda()
        temps = torch.zeros(bs, device="cuda")
        outs = [None, None]
        # the worker is launched first here: it does not wait for temperatures (passed explicitly) and rank 0's
        # to_host path synchronises its own stream
        for r in (1, 0):
            with torch.cuda.stream(streams[r]):
                streams[r].wait_stream(torch.cuda.default_stream())
                outs[r] = samplers[r].forward_sharded(logits[:, r * vocab // 2:(r + 1) * vocab // 2], temps, to_host=r == 0)
        torch.cuda.synchronize()
        want = logits.float().argmax(-1).cpu().tolist()
        assert not outs[0].is_cuda and outs[0].is_pinned() and outs[0].tolist() == want
        assert outs[1].is_cuda and outs[1].cpu().tolist() == want


def test_argument_validation(lib):
    from nano_vllm_learn_b200 import ops
    comms = _comms([0, 0])
    x = torch.zeros(2, 1024, device="cuda", dtype=torch.bfloat16)
    t = torch.zeros(2, device="cuda")
    with pytest.raises(ValueError):
        ops.sample_tokens_sharded(x, t, 1022, comms[0])            # offset not a multiple of 4
    with pytest.raises(AssertionError):
        ops.sample_tokens_sharded(x, None, 0, comms[0])            # rank 0 must bring the temperatures
