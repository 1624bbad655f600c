"""GPU: fused token sampler (csrc/sampler.cu) vs the reference Sampler.

Token ids are index work: the bar is exact equality.  Greedy rows never depend on noise.  Sampled rows are
compared on the SAME exponential draw: either the reference's own `exponential_` draw injected into the
kernel, or the kernel's Philox stream regenerated on the host and fed to the reference formula."""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import sampler_oracle as so

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "sampler.npz"))


@pytest.mark.parametrize("name", list(cases.SAMPLER_CASES))
def test_injected_reference_noise_matches_fixture(lib, name):
    from nano_vllm_learn_b200 import ops
    c = cases.make_sampler_case(name)
    noise = so.reference_noise(c["logits"].shape, c["noise_seed"])
    got = ops.sample_tokens(c["logits"].to(DEV), c["temperatures"].to(DEV), noise=noise.to(DEV))
    assert got.dtype == torch.int64 and got.shape == (c["logits"].shape[0],)
    assert got.cpu().tolist() == GOLD[f"{name}_tokens"].tolist()


@pytest.mark.parametrize("name", list(cases.SAMPLER_CASES))
def test_philox_mode_matches_reference_formula_on_the_same_draw(lib, name):
    from nano_vllm_learn_b200 import ops
    c = cases.make_sampler_case(name)
    bs, vocab = c["logits"].shape
    for seed, offset in ((1234, 1), (2 ** 40 + 17, 2 ** 33 + 5)):
        got = ops.sample_tokens(c["logits"].to(DEV), c["temperatures"].to(DEV), seed=seed, offset=offset)
        want = so.sampler_ref(c["logits"], c["temperatures"], so.philox_exponential(seed, offset, bs, vocab))
        assert got.cpu().tolist() == want.tolist()


def test_unaligned_rows_take_the_scalar_path(lib):
    from nano_vllm_learn_b200 import ops
    g = torch.Generator().manual_seed(5)
    big = torch.randn(7, 1003, generator=g).to(torch.bfloat16)
    view = big[:, 1:1000]                      # base pointer 2 bytes off, row stride 1003 elements
    t = torch.tensor([0.0, 1.0, 0.0, 0.8, 0.0, 1.5, 0.0])
    noise = so.reference_noise((7, 999), 77)
    got = ops.sample_tokens(big.to(DEV)[:, 1:1000], t.to(DEV), noise=noise.to(DEV))
    assert got.cpu().tolist() == so.sampler_ref(view, t, noise).tolist()


def test_full_vocab_greedy_batch_vs_reference_module(lib):
    """bench_my's configuration: temperature 0, Qwen3 vocabulary, bs 64 — against the reference class on the GPU."""
    from nano_vllm_learn_b200.sampler_cuda import Sampler
    g = torch.Generator().manual_seed(9)
    logits = (torch.randn(64, 151936, generator=g) * 3).to(torch.bfloat16).to(DEV)
    t = torch.zeros(64, device=DEV)
    got = Sampler()(logits, t)
    assert torch.equal(got, logits.float().argmax(-1))
    ref_cls = so.load_reference_sampler()
    if ref_cls is not None:
        assert torch.equal(got, ref_cls()(logits.clone(), t))
    # a second call reuses the self-resetting workspace counters
    assert torch.equal(Sampler()(logits, t), got)


def test_nan_and_inf_rows(lib):
    from nano_vllm_learn_b200 import ops
    x = torch.zeros(3, 4096)
    x[0, 100] = float("nan")
    x[0, 50] = float("inf")
    x[1, :] = float("-inf")
    x[1, 7] = -5.0
    x[2, 9] = float("inf")
    x[2, 4000] = float("inf")
    t = torch.zeros(3)
    got = ops.sample_tokens(x.to(DEV), t.to(DEV))
    assert got.cpu().tolist() == x.argmax(-1).tolist() == [100, 7, 9]


def test_sampled_distribution_matches_softmax(lib):
    """Philox mode draws from softmax(logits / T): chi-square over 4096 rows x 16 calls of a 24-way distribution."""
    from nano_vllm_learn_b200 import ops
    g = torch.Generator().manual_seed(3)
    row = torch.randn(24, generator=g) * 1.5
    temp = 0.8
    probs = torch.softmax(row / temp, -1).double()
    logits = row.repeat(4096, 1).to(DEV)
    t = torch.full((4096,), temp, device=DEV)
    counts = torch.zeros(24, dtype=torch.double)
    calls = 16
    for call in range(calls):
        tok = ops.sample_tokens(logits, t, seed=99, offset=call)
        counts += torch.bincount(tok.cpu(), minlength=24).double()
    n = 4096 * calls
    chi2 = float(((counts - n * probs) ** 2 / (n * probs)).sum())
    assert chi2 < 60.0, chi2          # 23 degrees of freedom: P(chi2 > 60) < 4e-5
    # different offsets must give different draws
    a = ops.sample_tokens(logits, t, seed=99, offset=100)
    b = ops.sample_tokens(logits, t, seed=99, offset=101)
    assert not torch.equal(a, b)


def test_argument_errors(lib):
    from nano_vllm_learn_b200 import ops
    with pytest.raises(RuntimeError):
        ops.sample_tokens(torch.zeros(2, 8), torch.zeros(2))          # CPU tensors: no fallback
    with pytest.raises(AssertionError):
        ops.sample_tokens(torch.zeros(2, 8, device=DEV, dtype=torch.float16), torch.zeros(2, device=DEV))
