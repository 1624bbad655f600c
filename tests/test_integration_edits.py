"""The reference-side binding (INTEGRATION.md edits A1-A8, tools/make_integration.py) applies to the UNMODIFIED
reference: every anchored replacement matches exactly once, the edited files still compile, and nothing outside the
documented files changes -- so `attn_backend="flash"` on the edited copy is the reference's stock code path.
Runs where the reference install (baseline/_ref) or /root/reference exists; no GPU, nothing is imported from the copy.
"""
import filecmp
import os
import py_compile
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def _ref_src():
    for cand in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if os.path.isdir(os.path.join(cand, "nanovllm")):
            return cand
    return None


@pytest.fixture(scope="module")
def edited(tmp_path_factory):
    src = _ref_src()
    if src is None:
        pytest.skip("reference tree not on this machine")
    import make_integration
    dst = str(tmp_path_factory.mktemp("nanovllm_cuda"))
    make_integration.make(src, dst)
    return src, dst, make_integration


def test_every_edit_applies_once_and_compiles(edited):
    src, dst, mi = edited
    touched = {rel for rel, _, _ in mi.EDITS}
    assert touched == {"nanovllm/config.py", "nanovllm/models/qwen3.py", "nanovllm/engine/model_runner.py",
                       "nanovllm/layers/linear.py", "nanovllm/layers/embed_head.py"}
    for rel in sorted(touched) + ["nanovllm/layers/attention_cuda.py"]:
        py_compile.compile(os.path.join(dst, rel), doraise=True)
    for rel, old, new in mi.EDITS:
        text = open(os.path.join(dst, rel)).read()
        assert new in text
        if old is not None:
            assert open(os.path.join(src, rel)).read().count(old) == 1


def test_nothing_else_changes(edited):
    src, dst, mi = edited
    touched = {os.path.normpath(rel) for rel, _, _ in mi.EDITS} | {os.path.normpath("nanovllm/layers/attention_cuda.py")}
    seen = 0
    for base, dirs, files in os.walk(os.path.join(src, "nanovllm")):
        dirs[:] = [d for d in dirs if d != "__pycache__"]
        for f in files:
            rel = os.path.normpath(os.path.relpath(os.path.join(base, f), src))
            if rel in touched or f.endswith(".pyc"):
                continue
            assert filecmp.cmp(os.path.join(base, f), os.path.join(dst, rel), shallow=False), rel
            seen += 1
    assert seen >= 15   # attention.py, attention_triton.py, attention_sdpa.py, block_manager.py, ... byte-identical


def test_other_backends_dispatch_is_untouched(edited):
    """the flash / sdpa / triton branches of the dispatch and RowParallelLinear.forward's body are the reference's lines"""
    src, dst, _ = edited
    ref = open(os.path.join(src, "nanovllm/models/qwen3.py")).read().splitlines()
    new = open(os.path.join(dst, "nanovllm/models/qwen3.py")).read().splitlines()
    assert [l for l in ref if l in new] == ref            # edits only ADD lines
    added = [l for l in new if l not in ref]
    assert all("cuda" in l or "forward_fused" in l or "block_size" in l for l in added), added
    ref = open(os.path.join(src, "nanovllm/layers/linear.py")).read()
    new = open(os.path.join(dst, "nanovllm/layers/linear.py")).read()
    assert new.startswith(ref) and "rowparallel_cuda" in new[len(ref):]
