"""Build the sm_100a C-ABI library in-tree with nvcc (no torch/ATen in the link).

`python nano-vllm-learn_b200/build.py` or `build_library()` writes
`nano-vllm-learn_b200/lib/libnvllm_attn.so`.  The .so is git-ignored but travels with the gpurun
snapshot.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_DIR = PKG_DIR / "lib"
LIB_PATH = LIB_DIR / "libnvllm_attn.so"
STAMP = LIB_DIR / "libnvllm_attn.stamp"
# measurement / debug aids (include/nvllm_tools.h): the product sources + tools/csrc, never loaded by the package
TOOLS_CSRC = PKG_DIR.parent / "tools" / "csrc"
TOOLS_LIB_PATH = LIB_DIR / "libnvllm_tools.so"
TOOLS_STAMP = LIB_DIR / "libnvllm_tools.stamp"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--use_fast_math",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "-shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA C-ABI library cannot be built")


def _sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest(tools: bool = False) -> str:
    h = hashlib.sha256()
    files = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.inl")) + [PKG_DIR.parent / "include" / "nvllm_attn.h"]
    if tools:
        files += list(TOOLS_CSRC.glob("*.cu")) + [PKG_DIR.parent / "include" / "nvllm_tools.h"]
    for f in sorted(files):
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_library(force: bool = False, verbose: bool = False) -> Path:
    digest = _digest()
    if not force and LIB_PATH.exists() and STAMP.exists() and STAMP.read_text().strip() == digest:
        return LIB_PATH
    LIB_DIR.mkdir(exist_ok=True)
    cmd = [_nvcc(), *NVCC_FLAGS]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += ["-o", str(LIB_PATH), *map(str, _sources())]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        print(proc.stderr)
    STAMP.write_text(digest)
    return LIB_PATH


def build_tools_library(force: bool = False) -> Path:
    """lib/libnvllm_tools.so = product sources with -DNVLLM_TOOLS_BUILD (traced prefill instantiation) + tools/csrc."""
    digest = _digest(tools=True)
    if not force and TOOLS_LIB_PATH.exists() and TOOLS_STAMP.exists() and TOOLS_STAMP.read_text().strip() == digest:
        return TOOLS_LIB_PATH
    LIB_DIR.mkdir(exist_ok=True)
    cmd = [_nvcc(), *NVCC_FLAGS, "-DNVLLM_TOOLS_BUILD", "-o", str(TOOLS_LIB_PATH), *map(str, _sources()),
           *map(str, sorted(TOOLS_CSRC.glob("*.cu")))]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{proc.stdout}\n{proc.stderr}")
    TOOLS_STAMP.write_text(digest)
    return TOOLS_LIB_PATH


if __name__ == "__main__":
    path = build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
    if "--tools" in sys.argv:
        print(build_tools_library(force="--force" in sys.argv))
