"""The device code of csrc/kernels/sparse.cu (K5) run on the host, thread by thread.

The kernels have no barriers, shuffles or spin-waits, so executing their threads one after the
other is a faithful run of their logic.  The test cuts the kernel bodies out of the real source
(and the optimizer math out of common.cuh), compiles them with g++ against tests/emu/cuda_shim.h
and runs tests/emu/sparse_emu.cpp: 1-8 ranks, three row widths (8 / 16 / 32 lanes per row),
fp32 and bf16 gradients, all four optimizers, duplicates, padding, short and stale staging,
one-CTA grids.  What it cannot see — alignment faults, the float4 atomics, peer addressing — is
what tests/test_zy_sparse_gpu.py runs on the GPU."""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _cut(text: str, start: str, end: str) -> str:
    a = text.index(start)
    return text[a:text.index(end, a)]


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_sparse_kernels_on_the_host(tmp_path):
    kernels = _cut((ROOT / "csrc/kernels/sparse.cu").read_text(),
                   "constexpr int kThreads", "// ---- host side")
    assert "sparse_apply_kernel" in kernels and "grid_for" in kernels and "<<<" not in kernels
    math = _cut((ROOT / "csrc/kernels/common.cuh").read_text(),
                "struct OptimCoef {", "}  // namespace psb")
    assert "update1" in math
    (tmp_path / "sparse_kernels.inc").write_text(kernels)
    (tmp_path / "optim_math.inc").write_text(math)
    exe = tmp_path / "sparse_emu"
    cc = subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Wno-unused-function",
                         "-Wno-unused-variable", "-I", str(tmp_path), "-I", str(ROOT / "tests/emu"),
                         "-I", str(ROOT / "include"), str(ROOT / "tests/emu/sparse_emu.cpp"),
                         "-o", str(exe)], capture_output=True, text=True)
    assert cc.returncode == 0, cc.stderr[-4000:]
    run = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, run.stdout[-2000:] + run.stderr[-2000:]
    assert "ok" in run.stdout
