"""In-tree build of the native pieces (no JIT cache: the .so travels with the repo).

* ``libpsb_kernels.so`` — every sm_100a CUDA kernel plus the symmetric-heap fabric
  (``csrc/kernels/*.cu``, ``csrc/fabric/*.cu``), C ABI in ``include/psb_kernels.h``.
* ``build/bin/{coordinator,parameter_server,worker_main}`` — the C++ control plane
  (``csrc/wire``, ``csrc/control``, ``csrc/ps``, ``csrc/worker``), CLI-compatible with the
  reference binaries (reference: CMakeLists.txt:116-174).

nvcc cross-compiles for sm_100a without a GPU, so this runs on the CPU-only build box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
NATIVE_DIR = Path(__file__).resolve().parent / "_native"
OBJ_DIR = ROOT / "build" / "obj"
BIN_DIR = ROOT / "build" / "bin"
LIB_PATH = NATIVE_DIR / "libpsb_kernels.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", str(ROOT / "include"),
]

CUDA_SOURCES = sorted((ROOT / "csrc" / "kernels").glob("*.cu")) + sorted(
    (ROOT / "csrc" / "fabric").glob("*.cu"))


def _nvcc() -> str:
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; the CUDA kernels cannot be built")
    return nvcc


def _stamp(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(str(p).encode())
        h.update(Path(p).read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _headers():
    return list((ROOT / "include").glob("*.h")) + list((ROOT / "csrc").rglob("*.cuh")) + list(
        (ROOT / "csrc").rglob("*.h"))


def build_kernels(force: bool = False, verbose: bool = False) -> Path:
    """Compile every .cu for sm_100a and link libpsb_kernels.so (incremental)."""
    NATIVE_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    stamp_file = NATIVE_DIR / "libpsb_kernels.stamp"
    stamp = _stamp(list(CUDA_SOURCES) + _headers())
    if not force and LIB_PATH.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return LIB_PATH
    nvcc = _nvcc()

    def compile_one(src: Path) -> Path:
        obj = OBJ_DIR / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(CUDA_SOURCES))) as ex:
        objs = list(ex.map(compile_one, CUDA_SOURCES))
    cmd = [nvcc, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-lcudart", "-Xlinker",
           "-rpath,/usr/local/cuda/lib64"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link of libpsb_kernels.so failed:\n{r.stdout}\n{r.stderr}")
    stamp_file.write_text(stamp)
    return LIB_PATH


def build_binaries(force: bool = False) -> Path:
    """Build the C++ control-plane binaries with make (CMakeLists.txt is equivalent)."""
    makefile = ROOT / "Makefile"
    if not makefile.exists():
        return BIN_DIR
    cmd = ["make", "-C", str(ROOT), "-j8", "binaries"]
    if force:
        cmd.append("-B")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"make binaries failed:\n{r.stdout[-4000:]}\n{r.stderr[-4000:]}")
    return BIN_DIR


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_kernels(force=force, verbose=verbose)
    build_binaries(force=force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB_PATH)
