"""Build ``yoho_b200/lib/libyoho_b200.so`` (sm_100a only) and the test-only C oracle.

``python -m yoho_b200.build`` or ``__graft_entry__.build()``.  The shared library is plain CUDA
behind a C ABI (include/yoho_b200.h): no torch / pybind types, statically linked cudart, so it can
be bound from ctypes, cgo-style FFI or an XLA custom call alike.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "yoho_b200" / "csrc"
LIB_DIR = ROOT / "yoho_b200" / "lib"
LIB = LIB_DIR / "libyoho_b200.so"
OBJ_DIR = ROOT / "build" / "obj"

SOURCES = ["api.cu", "frontend.cu", "elementwise.cu", "encoder_fp32.cu", "gemm_sm100.cu", "attention_sm100.cu", "attention3_sm100.cu", "encoder.cu", "yoho_ops.cu", "yoho_model.cu", "crosskv.cu", "decoder_ops.cu", "decoder.cu", "xla_ffi.cc"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
] + os.environ.get("YOHO_EXTRA_NVCC_FLAGS", "").split()   # e.g. -DYOHO_ATT_TIMING for the phase-timing debug build


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cc")) + list(CSRC.glob("*.cuh")) + [ROOT / "include" / "yoho_b200.h"]
    stamp = LIB_DIR / ".build_digest"
    digest = _digest(deps)
    if not force and LIB.exists() and stamp.exists() and stamp.read_text() == digest:
        return LIB
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: str) -> Path:
        obj = OBJ_DIR / (src + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(CSRC / src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as pool:
        objs = list(pool.map(compile_one, SOURCES))
    link = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB), *map(str, objs),
            "-cudart", "static", "-Xlinker", "--exclude-libs=ALL", "-ldl"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(digest)
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
