"""In-tree build of libavdiff_b200.so (sm_100a only) with nvcc.

    python -m multimodal_diffusion_b200.build [--force] [--verbose] [--debug-hooks]

nvcc cross-compiles without a GPU, so this runs in the CPU-only build container; the resulting
.so is git-ignored but travels with the working tree to the GPU box.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
ROOT = PKG.parent
LIB = PKG / "libavdiff_b200.so"
OBJ_DIR = ROOT / "build" / "obj"
SOURCES = ["avd_elementwise.cu", "avd_gemm.cu", "avd_gemm_ps.cu", "avd_gemm_chain.cu", "avd_gemm_ps2.cu", "avd_attn.cu", "avd_attn_fa.cu", "avd_stream.cu", "avd_codec.cu",
           "avd_vae.cu", "avd_api.cu"]
HEADERS = [CSRC / "avd_internal.h", CSRC / "avd_ptx.cuh", CSRC / "avd_gemm_common.cuh", CSRC / "avd_gemm_ps_epilogue.inc", CSRC / "avd_attn_common.cuh", ROOT / "include" / "avdiff_b200.h"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _stamp() -> str:
    h = hashlib.sha256()
    for p in [CSRC / s for s in SOURCES] + HEADERS + [Path(__file__)]:
        h.update(p.read_bytes())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False, debug_hooks: bool = False) -> Path:
    """debug_hooks: compile the in-kernel phase trace (avd_debug_trace) in; the default build has no debug hooks."""
    stamp_file = OBJ_DIR / "stamp.txt"
    stamp = _stamp() + ("+debug-hooks" if debug_hooks else "")
    if not force and LIB.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return LIB
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src: str) -> Path:
        obj = OBJ_DIR / (Path(src).stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *(["-DAVD_DEBUG_HOOKS"] if debug_hooks else []), "-c", str(CSRC / src), "-o", str(obj)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        (OBJ_DIR / (Path(src).stem + ".ptxas.log")).write_text(res.stderr)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            print(res.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-o", str(LIB), *[str(o) for o in objs], "-gencode", "arch=compute_100a,code=sm_100a"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    stamp_file.write_text(stamp)
    return LIB


if __name__ == "__main__":
    out = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, debug_hooks="--debug-hooks" in sys.argv)
    print(out)
