"""In-tree build of the CUDA library (``panvc3_b200/libpanvc3_b200.so``) for sm_100a with nvcc.

nvcc cross-compiles without a GPU; the built .so is git-ignored but travels to the GPU box with the repo snapshot."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpanvc3_b200.so")
SOURCES = ["pv3_api.cu", "pv3_pipeline.cu", "pv3_rewrite.cu", "pv3_rewrite_thread.cu", "pv3_rewrite_word.cu", "pv3_rewrite_seg.cu", "pv3_realign.cu", "pv3_realign_thread.cu", "pv3_splice.cu", "pv3_pack.cu", "pv3_msa_build.cu", "pv3_annotate.cu", "pv3_mapq.cu", "pv3_index_file.cc"]
HEADERS = ["pv3_device.cuh", "pv3_kernels.h", "pv3_runs.cuh", "pv3_k1w_core.h", "pv3_segment_core.h", "pv3_index_build.h", "pv3_ctx.h", os.path.join("..", "..", "include", "panvc3_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the panvc3_b200 CUDA library cannot be built (there is no CPU fallback)")


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-o", LIB, *[os.path.join(CSRC, s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building libpanvc3_b200.so")
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
