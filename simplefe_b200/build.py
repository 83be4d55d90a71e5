"""build.py -- compiles libsfegpu.so in-tree with nvcc for sm_100a (no torch involved).

The library is a plain C-ABI shared object (include/sfe_gpu.h); the static CUDA runtime is linked
in, so it has no dependency beyond the driver.  `python -m simplefe_b200.build` rebuilds it.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libsfegpu.so")

CU_SOURCES = ["capi.cu", "chain.cu", "blkconv_kernel.cu", "polyphase_kernel.cu", "util_kernels.cu", "toeplitz_mma_kernel.cu", "fused_bank_kernel.cu", "ring.cu", "bpsk_tx_kernel.cu", "bpsk_tx.cu", "feed.cu", "box.cu", "no_alloc.cu"]
CPP_SOURCES = ["schedule.cpp"]
HEADERS = ["kernels.h", "schedule.h", "sfe_common.cuh", "sfe_fft.cuh", "sfe_fft_plan.h", "../../include/sfe_gpu.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libsfegpu.so cannot be built (there is no CPU fallback)")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = _nvcc()
    # tuning builds only: every K3 flavour of the variant sweeps (scripts/bank_sweep.sh); objects kept apart
    sweep = os.environ.get("SFE_BANK_SWEEP") == "1"
    extra = os.environ.get("SFE_NVCC_EXTRA", "").split()     # tuning builds: extra -D flags (A/B runs on the GPU box)
    objdir = os.path.join(PKG, "build_sweep" if sweep else "build")
    os.makedirs(objdir, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    objs = []
    procs = []
    for src in CU_SOURCES + CPP_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src + ".o")
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            if src.endswith(".cu"):
                cmd = [nvcc] + NVCC_FLAGS + (["-DSFE_BANK_SWEEP"] if sweep else []) + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
            else:
                # host-only schedule code: as-written IEEE single precision, no FMA contraction
                cmd = [nvcc, "-O2", "-std=c++17", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
                       "-Xcompiler", "-fvisibility=hidden",
                       # the schedule walk is one tight float loop: pin its alignment, or its speed depends on what else is linked
                       "-Xcompiler", "-falign-functions=64", "-Xcompiler", "-falign-loops=32", "-x", "c++", "-c", s, "-o", o]
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- nvcc {src}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-o", LIB] + objs
        subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
