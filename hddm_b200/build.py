"""Build libhddm_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m hddm_b200.build [--force] [--verbose]

The library is self-contained (static cudart): it needs no PyTorch and exports exactly the
C ABI declared in include/hddm_b200.h.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
# HDDM_B200_LIB: load (and build) another file name -- kernel variants built with HDDM_B200_NVCC_EXTRA
# (space-separated extra nvcc flags, e.g. -DHDDM_LIKE_THREADS=128) can be measured side by side on the GPU box.
LIB = os.environ.get("HDDM_B200_LIB") or os.path.join(HERE, "libhddm_b200.so")
SOURCES = ["abi.cu", "abi_rl.cu", "abi_ppc.cu", "abi_mcmc.cu", "rl_kernels.cu", "ppc_kernels.cu", "sim_kernels.cu", "wfpt_kernels.cu", "cdfdif_kernels.cu", "gen_kernels.cu", "series_table.cu", "sort_kernels.cu"]
HEADERS = ["kernels.cuh", "abi_runtime.cuh", "wfpt_device.cuh", "wfpt_fast.cuh", "wfpt_grid.cuh", "wfpt_lean.cuh", os.path.join(ROOT, "include", "hddm_b200.h")]

NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "--threads", "0"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: hddm_b200 has no prebuilt or CPU fallback")


def up_to_date():
    if not os.path.exists(LIB):
        return False
    deps = [os.path.join(CSRC, s) for s in SOURCES] + [h if os.path.isabs(h) else os.path.join(CSRC, h)
                                                       for h in HEADERS]
    return os.path.getmtime(LIB) >= max(os.path.getmtime(d) for d in deps)


def build_library(force=False, verbose=False):
    if up_to_date() and not force:
        return LIB
    cmd = [_nvcc()] + NVCC_FLAGS + os.environ.get("HDDM_B200_NVCC_EXTRA", "").split() + \
          (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libhddm_b200.so")
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
