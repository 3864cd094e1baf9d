"""Build libdefl_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdefl_b200.so")
# the same sources with the retired engines compiled in (v1 group per stream, v2 lane-decoder pipeline): only the
# cross-check tests load it (tests/conftest.py gpu_lib_legacy); the product library does not carry them
LIB_LEGACY = os.path.join(HERE, "libdefl_b200_legacy.so")
SOURCES = ["defl_b200.cu"]
DEPS = sorted(f for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".inc"))) + \
       ["../../include/defl/infl.h", "../../include/defl/infl_batch.h", "../../include/defl/common.h", "../../include/defl/infl_wrap.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "-shared",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libdefl_b200.so cannot be built (there is no CPU fallback)")


def stale(lib=None):
    lib = lib or LIB
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    for d in DEPS:
        p = os.path.join(CSRC, d)
        if os.path.exists(p) and os.path.getmtime(p) > t:
            return True
    return False


def build(force=False, verbose=False, legacy=False):
    lib = LIB_LEGACY if legacy else LIB
    if not force and not stale(lib):
        return lib
    extra = os.environ.get("DEFL_NVCC_EXTRA", "").split()          # experiments only (e.g. -DDEFL_WARP_CTAS_PER_SM=10)
    if legacy:
        extra.append("-DDEFL_LEGACY_ENGINES")
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", lib] + [os.path.join(CSRC, s) for s in SOURCES]
    out = subprocess.run(cmd, capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stderr)
    return lib


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, legacy="--legacy" in sys.argv))
