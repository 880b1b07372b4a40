"""Build libchub.so (CUDA kernels + C ABI) in-tree for sm_100a with nvcc."""
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
SOURCES = ["chub_kernels.cu", "chub_engine.cu", "chub_noalloc.cu", "chub_hostpack.cpp"]
HEADERS = ["chub_kernels.cuh", "chub_device.cuh", "chub_hostpack.h", "chub_fastmath.h", os.path.join("..", "..", "include", "chub.h"),
           os.path.join("..", "..", "include", "chub_rng.h")]
OUT = os.path.join(CSRC, "libchub.so")
# -fmad=false: the kernels reproduce the reference's separately rounded multiply/add; fused ops are
# written explicitly as fma() where the design wants them (see csrc/chub_device.cuh).
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC,-fno-builtin-pow"]


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    return all(os.path.getmtime(os.path.join(CSRC, f)) <= t for f in SOURCES + HEADERS)


OUT_CHECKED = os.path.join(CSRC, "libchub_checked.so")   # the same sources with -DCHUB_CHECKS (include/chub.h chub_checks)


def build(force=False, verbose=False, checked=True):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if force or not up_to_date():
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + SOURCES + ["-o", OUT]
        subprocess.check_call(cmd, cwd=CSRC)
    if checked and (force or not os.path.exists(OUT_CHECKED) or os.path.getmtime(OUT_CHECKED) < os.path.getmtime(OUT)):
        subprocess.check_call([nvcc] + NVCC_FLAGS + ["-DCHUB_CHECKS"] + SOURCES + ["-o", OUT_CHECKED], cwd=CSRC)
    return OUT


if __name__ == "__main__":
    print(build(force=True, verbose=True))
