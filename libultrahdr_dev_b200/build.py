"""In-tree build of the CUDA library (sm_100a only): libultrahdr_dev_b200/libuhdr_b200.so."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libuhdr_b200.so")
SOURCES = ["capi.cu", "generate.cu", "apply.cu", "apply_fast.cu", "apply_pipe.cu", "kat.cu", "jpeg_idct.cu", "jpeg_huff.cu", "jpeg_enc.cu", "tables.cu", "tonemap.cu", "convert_yuv.cu", "sdr_pack.cu"]
HEADERS = ["uhdr_math.cuh", "uhdr_kernels.cuh", "uhdr_io.cuh", "uhdr_convert.cuh", "apply_q_common.cuh", os.path.join("..", "..", "include", "uhdr_b200.h"),
           os.path.join("..", "..", "include", "uhdr_b200_test.h")]
# -fmad=false: plain a*b+c keeps the reference's two roundings (x86-64 baseline has no FMA
# contraction); fmaf() is explicit where contraction is wanted. See csrc/uhdr_math.cuh.
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC", "-shared",
]


def _nvcc():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: cannot build libuhdr_b200.so")


def up_to_date():
    if not os.path.exists(LIB):
        return False
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return all(os.path.getmtime(d) <= t for d in deps)


def build(force=False, verbose=False):
    """Compile every kernel for sm_100a into one shared library. Returns its path."""
    if not force and up_to_date():
        return LIB
    extra = os.environ.get("UHDR_NVCC_EXTRA", "").split()  # experiments only, e.g. -DUHDR_APPLY_MUFU_CHANNELS=2
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + [os.path.join(CSRC, f) for f in SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    import sys

    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
