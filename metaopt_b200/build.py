"""Build recipe for libob200.so (nvcc, sm_100a only).  Run: python -m metaopt_b200.build"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
SOURCES = ["ob200_ctx.cu", "ob200_sample.cu", "ob200_fit.cu", "ob200_score.cu", "ob200_bound.cu"]
LIB = os.path.join(HERE, "libob200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC",
         "-I", os.path.join(ROOT, "include"), "-I", CSRC]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "ob200.h"),
                                                                 os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, defines=(), out=None):
    """Compile every .cu for sm_100a (cross-compiles without a GPU) and link libob200.so in-tree.

    `defines`/`out` build a tuning variant (e.g. defines=["SC_BK=32"], out="libob200_bk32.so")."""
    lib = os.path.join(HERE, out) if out else LIB
    if not force and not out and not needs_build():
        return lib
    tag = ("_" + os.path.splitext(out)[0]) if out else ""
    objs = []
    for src in SOURCES:
        obj = os.path.join(CSRC, src.replace(".cu", tag + ".o"))
        cmd = [NVCC] + FLAGS + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + \
            ["-c", os.path.join(CSRC, src), "-o", obj]
        subprocess.run(cmd, check=True)
        objs.append(obj)
    subprocess.run([NVCC, "-shared", "-o", lib] + objs + ["-lcudart"], check=True)
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
