"""Build recipe for libm4ri_b200.so (explicit nvcc, sm_100a only, in-tree output).

    python m4ri-rust_b200/build.py [--force] [--verbose]

The library is plain C++/CUDA behind a C ABI (include/m4ri_b200.h); it does not
link torch. cudart is linked statically, so the .so loads on a box without a GPU
(the product entry points then die loudly on first use).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libm4ri_b200.so")
SOURCES = ["api.cu", "kernels_aux.cu", "kernels_m4rm.cu", "kernels_mma.cu", "kernels_mma2.cu", "mzd_host.cpp"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3,-Wall",
    "-Xptxas", "-v",
]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f != "microbench.cu"]
    deps.append(os.path.join(HERE, "..", "include", "m4ri_b200.h"))
    deps.append(os.path.abspath(__file__))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, os.path.splitext(src)[0] + ".o")
        objs.append(obj)
        cmd = ["nvcc"] + NVCC_FLAGS + ["-c", "-o", obj, os.path.join(CSRC, src)]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"==== {src}\n{out}")
        failed |= p.returncode != 0
    with open(os.path.join(LIBDIR, "build.log"), "w") as f:
        f.write("\n".join(log))
    if failed or verbose:
        sys.stderr.write("\n".join(log))
    if failed:
        raise RuntimeError("nvcc failed, see m4ri-rust_b200/lib/build.log")
    subprocess.check_call(["nvcc", "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return LIB


MICROBENCH = os.path.join(LIBDIR, "microbench")


def build_microbench(force=False):
    """Standalone peak microbenchmarks (LOP3, shared-memory loads, IMMA, HBM copy)."""
    src = os.path.join(CSRC, "microbench.cu")
    if not force and os.path.exists(MICROBENCH) and os.path.getmtime(MICROBENCH) >= os.path.getmtime(src):
        return MICROBENCH
    os.makedirs(LIBDIR, exist_ok=True)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
                           "-std=c++17", "-o", MICROBENCH, src])
    return MICROBENCH


if __name__ == "__main__":
    build_microbench(force="--force" in sys.argv)
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
