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
SOURCES = ["api.cu", "elim.cu", "kernels_aux.cu", "kernels_m4rm.cu", "kernels_mma.cu", "kernels_mma2.cu", "kernels_mma4.cu", "kernels_mma5.cu", "kernels_mma6.cu", "noalloc.cu", "mzd_host.cpp"]
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


PERF_LIB = os.path.join(LIBDIR, "libm4ri_b200_perf.so")


def build(force=False, verbose=False, perf=False):
    """perf=True builds libm4ri_b200_perf.so with -DM4RI_B200_PERF_HOOKS: the tensor kernels then
    honour M4RI_B200_DBG (bit 0: no operand expansion, bit 2: no accumulator drain) for the
    isolation experiments of tools/gpu_dbg.sh. The shipped library has no such switch; select the
    perf build with M4RI_B200_LIB=<path>."""
    if perf:
        return _build(PERF_LIB, os.path.join(LIBDIR, "perf"), ["-DM4RI_B200_PERF_HOOKS"], verbose)
    if not force and not _stale():
        return LIB
    return _build(LIB, LIBDIR, [], verbose)


def _build(LIB, OBJDIR, extra, verbose):
    os.makedirs(OBJDIR, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(OBJDIR, os.path.splitext(src)[0] + ".o")
        objs.append(obj)
        cmd = ["nvcc"] + NVCC_FLAGS + extra + ["-c", "-o", obj, os.path.join(CSRC, src)]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"==== {src}\n{out}")
        failed |= p.returncode != 0
    with open(os.path.join(OBJDIR, "build.log"), "w") as f:
        f.write("\n".join(log))
    if failed or verbose:
        sys.stderr.write("\n".join(log))
    if failed:
        raise RuntimeError("nvcc failed, see m4ri-rust_b200/lib/build.log")
    subprocess.check_call(["nvcc", "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return LIB


def build_variant(name, defs, source="kernels_mma6.cu"):
    """A/B builds of one kernel file: lib/exp/<name>.so = the current objects of the library with
    `source` recompiled under the given -D definitions (tools/mma6_explore.sh)."""
    build()
    d = os.path.join(LIBDIR, "exp")
    os.makedirs(d, exist_ok=True)
    obj = os.path.join(d, name + ".o")
    subprocess.check_call(["nvcc"] + [f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")] + ["-D" + x for x in defs] +
                          ["-c", "-o", obj, os.path.join(CSRC, source)])
    objs = [os.path.join(LIBDIR, os.path.splitext(src)[0] + ".o") for src in SOURCES if src != source] + [obj]
    out = os.path.join(d, name + ".so")
    subprocess.check_call(["nvcc", "-shared", "-o", out] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return out


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


CPP_TEST = os.path.join(LIBDIR, "test_friendly")


def build_cpp_tests(force=False):
    """The reference's unit tests replayed through the C++ mirror of its operator API
    (include/m4ri_b200.hpp, tests/cpp/test_friendly.cpp), linked against the library."""
    src = os.path.join(HERE, "..", "tests", "cpp", "test_friendly.cpp")
    inc = os.path.join(HERE, "..", "include")
    deps = [src, os.path.join(inc, "m4ri_b200.hpp"), os.path.join(inc, "m4ri_b200.h"), LIB]
    if not force and os.path.exists(CPP_TEST) and all(os.path.getmtime(CPP_TEST) >= os.path.getmtime(d) for d in deps):
        return CPP_TEST
    for extra in ([], ["-DM4RI_RUST_M4RM_MUL"], ["-DM4RI_RUST_NAIVE_MUL"]):
        out = CPP_TEST + {0: "", 1: "_m4rm", 2: "_naive"}[[[], ["-DM4RI_RUST_M4RM_MUL"], ["-DM4RI_RUST_NAIVE_MUL"]].index(extra)]
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I", inc] + extra + [src, "-o", out,
                               "-L", LIBDIR, "-lm4ri_b200", "-Wl,-rpath,$ORIGIN"])
    return CPP_TEST


if __name__ == "__main__":
    if "--perf" in sys.argv:
        print(build(perf=True, verbose="--verbose" in sys.argv))
        sys.exit(0)
    build_microbench(force="--force" in sys.argv)
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
    build_cpp_tests(force="--force" in sys.argv)
