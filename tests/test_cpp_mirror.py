"""The C++ mirror of m4ri-rust's operator API (include/m4ri_b200.hpp) replaying the
reference's own unit tests (tests/cpp/test_friendly.cpp), one binary per `mul_impl!`
strategy (binary_matrix.rs:53-95). On a GPU box all checks must pass; on a CPU box
the binaries must refuse loudly (no CPU fallback)."""
import importlib.util
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _binaries():
    spec = importlib.util.spec_from_file_location("m4b_build", os.path.join(ROOT, "m4ri-rust_b200", "build.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    b.build()
    base = b.build_cpp_tests()
    return [base, base + "_m4rm", base + "_naive"]


def test_cpp_mirror_builds_and_refuses_without_gpu(m4):
    bins = _binaries()
    assert all(os.path.exists(p) for p in bins)
    if m4.device_count() > 0:
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([bins[0]], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "no CUDA device" in r.stderr


@pytest.mark.gpu
def test_reference_unit_tests_through_cpp_mirror(gpu):
    for p in _binaries():
        r = subprocess.run([p], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, f"{os.path.basename(p)}:\n{r.stdout}\n{r.stderr}"
        assert "0 failed checks" in r.stdout
        launches = int(r.stdout.strip().splitlines()[-1].split(",")[-1].split()[0])
        assert launches > 0
