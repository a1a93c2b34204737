"""The Rust FFI crate (rust/m4ri-sys-b200) is shipped as source — there is no cargo in the
image — so its extern block is checked here against the two things it must agree with: the C
header (same names, same number of arguments) and the shared library's export table."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_RS = os.path.join(ROOT, "rust", "m4ri-sys-b200", "src", "lib.rs")
HEADER = os.path.join(ROOT, "include", "m4ri_b200.h")


def _args(arglist):
    arglist = arglist.strip()
    if arglist in ("", "void"):
        return 0
    return len([a for a in arglist.split(",") if a.strip()])


def rust_decls():
    src = open(LIB_RS).read()
    block = src[src.index('extern "C" {'):]
    block = block[: block.index("\n}\n")]
    return {m.group(1): _args(m.group(2))
            for m in re.finditer(r"pub fn (\w+)\s*\(([^)]*)\)", block, re.S)}


def header_decls():
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return {m.group(1): _args(m.group(2))
            for m in re.finditer(r"\b(_?mzd_\w+|[mM]zp_\w+|m4ri_b200_\w+)\s*\(([^;{]*?)\)\s*;", src, re.S)}


def test_every_rust_declaration_is_in_the_header_with_the_same_arity():
    rust, hdr = rust_decls(), header_decls()
    assert len(rust) >= 40
    for name, n in rust.items():
        assert name in hdr, f"{name} is not declared in include/m4ri_b200.h"
        assert hdr[name] == n, f"{name}: {n} arguments in lib.rs, {hdr[name]} in the header"
    # the whole reference-facing multiplication path is bound
    for name in ("mzd_mul", "mzd_addmul", "mzd_mul_m4rm", "mzd_addmul_m4rm", "_mzd_mul_m4rm", "mzd_mul_naive",
                 "mzd_addmul_naive", "_mzd_mul_naive", "_mzd_mul_va", "mzd_mul_v", "mzd_transpose", "mzd_add",
                 "mzd_init", "mzd_free", "mzd_copy", "mzd_equal", "mzd_randomize", "mzd_set_ui"):
        assert name in rust
    # and the decomposition surface of m4ri-sys/src/ple.rs and mzp.rs
    for name in ("mzd_ple", "mzd_pluq", "mzp_init", "mzp_free", "mzp_init_window", "Mzp_free_window", "mzp_copy",
                 "mzp_set_ui", "mzd_apply_p_left", "mzd_apply_p_left_trans", "mzd_apply_p_right",
                 "mzd_apply_p_right_trans", "mzp_print"):
        assert name in rust


def test_every_rust_declaration_is_exported(m4):
    so = os.path.join(ROOT, "m4ri-rust_b200", "lib", "libm4ri_b200.so")
    out = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    missing = [n for n in rust_decls() if n not in exported]
    assert not missing, missing


def test_struct_fields_follow_the_header_order():
    src = open(LIB_RS).read()
    body = src[src.index("pub struct Mzd {"):]
    body = body[: body.index("}")]
    fields = re.findall(r"pub (\w+):", body)
    assert fields == ["nrows", "ncols", "width", "rowstride", "offset_vector", "row_offset", "flags",
                      "blockrows_log", "padding", "high_bitmask", "blocks", "rows"]


# Every C symbol m4ri-rust's own sources call (grep of m4ri-rust/src for mzd_* on the reference
# checkout: friendly/binary_matrix.rs and friendly/binary_vector.rs; mzd_read_bit/mzd_write_bit are
# inline Rust in m4ri-sys, mirrored in lib.rs). A user switching crates must find all of them.
M4RI_RUST_USES = ("mzd_init", "mzd_free", "mzd_add", "mzd_equal", "mzd_mul_naive", "mzd_transpose", "mzd_stack",
                  "mzd_solve_left", "mzd_set_ui", "mzd_randomize", "mzd_mul_m4rm", "mzd_mul", "mzd_inv_m4ri",
                  "mzd_echelonize", "mzd_copy", "mzd_concat")


def test_every_symbol_m4ri_rust_calls_is_bound_and_exported(m4):
    rust = rust_decls()
    so = os.path.join(ROOT, "m4ri-rust_b200", "lib", "libm4ri_b200.so")
    out = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    for name in M4RI_RUST_USES:
        assert name in rust, f"{name} missing from rust/m4ri-sys-b200/src/lib.rs"
        assert name in exported, f"{name} not exported by libm4ri_b200.so"
    src = open(LIB_RS).read()
    assert "pub unsafe fn mzd_read_bit" in src and "pub unsafe fn mzd_write_bit" in src


def test_no_second_allocator_is_linked():
    for f in ("Cargo.toml", "build.rs"):
        text = open(os.path.join(ROOT, "rust", "m4ri-sys-b200", f)).read()
        assert "full-m4ri" not in text and "dylib=m4ri\"" not in text
