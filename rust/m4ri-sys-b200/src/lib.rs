//! `extern "C"` surface of `libm4ri_b200.so` for the multiplication path of m4ri-rust and its
//! immediate consumers (concat/stack, echelon form, inverse, solve) — every `m4ri-sys` symbol
//! that `m4ri-rust/src/friendly/*.rs` references is exported by the one library, so no second
//! allocator is ever linked in (mixing this library with a system libm4ri is unsupported:
//! `mzd_free` dies on a descriptor it did not allocate).
//!
//! The names, argument order and the 64-byte matrix descriptor are the ones `m4ri-sys`
//! declares for the M4RI C library (`include/m4ri_b200.h` cites the `m4ri-sys` line next to
//! every prototype), so `m4ri-rust`'s `BinMatrix`/`BinVector` operators run unchanged on top
//! of this crate: `&a * &b` still expands to one `mzd_mul(NULL, a, b, 0)` call.
//!
//! SOURCE ONLY — never compiled in the build image (no `cargo`/`rustc` there).
#![allow(non_camel_case_types, non_snake_case)]

use libc::{c_char, c_double, c_int, c_uint, c_ulonglong, c_void, size_t};

pub type Rci = c_int;
pub type Wi = c_int;
pub type Word = u64;

/// One block of matrix words (pinned host memory in this implementation).
#[repr(C)]
pub struct MzdBlock {
    pub size: size_t,
    pub begin: *mut Word,
    pub end: *mut Word,
}

/// The 64-byte descriptor. Offsets: nrows 0, ncols 4, width 8, rowstride 12, offset_vector 16,
/// row_offset 20, flags 24, blockrows_log 25, padding 26..40, high_bitmask 40, blocks 48, rows 56.
#[repr(C)]
pub struct Mzd {
    pub nrows: Rci,
    pub ncols: Rci,
    pub width: Wi,
    pub rowstride: Wi,
    pub offset_vector: Wi,
    pub row_offset: Wi,
    pub flags: u8,
    pub blockrows_log: u8,
    pub padding: [u8; 14],
    pub high_bitmask: Word,
    pub blocks: *mut MzdBlock,
    pub rows: *mut *mut Word,
}

/// Permutation as a list of transpositions (`values[i] = j`: swap i and j, for i = 0, 1, ... in
/// turn). `m4ri-sys/src/mzp.rs:9-12` keeps it opaque; this is M4RI's published two-field layout.
#[repr(C)]
pub struct Mzp {
    pub values: *mut Rci,
    pub length: Rci,
}

/// Opaque device-resident matrix (same packing, even row stride, lives in HBM).
#[repr(C)]
pub struct DMat {
    _private: [u8; 0],
}

pub const ALGO_AUTO: c_int = 0;
pub const ALGO_M4RM: c_int = 1;
pub const ALGO_MMA: c_int = 2;
pub const ALGO_SIMPLE: c_int = 3;

extern "C" {
    // ---- allocator set: one allocator owns every matrix the path touches ----
    pub fn mzd_init(r: Rci, c: Rci) -> *mut Mzd;
    pub fn mzd_free(a: *mut Mzd);
    pub fn mzd_init_window(m: *mut Mzd, lowr: Rci, lowc: Rci, highr: Rci, highc: Rci) -> *mut Mzd;
    pub fn mzd_copy(dst: *mut Mzd, a: *const Mzd) -> *mut Mzd;
    pub fn mzd_equal(a: *const Mzd, b: *const Mzd) -> c_int;
    pub fn mzd_set_ui(a: *mut Mzd, value: c_uint);
    pub fn mzd_randomize(a: *mut Mzd);
    pub fn mzd_transpose(dst: *mut Mzd, a: *const Mzd) -> *mut Mzd;
    pub fn mzd_add(c: *mut Mzd, a: *const Mzd, b: *const Mzd) -> *mut Mzd;
    pub fn mzd_sub(c: *mut Mzd, a: *const Mzd, b: *const Mzd) -> *mut Mzd;
    /// [A | B] (`BinMatrix::augmented`, binary_matrix.rs:271-276) and A over B (`stacked`, 279-283).
    pub fn mzd_concat(c: *mut Mzd, a: *const Mzd, b: *const Mzd) -> *mut Mzd;
    pub fn mzd_stack(c: *mut Mzd, a: *const Mzd, b: *const Mzd) -> *mut Mzd;

    // ---- products (C == NULL: the library allocates, the caller frees with mzd_free) ----
    pub fn mzd_mul(c: *mut Mzd, a: *const Mzd, b: *const Mzd, cutoff: c_int) -> *mut Mzd;
    pub fn mzd_addmul(c: *mut Mzd, a: *const Mzd, b: *const Mzd, cutoff: c_int) -> *mut Mzd;
    pub fn mzd_mul_m4rm(c: *mut Mzd, a: *const Mzd, b: *const Mzd, k: c_int) -> *mut Mzd;
    pub fn mzd_addmul_m4rm(c: *mut Mzd, a: *const Mzd, b: *const Mzd, k: c_int) -> *mut Mzd;
    pub fn _mzd_mul_m4rm(c: *mut Mzd, a: *const Mzd, b: *const Mzd, k: c_int, clear: c_int) -> *mut Mzd;
    pub fn mzd_mul_naive(c: *mut Mzd, a: *const Mzd, b: *const Mzd) -> *mut Mzd;
    pub fn mzd_addmul_naive(c: *mut Mzd, a: *const Mzd, b: *const Mzd) -> *mut Mzd;
    pub fn _mzd_mul_naive(c: *mut Mzd, a: *const Mzd, bt: *const Mzd, clear: c_int) -> *mut Mzd;
    pub fn _mzd_mul_va(c: *mut Mzd, v: *const Mzd, a: *const Mzd, clear: c_int) -> *mut Mzd;
    /// y = A*x with x (ncols x 1) and y (nrows x 1): the shape `BinMatrix::mul_slice` builds.
    pub fn mzd_mul_v(y: *mut Mzd, a: *const Mzd, x: *const Mzd) -> *mut Mzd;

    // ---- consumers of the product: reduced row echelon form on the GPU (csrc/elim.cu) ----
    /// Reduced row echelon form in place; returns the rank (`full` is accepted and ignored:
    /// the result is always the fully reduced form, which is also a valid echelon form).
    pub fn mzd_echelonize(a: *mut Mzd, full: c_int) -> Rci;
    pub fn mzd_echelonize_pluq(a: *mut Mzd, full: c_int) -> Rci;
    pub fn mzd_echelonize_m4ri(a: *mut Mzd, full: c_int, k: c_int) -> Rci;
    /// NULL when `src` is singular.
    pub fn mzd_inv_m4ri(dst: *mut Mzd, src: *const Mzd, k: c_int) -> *mut Mzd;
    /// Solves A X = B, X in the first ncols(A) rows of B; 0 on success, -1 when inconsistent.
    pub fn mzd_solve_left(a: *mut Mzd, b: *mut Mzd, cutoff: c_int, inconsistency_check: c_int) -> c_int;
    /// `mzd.rs:218`; NULL when `a` is singular. `identity` is accepted and not needed.
    pub fn mzd_invert_naive(inv: *mut Mzd, a: *const Mzd, identity: *const Mzd) -> *mut Mzd;
    /// `brilliantrussian.rs:199`; returns the rank.
    pub fn mzd_top_echelonize_m4ri(m: *mut Mzd, k: c_int) -> Rci;
    /// `mp.rs:6,8`: M4RI's OpenMP products — the same GPU calls as `mzd_mul` / `mzd_addmul`.
    pub fn mzd_mul_mp(c: *mut Mzd, a: *const Mzd, b: *const Mzd, cutoff: c_int) -> *mut Mzd;
    pub fn mzd_addmul_mp(c: *mut Mzd, a: *const Mzd, b: *const Mzd, cutoff: c_int) -> *mut Mzd;

    // ---- row / column utilities (`mzd.rs:130-240`): host word work ----
    pub fn mzd_row_swap(matrix: *mut Mzd, rowa: Rci, rowb: Rci);
    pub fn mzd_copy_row(b: *mut Mzd, i: Rci, a: *const Mzd, j: Rci);
    pub fn mzd_col_swap(matrix: *mut Mzd, cola: Rci, colb: Rci);
    pub fn mzd_submatrix(s: *mut Mzd, m: *const Mzd, lowr: Rci, lowc: Rci, highr: Rci, highc: Rci) -> *mut Mzd;
    pub fn mzd_is_zero(a: *const Mzd) -> c_int;
    pub fn mzd_row_clear_offset(m: *mut Mzd, row: Rci, coloffset: Rci);

    // ---- PLE / PLUQ decomposition and permutations (m4ri-sys ple.rs, mzp.rs; csrc/elim.cu) ----
    /// A = P L E in place, returns the rank; P has length nrows, Q length ncols (`ple.rs:65`).
    pub fn mzd_ple(a: *mut Mzd, p: *mut Mzp, q: *mut Mzp, cutoff: c_int) -> Rci;
    /// A = P L U Q in place, returns the rank (`ple.rs:35`).
    pub fn mzd_pluq(a: *mut Mzd, p: *mut Mzp, q: *mut Mzp, cutoff: c_int) -> Rci;
    /// Solves (P L U Q) X = B from the packed factors and the rank `mzd_pluq` returned (`solve.rs:53`).
    pub fn mzd_pluq_solve_left(a: *const Mzd, rank: Rci, p: *const Mzp, q: *const Mzp, b: *mut Mzd,
                               cutoff: c_int, inconsistency_check: c_int) -> c_int;
    pub fn mzp_init(length: Rci) -> *mut Mzp;
    pub fn mzp_free(p: *mut Mzp);
    /// (`mzp.rs:26` declares no return value; the window is the returned pointer, as in M4RI.)
    pub fn mzp_init_window(p: *mut Mzp, begin: Rci, end: Rci) -> *mut Mzp;
    pub fn mzp_free_window(condemned: *mut Mzp);
    /// The spelling `mzp.rs:29` declares; same function.
    pub fn Mzp_free_window(condemned: *mut Mzp);
    pub fn mzp_copy(p: *mut Mzp, q: *const Mzp) -> *mut Mzp;
    pub fn mzp_set_ui(p: *mut Mzp, value: c_uint);
    pub fn mzd_apply_p_left(a: *mut Mzd, p: *const Mzp);
    pub fn mzd_apply_p_left_trans(a: *mut Mzd, p: *const Mzp);
    pub fn mzd_apply_p_right(a: *mut Mzd, p: *const Mzp);
    pub fn mzd_apply_p_right_trans(a: *mut Mzd, p: *const Mzp);
    pub fn mzp_print(p: *const Mzp);

    // ---- engine control ----
    pub fn m4ri_b200_set_algo(algo: c_int);
    pub fn m4ri_b200_get_algo() -> c_int;
    pub fn m4ri_b200_set_strassen_levels(levels: c_int);
    pub fn m4ri_b200_set_devices(n: c_int);
    pub fn m4ri_b200_get_devices() -> c_int;
    pub fn m4ri_b200_host_cache_trim();
    pub fn m4ri_b200_device_count() -> c_int;
    pub fn m4ri_b200_device_info(name: *mut c_char, name_len: c_int, sm_count: *mut c_int,
                                 cc_major: *mut c_int, cc_minor: *mut c_int) -> c_int;
    pub fn m4ri_b200_set_stream(stream: *mut c_void);
    pub fn m4ri_b200_use_own_stream();
    pub fn m4ri_b200_synchronize();
    pub fn m4ri_b200_launch_count() -> c_ulonglong;
    pub fn m4ri_b200_version() -> *const c_char;
    pub fn m4ri_b200_pipeline_plan(m: c_int, k: c_int, n: c_int, clear: c_int,
                                   row_chunks: *mut c_int, col_blocks: *mut c_int);

    // ---- device-resident matrices: chained products without the PCIe round trip ----
    pub fn m4ri_b200_dmat_init(nrows: Rci, ncols: Rci) -> *mut DMat;
    pub fn m4ri_b200_dmat_wrap(dptr: *mut c_void, nrows: Rci, ncols: Rci, rowstride: Wi) -> *mut DMat;
    pub fn m4ri_b200_dmat_free(m: *mut DMat);
    pub fn m4ri_b200_dmat_nrows(m: *const DMat) -> Rci;
    pub fn m4ri_b200_dmat_ncols(m: *const DMat) -> Rci;
    pub fn m4ri_b200_dmat_rowstride(m: *const DMat) -> Wi;
    pub fn m4ri_b200_dmat_ptr(m: *const DMat) -> *mut c_void;
    pub fn m4ri_b200_upload(d: *mut DMat, h: *const Mzd);
    pub fn m4ri_b200_download(h: *mut Mzd, d: *const DMat);
    pub fn m4ri_b200_dmul(c: *mut DMat, a: *const DMat, b: *const DMat, clear: c_int, algo: c_int);
    pub fn m4ri_b200_dmul_abt(c: *mut DMat, a: *const DMat, bt: *const DMat, clear: c_int, algo: c_int);
    pub fn m4ri_b200_dmul_v(y: *mut DMat, a: *const DMat, x: *const DMat, clear: c_int);
    pub fn m4ri_b200_dmul_va(c: *mut DMat, v: *const DMat, a: *const DMat, clear: c_int);
    /// Y (b x nrows(A)) = rows of X (b x ncols(A)) each multiplied by A: b mat-vecs per pass over A.
    pub fn m4ri_b200_dmul_v_batched(y: *mut DMat, a: *const DMat, x: *const DMat, clear: c_int);
    /// Reduced row echelon form of a device matrix in place; returns the rank.
    pub fn m4ri_b200_drref(w: *mut DMat, pivot_cols: c_int, pivot_columns: *mut c_int, max_pivots: c_int) -> c_int;
    /// PLE (`pluq == 0`) or PLUQ of a device matrix in place; `p` (nrows) and `q` (ncols) are host arrays.
    pub fn m4ri_b200_dple(w: *mut DMat, p: *mut Rci, q: *mut Rci, pluq: c_int) -> c_int;
    pub fn m4ri_b200_dtranspose(dst: *mut DMat, a: *const DMat);
    pub fn m4ri_b200_dadd(c: *mut DMat, a: *const DMat, b: *const DMat);
    pub fn m4ri_b200_dcopy(dst: *mut DMat, a: *const DMat);
    pub fn m4ri_b200_dchecksum(m: *const DMat) -> u64;
    pub fn m4ri_b200_timing_ms() -> c_double;
}

/// Bit (row, col) of a host matrix: bit `col % 64` of word `col / 64` of `rows[row]`, LSB first.
///
/// # Safety
/// `m` must come from this library and `row`/`col` must be in range.
#[inline]
pub unsafe fn mzd_read_bit(m: *const Mzd, row: Rci, col: Rci) -> c_int {
    let words = *(*m).rows.offset(row as isize);
    ((*words.offset((col / 64) as isize) >> (col % 64)) & 1) as c_int
}

/// Sets bit (row, col) of a host matrix to `value & 1` (m4ri-sys/src/mzd.rs `mzd_write_bit`).
///
/// # Safety
/// `m` must come from this library and `row`/`col` must be in range.
#[inline]
pub unsafe fn mzd_write_bit(m: *mut Mzd, row: Rci, col: Rci, value: c_int) {
    let words = *(*m).rows.offset(row as isize);
    let w = words.offset((col / 64) as isize);
    let bit = 1u64 << (col % 64);
    *w = (*w & !bit) | (((value & 1) as u64) << (col % 64));
}

#[cfg(test)]
mod tests {
    use super::*;
    use std::mem::size_of;
    use std::ptr;

    #[test]
    fn descriptor_is_64_bytes() {
        assert_eq!(size_of::<Mzd>(), 64);
    }

    #[test]
    fn identity_times_identity() {
        unsafe {
            let a = mzd_init(8, 8);
            mzd_set_ui(a, 1);
            let c = mzd_mul(ptr::null_mut(), a, a, 0);
            assert!(!c.is_null());
            assert_eq!(mzd_equal(a, c), 1);
            mzd_free(c);
            mzd_free(a);
        }
    }
}
