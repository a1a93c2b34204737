/*
 * m4ri_b200.h — C-ABI of libm4ri_b200.so, the B200-native drop-in for the GF(2)
 * matrix-multiplication path of thomwiggers/m4ri-rust.
 *
 * Every `mzd_*` symbol below has the name, argument order and meaning of the
 * `extern "C"` prototype that m4ri-sys declares for the M4RI C library; the
 * reference declaration each one replaces is cited as `file:line` relative to
 * the reference checkout. All pointers in the `mzd_*` functions are HOST
 * pointers: host memory is the source of truth on entry and is valid on return
 * (m4ri-rust writes matrix words through `rows[]` without telling the library,
 * m4ri-rust/src/friendly/binary_matrix.rs:137-159).
 *
 * The `m4ri_b200_*` symbols are additive: device-resident operands, stream
 * control and algorithm selection for benchmarking and chained products.
 *
 * Error model (M4RI-compatible, `m4ri_die`): dimension mismatches and CUDA
 * failures print to stderr and abort(). There is no CPU fallback for any
 * multiplication entry point: without a usable sm_100 GPU they die loudly.
 */
#ifndef M4RI_B200_H
#define M4RI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- scalar types: m4ri-sys/src/misc.rs:5-26 -------------------------------- */
typedef int rci_t;      /* Rci  = c_int  (misc.rs:5)  */
typedef int wi_t;       /* Wi   = c_int  (misc.rs:14) */
typedef int BIT;        /* BIT  = c_int  (misc.rs:8)  */
typedef uint64_t word;  /* Word = u64    (misc.rs:11) */

#define m4ri_radix 64                          /* misc.rs:18 */
#define m4ri_one ((word)1)                     /* misc.rs:22 */
#define m4ri_ffff ((word)0xffffffffffffffffULL) /* misc.rs:26 */

/* ---- matrix descriptor: m4ri-sys/src/mzd.rs:16-79 --------------------------- */
typedef struct mzd_block_t { /* mzd.rs:16-21 */
  size_t size;               /* bytes in this block */
  word *begin;
  word *end;
} mzd_block_t;

/* Exactly 64 bytes (asserted by the reference test mzd.rs:385). Bit (r,c) is
 * (rows[r][c/64] >> (c%64)) & 1  (mzd.rs:263-269). */
typedef struct mzd_t {
  rci_t nrows;             /*  0 */
  rci_t ncols;             /*  4 */
  wi_t width;              /*  8  ceil(ncols/64) */
  wi_t rowstride;          /* 12  (width<3 || width even) ? width : width+1   (mzd.rs:35-37) */
  wi_t offset_vector;      /* 16  rows[0] == blocks[0].begin + offset_vector  (mzd.rs:41-44) */
  wi_t row_offset;         /* 20 */
  uint8_t flags;           /* 24  mzd.rs:82-94 */
  uint8_t blockrows_log;   /* 25 */
  uint8_t padding[14];     /* 26  mzd.rs:67-71 */
  word high_bitmask;       /* 40  valid bits of word width-1 */
  mzd_block_t *blocks;     /* 48 */
  word **rows;             /* 56 */
} mzd_t;

#define mzd_flag_nonzero_excess 0x2       /* mzd.rs:82 */
#define mzd_flag_windowed_zerooffset 0x4  /* mzd.rs:85 */
#define mzd_flag_windowed_zeroexcess 0x8  /* mzd.rs:88 */
#define mzd_flag_windowed_ownsblocks 0x10 /* mzd.rs:91 */
#define mzd_flag_multiple_blocks 0x20     /* mzd.rs:94 */

/* ---- allocator set (one allocator owns every mzd_t the path touches) -------- */
mzd_t *mzd_init(rci_t r, rci_t c);                       /* mzd.rs:98  */
void mzd_free(mzd_t *A);                                 /* mzd.rs:102 */
mzd_t *mzd_init_window(mzd_t *M, rci_t lowr, rci_t lowc, /* mzd.rs:121-127 (lowc % 64 == 0) */
                       rci_t highr, rci_t highc);
mzd_t *mzd_copy(mzd_t *DST, const mzd_t *A);             /* mzd.rs:192 */
int mzd_equal(const mzd_t *A, const mzd_t *B);           /* mzd.rs:187 */
void mzd_set_ui(mzd_t *A, unsigned int value);           /* mzd.rs:198 */
void mzd_randomize(mzd_t *A);                            /* mzd.rs:184 */
mzd_t *mzd_transpose(mzd_t *DST, const mzd_t *A);        /* mzd.rs:148 */
mzd_t *mzd_add(mzd_t *C, const mzd_t *A, const mzd_t *B);/* mzd.rs:223 (C may alias A) */
mzd_t *mzd_sub(mzd_t *C, const mzd_t *A, const mzd_t *B);/* mzd.rs:230 (alias of add)  */
/* [A B] and [A; B] (BinMatrix::augmented / stacked, binary_matrix.rs:221-244): host copies, so that
 * an unmodified m4ri-rust links against this library alone. */
/* Row / column utilities m4ri-sys binds next to them: host word work on the caller's matrix. */
void mzd_row_swap(mzd_t *M, rci_t rowa, rci_t rowb);                 /* mzd.rs:130 */
void mzd_copy_row(mzd_t *B, rci_t i, const mzd_t *A, rci_t j);       /* mzd.rs:141 */
void mzd_col_swap(mzd_t *M, rci_t cola, rci_t colb);                 /* mzd.rs:144 */
mzd_t *mzd_submatrix(mzd_t *S, const mzd_t *M, rci_t lowr, rci_t lowc, rci_t highr, rci_t highc); /* mzd.rs:205 (copy; any lowc) */
int mzd_is_zero(const mzd_t *A);                                     /* mzd.rs:233 (declared there without its int result) */
void mzd_row_clear_offset(mzd_t *M, rci_t row, rci_t coloffset);     /* mzd.rs:240 */
mzd_t *mzd_concat(mzd_t *C, const mzd_t *A, const mzd_t *B); /* mzd.rs:195 */
mzd_t *mzd_stack(mzd_t *C, const mzd_t *A, const mzd_t *B);  /* mzd.rs:201 */

/* ---- the hot path: every one of these runs on the GPU ----------------------- */
/* C = A*B. C == NULL: allocate. cutoff is a hint (the result does not depend on it). */
mzd_t *mzd_mul(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff);      /* strassen.rs:18 */
/* C ^= A*B. */
mzd_t *mzd_addmul(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff);   /* strassen.rs:31 */
/* M4RI's multi-core forms of the two: the same GPU calls. */
mzd_t *mzd_mul_mp(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff);   /* mp.rs:6 */
mzd_t *mzd_addmul_mp(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff); /* mp.rs:8 */
mzd_t *mzd_mul_m4rm(mzd_t *C, const mzd_t *A, const mzd_t *B, int k);      /* brilliantrussian.rs:216 */
mzd_t *mzd_addmul_m4rm(mzd_t *C, const mzd_t *A, const mzd_t *B, int k);   /* brilliantrussian.rs:224 */
/* upstream-M4RI-shaped, named by the north star; not declared by m4ri-sys */
mzd_t *_mzd_mul_m4rm(mzd_t *C, const mzd_t *A, const mzd_t *B, int k, int clear);
mzd_t *mzd_mul_naive(mzd_t *C, const mzd_t *A, const mzd_t *B);            /* mzd.rs:152 */
mzd_t *mzd_addmul_naive(mzd_t *C, const mzd_t *A, const mzd_t *B);         /* mzd.rs:173 */
/* C (^)= A * Bt^T with B pre-transposed; clear != 0 zeroes C first. */
mzd_t *_mzd_mul_naive(mzd_t *C, const mzd_t *A, const mzd_t *Bt, int clear); /* mzd.rs:163-168 */
/* C (^)= v*A, v has few rows (normally 1). */
mzd_t *_mzd_mul_va(mzd_t *C, const mzd_t *v, const mzd_t *A, int clear);   /* mzd.rs:181 */
/* y = A*x, x is ncols(A) x 1, y is nrows(A) x 1 (the shape BinMatrix::mul_slice
 * produces, binary_matrix.rs:426-427). Named by the north star. */
mzd_t *mzd_mul_v(mzd_t *y, const mzd_t *A, const mzd_t *x);

/* ---- consumers of the product: blocked Gauss-Jordan on the GPU whose trailing
 * update is one accumulating product per column block (csrc/elim.cu) ---------------------------- */
/* (Reduced) row echelon form in place, returns the rank. The REDUCED form is returned for either
 * value of `full` (it is an upper triangular form, and unlike that one it is unique).
 * DIFFERS FROM M4RI: with full = 0 M4RI stops at an upper triangular form and leaves the entries
 * above the pivots as they fall; a caller that relied on exactly those entries sees different (still
 * valid echelon) rows here. m4ri-rust only ever reads the rank or passes full = 1
 * (binary_matrix.rs:247-261), so it cannot tell. */
rci_t mzd_echelonize(mzd_t *A, int full);               /* echelonform.rs:16 (BinMatrix::echelonize/rank, binary_matrix.rs:247-261) */
rci_t mzd_echelonize_pluq(mzd_t *A, int full);          /* echelonform.rs:27 */
rci_t mzd_echelonize_m4ri(mzd_t *A, int full, int k);   /* echelonform.rs:36 */
rci_t mzd_top_echelonize_m4ri(mzd_t *A, int k);         /* brilliantrussian.rs:199 (declared there without its rank result) */
/* dst = src^-1 (dst NULL: allocate); NULL when src is singular. */
mzd_t *mzd_inv_m4ri(mzd_t *dst, const mzd_t *src, int k); /* brilliantrussian.rs:208 (BinMatrix::inverted, binary_matrix.rs:264-268) */
mzd_t *mzd_invert_naive(mzd_t *inv, const mzd_t *A, const mzd_t *identity); /* mzd.rs:218 (identity is not needed) */
/* Solves A X = B, X stored in the first ncols(A) rows of B (free variables zero, other rows of B
 * zeroed); 0 = solved, -1 = inconsistent (B untouched). A is left as it was.
 * DIFFERS FROM M4RI: M4RI factors in place and returns with A overwritten by its PLUQ factors;
 * here A is only read. m4ri-rust clones A before the call and drops the clone
 * (binary_matrix.rs:582-586), so both behaviours look the same to it. */
int mzd_solve_left(mzd_t *A, mzd_t *B, int cutoff, int inconsistency_check); /* solve.rs:23 (solve_left, binary_matrix.rs:582-586) */

/* ---- PLE / PLUQ decomposition and permutations (m4ri-sys/src/ple.rs, mzp.rs) ----
 * mzp.rs:9-12 keeps Mzp opaque; the layout is M4RI's published mzp_t. P->values[i] = j means
 * "swap i and j", applied for i = 0, 1, ... in turn (LAPACK style), values[i] >= i. */
typedef struct mzp_t {
  rci_t *values;
  rci_t length;
} mzp_t;
mzp_t *mzp_init(rci_t length);                             /* mzp.rs:18  identity permutation */
void mzp_free(mzp_t *P);                                   /* mzp.rs:21 */
mzp_t *mzp_init_window(mzp_t *P, rci_t begin, rci_t end);  /* mzp.rs:26  shares P's values */
void mzp_free_window(mzp_t *condemned);                    /* mzp.rs:29 (declared there as `Mzp_free_window`, exported too) */
void Mzp_free_window(mzp_t *condemned);
mzp_t *mzp_copy(mzp_t *P, const mzp_t *Q);                 /* mzp.rs:33  P NULL: allocate */
void mzp_set_ui(mzp_t *P, unsigned int value);             /* mzp.rs:36  identity (value is ignored, as in M4RI) */
void mzp_print(const mzp_t *P);                            /* mzp.rs:51 */
/* Row / column permutations of a host matrix, done on the GPU (one gather pass):
 * left: rows i <-> P[i] for i = 0 .. len-1 in turn; left_trans: the same swaps in reverse order;
 * right_trans: columns i <-> P[i] for i = 0 .. len-1 in turn; right: in reverse order. */
void mzd_apply_p_left(mzd_t *A, const mzp_t *P);           /* mzp.rs:39 */
void mzd_apply_p_left_trans(mzd_t *A, const mzp_t *P);     /* mzp.rs:42 */
void mzd_apply_p_right(mzd_t *A, const mzp_t *P);          /* mzp.rs:45 */
void mzd_apply_p_right_trans(mzd_t *A, const mzp_t *P);    /* mzp.rs:48 */
/* A = P L E in place, returns the rank r (ple.rs:65). P (length nrows) and Q (length ncols) are
 * overwritten. Afterwards row i < r holds L[i][0..i) in columns [0, i) and row i of the echelon
 * form E from its pivot column Q[i] on (E[i][Q[i]] = 1); rows i >= r hold L[i][0..r) in columns
 * [0, r). L has a unit diagonal (not stored). Q[0] < Q[1] < ... < Q[r-1] is the column rank
 * profile, Q[i] = i beyond. mzd_apply_p_left(A_input, P) == L * E. cutoff is accepted and ignored. */
rci_t mzd_ple(mzd_t *A, mzp_t *P, mzp_t *Q, int cutoff);
/* A = P L U Q in place, returns the rank (ple.rs:35): as mzd_ple, with U = E after the column
 * swaps i <-> Q[i], i = 0 .. r-1 in turn (mzd_apply_p_right_trans), upper triangular r x ncols. */
rci_t mzd_pluq(mzd_t *A, mzp_t *P, mzp_t *Q, int cutoff);
/* Solves (P L U Q) X = B from the packed factors mzd_pluq left in A and its rank (solve.rs:53): X
 * overwrites the first ncols(A) rows of B (free variables zero, other rows zeroed); 0 = solved,
 * -1 = inconsistent (B untouched). B needs at least max(nrows(A), ncols(A)) rows. */
int mzd_pluq_solve_left(const mzd_t *A, rci_t rank, const mzp_t *P, const mzp_t *Q, mzd_t *B, int cutoff,
                        int inconsistency_check);

/* ---- additive API: algorithm choice, device-resident operands --------------- */
enum m4ri_b200_algo {
  M4RI_B200_ALGO_AUTO = 0,  /* library picks (shape based)                        */
  M4RI_B200_ALGO_M4RM = 1,  /* path (a): shared-memory Four-Russians, LOP3 gather  */
  M4RI_B200_ALGO_MMA = 2,   /* path (b): tcgen05 MMA (FP4 default, int8 variants)  */
  M4RI_B200_ALGO_SIMPLE = 3 /* AND+popc row.row kernel (what *_naive maps to)      */
};
/* Force an algorithm for mzd_mul/mzd_addmul (AUTO restores the default). Also
 * settable with the environment variable M4RI_B200_ALGO=auto|m4rm|mma|simple. */
void m4ri_b200_set_algo(int algo);
int m4ri_b200_get_algo(void);
/* Strassen-Winograd levels applied above the leaf kernel on device products:
 * -1 = automatic from `cutoff`/shape, 0 = none. */
void m4ri_b200_set_strassen_levels(int levels);
/* The persistent product kernels normally take every SM. A collective that should overlap a
 * product (the NCCL broadcast of the next k-chunk of B, sharded.py) needs a few SMs of its own:
 * sms > 0 caps the SMs the product kernels launched after this call use (0 = all). */
void m4ri_b200_set_sm_limit(int sms);
/* GPUs one host-ABI product may use (default 1; also M4RI_B200_DEVICES). With n > 1 a large
 * mzd_mul/mzd_addmul/_mzd_mul_naive on host operands is cut into n row-blocks of A and C, one per
 * GPU of this process (devices 0..n-1), each running the copy/compute pipeline on its block over
 * its own PCIe link; the blocks are independent (no collective). get returns the effective count
 * (clamped to the devices present, and to 8). */
void m4ri_b200_set_devices(int n);
int m4ri_b200_get_devices(void);
/* Block plan the host-ABI products use to overlap PCIe copies with the kernels on an
 * m x k x n product: row-chunks of A and column blocks of B (pure host arithmetic, no device
 * needed; M4RI_B200_PIPE_ROWS / M4RI_B200_PIPE_COLS override it). */
void m4ri_b200_pipeline_plan(int m, int k, int n, int clear, int *row_chunks, int *col_blocks);
/* Kernel variant behind M4RI_B200_ALGO_MMA (0 restores the default; also
 * settable with M4RI_B200_MMA=v1|v2cg1|v2|fp4cg1|fp4|fp4ldgcg1|fp4ldg|fp4s256cg1|fp4s256): benchmarking knob. */
enum m4ri_b200_mma_variant {
  M4RI_B200_MMA_DEFAULT = 0,
  M4RI_B200_MMA_V1 = 1,        /* one CTA per 128x128 tile, operands expanded into shared memory */
  M4RI_B200_MMA_V2_SINGLE = 2, /* TMA-staged bits, A in tensor memory, cta_group::1              */
  M4RI_B200_MMA_V2_PAIR = 3,   /* the same on CTA pairs (cta_group::2), 256x256 tiles            */
  M4RI_B200_MMA_FP4_SINGLE = 4,/* bits -> E2M1 nibbles, kind::mxf4 (2x the int8 rate), cta_group::1 */
  M4RI_B200_MMA_FP4_PAIR = 5,  /* the same on CTA pairs, 256x256 tiles                           */
  M4RI_B200_MMA_FP4_LDG_SINGLE = 6, /* FP4, packed bits by direct 256-bit global loads (no TMA)  */
  M4RI_B200_MMA_FP4_LDG_PAIR = 7,   /* the same on CTA pairs (round 1's default)                 */
  M4RI_B200_MMA_FP4_S256_SINGLE = 8,/* FP4, 256-bit stages x 6, 16 expander + 8 epilogue warps   */
  M4RI_B200_MMA_FP4_S256_PAIR = 9   /* the same on CTA pairs (default)                           */
};
void m4ri_b200_set_mma_variant(int variant);

/* Number of usable CUDA devices (0 when there is none; never aborts). */
int m4ri_b200_device_count(void);
/* Name/SM count of the device the calling thread's context uses. */
int m4ri_b200_device_info(char *name, int name_len, int *sm_count, int *cc_major, int *cc_minor);
/* Use `stream` (a cudaStream_t) for all device-API work issued by this thread;
 * NULL selects the library's own stream. */
/* mzd_free keeps blocks of 256 KiB or more (pinned host memory is expensive to allocate) for the
 * next mzd_init of about the same size, up to M4RI_B200_HOST_CACHE_MB (default 4096) in total.
 * trim returns them to the system; stats reports hits, misses and the bytes currently held. */
void m4ri_b200_host_cache_trim(void);
void m4ri_b200_host_cache_stats(unsigned long long *hits, unsigned long long *misses,
                                unsigned long long *bytes_held);
void m4ri_b200_set_stream(void *stream);
/* Back to the calling thread's own non-blocking stream (the default before any set_stream). */
void m4ri_b200_use_own_stream(void);
void m4ri_b200_synchronize(void);
/* Count of kernels the library launched since process start (bench: gpu_launches). */
unsigned long long m4ri_b200_launch_count(void);
const char *m4ri_b200_version(void);

/* A device-resident matrix: same 64-bit-word row-major LSB-first packing as
 * mzd_t, rowstride in words (>= width, even), excess bits always zero. */
typedef struct m4ri_b200_dmat m4ri_b200_dmat_t;
m4ri_b200_dmat_t *m4ri_b200_dmat_init(rci_t nrows, rci_t ncols);
/* Wrap caller-owned device memory (e.g. a torch tensor): not freed by dmat_free. */
m4ri_b200_dmat_t *m4ri_b200_dmat_wrap(void *dptr, rci_t nrows, rci_t ncols, wi_t rowstride);
void m4ri_b200_dmat_free(m4ri_b200_dmat_t *M);
rci_t m4ri_b200_dmat_nrows(const m4ri_b200_dmat_t *M);
rci_t m4ri_b200_dmat_ncols(const m4ri_b200_dmat_t *M);
wi_t m4ri_b200_dmat_rowstride(const m4ri_b200_dmat_t *M);
void *m4ri_b200_dmat_ptr(const m4ri_b200_dmat_t *M);
/* Async on the current stream; host memory should be pinned (mzd_init's is). */
void m4ri_b200_upload(m4ri_b200_dmat_t *D, const mzd_t *H);
void m4ri_b200_download(mzd_t *H, const m4ri_b200_dmat_t *D);
/* C (^)= A*B on device, async on the current stream. algo as enum above. */
void m4ri_b200_dmul(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *B,
                    int clear, int algo);
/* C (^)= A*B on device like m4ri_b200_dmul, for callers that must not allocate. It has no
 * counterpart in the reference (M4RI allocates its temporaries freely; m4ri-sys declares nothing
 * like it). m4ri_b200_dmul takes B^T from the CUDA memory pool; this entry point keeps B^T in a
 * fixed array inside the library's device image instead and calls no cudaMalloc / cudaFree /
 * pool routine at all. It is meant for hosts whose device memory is owned by someone else (a
 * torch program passing its tensors, a service that filled the GPU with its own matrices), where
 * an allocation inside the library fails or takes memory the owner budgeted, and for tight
 * loops of small products that cannot afford cudaMalloc's device-wide synchronisation.
 *   C, A, B   device pointers to m x n, m x k, k x n bit matrices in the mzd_t packing (64-bit
 *             words, row-major, LSB first, bits beyond ncols zero), 16-byte aligned; C must not
 *             overlap A or B
 *   *_rowstride  words between rows: even, >= ceil(ncols/64)
 *   clear     non-zero: C = A*B; zero: C ^= A*B
 *   stream    a cudaStream_t, or NULL for the calling thread's library stream (the one every
 *             m4ri_b200_d* call uses; a thread's first such call creates it)
 * Limit: n * roundup(ceil(k/64), 4) <= M4RI_B200_NOALLOC_SCRATCH_WORDS, i.e. B up to 8192 x 8192
 * bits; any m. The array (8 MiB) exists on every GPU the process uses and stays until exit. No
 * Strassen level is applied (those need temporaries), so large products run a few percent
 * slower than m4ri_b200_dmul.
 * The first call on each device looks up that device's copy of the array; make that call
 * before any stream capture. All calls on one device share the array: two calls whose kernels
 * can overlap on the GPU (different streams, no dependency between them) race on it. Calls on
 * one stream are ordered and safe; the array is re-initialised at the start of every call.
 * Asynchronous on `stream`. Returns 0, or without launching anything: -1 B beyond the limit,
 * -2 misaligned operand or odd row stride, -3 negative size / NULL operand / stride below the row
 * width (each with a message on stderr). CUDA errors abort like everywhere else in the library. */
#define M4RI_B200_NOALLOC_SCRATCH_WORDS (1u << 20)
int m4ri_b200_mul_noalloc(word *C, wi_t c_rowstride, const word *A, wi_t a_rowstride, const word *B,
                          wi_t b_rowstride, rci_t m, rci_t k, rci_t n, int clear, void *stream);
/* C (^)= A*Bt^T on device (both operands k-contiguous). */
void m4ri_b200_dmul_abt(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A,
                        const m4ri_b200_dmat_t *Bt, int clear, int algo);
/* y (^)= A*x, x packed as a 1 x ncols(A) row, y packed as a 1 x nrows(A) row
 * (packed words in, packed words out: SURVEY 8f.2). */
void m4ri_b200_dmul_v(m4ri_b200_dmat_t *y, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *x,
                      int clear);
/* Batched mat-vec (BASELINE configs[4] "batched mzd_mul_v"): row q of Y (1 x nrows(A), packed) (^)=
 * A * x_q for the b packed vectors x_q = rows of X (b x ncols(A)); Y is b x nrows(A). One pass over A
 * per 8 vectors (HBM-bound); more than 32 vectors go to the tensor path as Y = X * A^T. The
 * reference's mat-vec surface is mul_slice / Mul<&BinVector> (binary_matrix.rs:416-431,528-550),
 * one vector per call. */
void m4ri_b200_dmul_v_batched(m4ri_b200_dmat_t *Y, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *X,
                              int clear);
/* C (^)= v*A, v is r x nrows(A) with small r. */
void m4ri_b200_dmul_va(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *v, const m4ri_b200_dmat_t *A,
                       int clear);
void m4ri_b200_dtranspose(m4ri_b200_dmat_t *DST, const m4ri_b200_dmat_t *A);
/* A^T written to ndst (1..8) destinations at once; A and the destinations may be peer-GPU
 * memory mapped into this process (fused broadcast + transpose of the sharded product). */
void m4ri_b200_dtranspose_scatter(m4ri_b200_dmat_t *const *DST, int ndst, const m4ri_b200_dmat_t *A);
void m4ri_b200_dadd(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *B);
void m4ri_b200_dcopy(m4ri_b200_dmat_t *DST, const m4ri_b200_dmat_t *A);
/* Fill with splitmix64(seed) words, row by row, excess bits masked: the same
 * stream of words the oracle's generator produces (oracle/gf2_oracle.c). */
void m4ri_b200_dfill_random(m4ri_b200_dmat_t *M, uint64_t seed);
/* Same, starting at output number `first_index` of the stream (row-block shards
 * of one big matrix generate exactly their own rows). */
void m4ri_b200_dfill_random_at(m4ri_b200_dmat_t *M, uint64_t seed, uint64_t first_index);
/* Per-kernel device timing of the dominant product kernel (CUDA events on the
 * launching stream). reset() enables recording and clears it; ms() ends the recording, synchronises and
 * returns the summed duration of the recorded launches; count() their number;
 * name() the kernel they belong to. */
void m4ri_b200_timing_reset(void);
double m4ri_b200_timing_ms(void);
int m4ri_b200_timing_count(void);
const char *m4ri_b200_timing_name(void);
/* Sum of 2*m*n*k over the recorded launches (Strassen leaves are smaller than the job). */
double m4ri_b200_timing_ops(void);
/* Diagnostics for the FP4 variant: C = A*Bt^T with the fp32 accumulators of the tensor
 * cores copied to acc (device pointer, nrows(C) x ncols(C) floats, row-major): every value
 * must be exactly (number of common set bits)/4. cg = 1 or 2 CTAs per MMA. */
void m4ri_b200_dmul_abt_fp4_debug(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A,
                                  const m4ri_b200_dmat_t *Bt, void *acc, int cg);
/* Reduced row echelon form of a device matrix in place; pivots are searched in columns
 * [0, pivot_cols) only (0 = all), the other columns are carried along. Returns the rank and the
 * first max_pivots pivot columns (increasing). Synchronises the stream once per column block. */
int m4ri_b200_drref(m4ri_b200_dmat_t *W, int pivot_cols, int *pivot_columns, int max_pivots);
/* PLE (pluq = 0) or PLUQ (pluq != 0) decomposition of a device matrix in place, storage as
 * mzd_ple / mzd_pluq; P (nrows entries) and Q (ncols entries) are HOST arrays. Returns the rank. */
int m4ri_b200_dple(m4ri_b200_dmat_t *W, rci_t *P, rci_t *Q, int pluq);
/* 64-bit XOR-fold checksum of all words (position-mixed), computed on device. */
uint64_t m4ri_b200_dchecksum(const m4ri_b200_dmat_t *M);
/* 1 iff every valid word is zero (OR-reduction on device). */
int m4ri_b200_dis_zero(const m4ri_b200_dmat_t *M);
/* Flip bit (r,c) on device (fault injection for the harness). */
void m4ri_b200_dflip_bit(m4ri_b200_dmat_t *M, rci_t r, rci_t c);
/* 1 when `algo` (enum above) is built into this library for the current device. */
int m4ri_b200_algo_available(int algo);
/* Freivalds over GF(2) with 64 packed random vectors: returns 1 iff
 * A*(B*X) == C*X for X = random(ncols(B) x 64, seed). False accept 2^-64. */
int m4ri_b200_dfreivalds(const m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A,
                         const m4ri_b200_dmat_t *B, uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif /* M4RI_B200_H */
