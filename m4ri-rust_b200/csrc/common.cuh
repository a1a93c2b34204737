// common.cuh — shared declarations of libm4ri_b200 (internal, not installed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>

#include "../../include/m4ri_b200.h"

namespace m4b {

// M4RI's error model (m4ri_die): print and abort. SURVEY 8b "Errors".
[[noreturn]] void die(const char *fmt, ...);

#define M4B_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess)                                                               \
      ::m4b::die("m4ri_b200: CUDA error %s (%d) at %s:%d: %s\n", cudaGetErrorName(_e),  \
                 (int)_e, __FILE__, __LINE__, cudaGetErrorString(_e));                   \
  } while (0)

// Called after every kernel launch: counts it and surfaces launch errors.
extern std::atomic<unsigned long long> g_launches;
#define M4B_LAUNCHED()                 \
  do {                                 \
    ::m4b::g_launches.fetch_add(1);    \
    M4B_CUDA(cudaGetLastError());      \
  } while (0)

// Makes sure a usable device exists (dies loudly otherwise) and returns the
// stream the calling thread issues work on.
cudaStream_t current_stream();
void require_device();
int sm_count();
bool has_tcgen05();  // compute capability 10.x

// Function attributes (opt-in dynamic shared memory) are per device: true the first time the
// calling site is reached with this device current (a process may drive several GPUs).
static inline bool first_use_on_this_device(std::atomic<unsigned long long> &seen) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
  const unsigned long long bit = 1ull << dev;
  return !(seen.fetch_or(bit) & bit);
}

// Event pair around the dominant product kernel (api.cu; no-ops until enabled).
// `ops` = the 2*m*n*k bit-operations of this launch (summed for the roofline).
// mzd_init without the O(size) zeroing for results whose every valid word is about to be
// downloaded over (mzd_host.cpp)
mzd_t *mzd_init_impl(rci_t r, rci_t c, int zero);
void timing_begin(const char *kernel_name, cudaStream_t s, double ops);
void timing_end(cudaStream_t s);

static inline int width_of(int ncols) { return (ncols + 63) / 64; }
static inline word high_mask(int ncols) {
  int r = ncols % 64;
  return r ? (((word)1 << r) - 1) : ~(word)0;
}
// Device row stride (64-bit words): even and >= 2 so every row is 16-byte aligned.
static inline int dev_rowstride(int ncols) {
  int w = width_of(ncols);
  w = (w + 1) & ~1;
  return w < 2 ? 2 : w;
}

}  // namespace m4b

// Device-resident matrix (opaque at the ABI).
struct m4ri_b200_dmat {
  word *d;        // first word of row 0
  rci_t nrows, ncols;
  wi_t width;     // ceil(ncols/64)
  wi_t rowstride; // words between rows (even)
  void *base;     // allocation to free (NULL when wrapped / window)
};

namespace m4b {
typedef m4ri_b200_dmat DMat;

// ---- api.cu: device matrices and the device product, shared with elim.cu -------
DMat *dmat_new(rci_t nrows, rci_t ncols, cudaStream_t s);   // zeroed, stream-ordered pool
void dmat_delete(DMat *M, cudaStream_t s);
DMat dwin(const DMat *M, int r0, int c0, int nr, int nc);   // window (c0 % 64 == 0), shares the words
void upload(DMat *D, const mzd_t *H, cudaStream_t s);
void download(mzd_t *H, const DMat *D, cudaStream_t s);
// C (^)= A*B on device (algo: enum m4ri_b200_algo; cutoff: Strassen hint, 0 = tuned)
void dmul(DMat *C, const DMat *A, const DMat *B, int clear, int algo, cudaStream_t s, int cutoff = 0);

// ---- kernels_aux.cu ----------------------------------------------------------
void k_zero(DMat *M, cudaStream_t s);
void k_mask_excess(DMat *M, cudaStream_t s);
// last[i*stride] = (last[i*stride] & keep) | (host_col[i] & ~keep) for i < nrows (window results)
void k_merge_excess(word *last, int stride, const word *host_col, int nrows, word keep, cudaStream_t s);
void k_add(DMat *C, const DMat *A, const DMat *B, cudaStream_t s);
void k_copy(DMat *D, const DMat *A, cudaStream_t s);
void k_transpose(DMat *D, const DMat *A, cudaStream_t s);
void k_transpose_scatter(DMat *const *D, int ndst, const DMat *A, cudaStream_t s);
void k_fill_random(DMat *M, uint64_t seed, uint64_t first_index, cudaStream_t s);
uint64_t k_checksum(const DMat *M, cudaStream_t s);
bool k_is_zero(const DMat *M, cudaStream_t s);
void k_flip_bit(DMat *M, int r, int c, cudaStream_t s);
// y (^)= A*x ; x, y packed rows
void k_mul_v(DMat *y, const DMat *A, const DMat *x, int clear, cudaStream_t s);
// Y (^)= X*A^T for a few packed vectors (rows of X): one pass over A per 8 vectors
void k_mul_v_batched(DMat *Y, const DMat *A, const DMat *X, int clear, cudaStream_t s);
// C (^)= v*A
void k_mul_va(DMat *C, const DMat *v, const DMat *A, int clear, cudaStream_t s);
// C (^)= A*Bt^T by AND + popcount parity
void k_mul_abt_simple(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s);

// ---- kernels_m4rm.cu ---------------------------------------------------------
// C (^)= A*B, Four-Russians tables in shared memory
void k_mul_m4rm(DMat *C, const DMat *A, const DMat *B, int clear, cudaStream_t s);

// ---- kernels_mma.cu ----------------------------------------------------------
// C (^)= A*Bt^T, bits -> int8 in shared memory / TMEM, tcgen05 MMA, parity
// dispatcher (api.cu): picks the variant set by m4ri_b200_set_mma_variant
void k_mul_abt_mma(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s);
// v1: one CTA per 128x128 tile, both operands expanded into shared memory (SS form)
void k_mul_abt_mma1(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s);
// ---- kernels_mma2.cu: TMA-staged bits, A in tensor memory (TS form), cg CTAs per MMA
void k_mul_abt_mma2(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, cudaStream_t s);
// ---- kernels_mma4.cu: bits -> E2M1 nibbles, tcgen05 kind::mxf4 with unit scale factors
void k_mul_abt_mma4(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, float *raw_acc,
                    cudaStream_t s);
// ---- kernels_mma5.cu: the same contraction, packed bits by direct global loads
void k_mul_abt_mma5(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, cudaStream_t s);
// ---- kernels_mma6.cu: the same contraction, 256-bit stages x 6, 16 expander + 8 epilogue warps (default)
void k_mul_abt_mma6_group(int count, DMat *const *C1, DMat *const *C2, const DMat *const *A, const DMat *const *Bt,
                          int cg, cudaStream_t s);
void k_quad_sums(DMat *const out[5], const DMat q[4], const int pairs[5][2], cudaStream_t s);
void k_mul_abt_mma6(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, cudaStream_t s, DMat *C2 = nullptr,
                    int clear2 = 0);
bool mma_shape_ok(const DMat *C, const DMat *A, const DMat *Bt);
bool mma_available();

}  // namespace m4b
