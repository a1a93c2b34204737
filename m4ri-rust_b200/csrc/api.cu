// api.cu — context, device-resident matrices and the multiplication entry
// points of the C-ABI (include/m4ri_b200.h). Each mzd_* product does what the
// reference's FFI call does (allocate C when NULL, validate shapes, die on
// mismatch — SURVEY 8b) with the arithmetic on the GPU:
//   host mzd_t --H2D--> device mirrors --kernels--> device C --D2H--> host mzd_t
// synchronously, because m4ri-rust reads the result words right after the call
// returns (m4ri-rust/src/friendly/binary_matrix.rs:459-472).
#include <stdarg.h>
#include <string.h>

#include <condition_variable>
#include <functional>
#include <chrono>
#include <mutex>
#include <thread>
#include <utility>
#include <vector>

#include "common.cuh"

namespace m4b {

std::atomic<unsigned long long> g_launches{0};

void die(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vfprintf(stderr, fmt, ap);
  va_end(ap);
  fflush(stderr);
  abort();
}

// ---- per-kernel timing (bench.py roofline) ------------------------------------------
static std::mutex g_tmu;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_tev;
static bool g_timing_on = false;
static std::thread::id g_timing_thread;  // only launches of the thread that called timing_reset are recorded
static const char *g_tname = "";
static double g_tops = 0;
void timing_begin(const char *kernel_name, cudaStream_t s, double ops) {
  if (!g_timing_on) return;
  std::lock_guard<std::mutex> g(g_tmu);
  if (std::this_thread::get_id() != g_timing_thread || g_tev.size() >= 4096) return;
  cudaEvent_t a, b;
  M4B_CUDA(cudaEventCreate(&a));
  M4B_CUDA(cudaEventCreate(&b));
  M4B_CUDA(cudaEventRecord(a, s));
  g_tev.push_back({a, b});
  g_tname = kernel_name;
  g_tops += ops;
}
void timing_end(cudaStream_t s) {
  if (!g_timing_on) return;
  std::lock_guard<std::mutex> g(g_tmu);
  if (std::this_thread::get_id() != g_timing_thread || g_tev.empty()) return;
  M4B_CUDA(cudaEventRecord(g_tev.back().second, s));
}

// ---- context -------------------------------------------------------------------
static std::once_flag g_once;
static int g_ndev = 0, g_sms = 0, g_cc_major = 0, g_cc_minor = 0;
static char g_name[256] = "";
static std::atomic<int> g_algo{-1};
static std::atomic<int> g_strassen{-1};
static std::atomic<int> g_mma_variant{0};

static void probe() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    (void)cudaGetLastError();
    n = 0;
  }
  g_ndev = n;
  if (n > 0) {
    int dev = 0;
    cudaDeviceProp p;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaGetDeviceProperties(&p, dev) == cudaSuccess) {
      g_sms = p.multiProcessorCount;
      g_cc_major = p.major;
      g_cc_minor = p.minor;
      snprintf(g_name, sizeof(g_name), "%s", p.name);
      (void)cudaGetLastError();
    } else {
      (void)cudaGetLastError();
      g_ndev = 0;
    }
  }
  const char *e = getenv("M4RI_B200_ALGO");
  int a = M4RI_B200_ALGO_AUTO;
  if (e) {
    if (!strcmp(e, "m4rm")) a = M4RI_B200_ALGO_M4RM;
    else if (!strcmp(e, "mma")) a = M4RI_B200_ALGO_MMA;
    else if (!strcmp(e, "simple")) a = M4RI_B200_ALGO_SIMPLE;
  }
  g_algo.store(a);
  const char *sl = getenv("M4RI_B200_STRASSEN");
  if (sl && *sl) g_strassen.store(atoi(sl));
}

void require_device() {
  std::call_once(g_once, probe);
  if (g_ndev <= 0)
    die("m4ri_b200: no CUDA device is available; this library has no CPU fallback for the "
        "GF(2) multiplication path\n");
  if (g_cc_major < 10)
    die("m4ri_b200: device '%s' is sm_%d%d; the kernels are built for sm_100a (B200) only\n",
        g_name, g_cc_major, g_cc_minor);
}

static std::atomic<int> g_sm_limit{0};
int sm_count() {
  require_device();
  const int lim = g_sm_limit.load();
  return (lim > 0 && lim < g_sms) ? lim : g_sms;
}
bool has_tcgen05() {
  require_device();
  return g_cc_major == 10;
}

static thread_local cudaStream_t t_user_stream = nullptr;
static thread_local bool t_use_user_stream = false;
static thread_local cudaStream_t t_own_stream = nullptr;

cudaStream_t current_stream() {
  require_device();
  if (t_use_user_stream) return t_user_stream;
  if (!t_own_stream) M4B_CUDA(cudaStreamCreateWithFlags(&t_own_stream, cudaStreamNonBlocking));
  return t_own_stream;
}

// ---- path (b) variant ----
static int mma_variant() {
  int v = g_mma_variant.load();
  if (v == 0) {
    const char *e = getenv("M4RI_B200_MMA");
    v = !e ? M4RI_B200_MMA_FP4_S256_PAIR
        : !strcmp(e, "v1") ? M4RI_B200_MMA_V1
        : !strcmp(e, "v2cg1") ? M4RI_B200_MMA_V2_SINGLE
        : !strcmp(e, "fp4cg1") ? M4RI_B200_MMA_FP4_SINGLE
        : !strcmp(e, "v2") ? M4RI_B200_MMA_V2_PAIR
        : !strcmp(e, "fp4ldg") ? M4RI_B200_MMA_FP4_LDG_PAIR
        : !strcmp(e, "fp4ldgcg1") ? M4RI_B200_MMA_FP4_LDG_SINGLE
        : !strcmp(e, "fp4s256cg1") ? M4RI_B200_MMA_FP4_S256_SINGLE
        : !strcmp(e, "fp4") ? M4RI_B200_MMA_FP4_PAIR : M4RI_B200_MMA_FP4_S256_PAIR;
  }
  return v;
}
void k_mul_abt_mma(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s) {
  const int v = mma_variant();
  if (v == M4RI_B200_MMA_V1) k_mul_abt_mma1(C, A, Bt, clear, s);
  else if (v == M4RI_B200_MMA_FP4_S256_SINGLE || v == M4RI_B200_MMA_FP4_S256_PAIR)
    k_mul_abt_mma6(C, A, Bt, clear, v == M4RI_B200_MMA_FP4_S256_SINGLE ? 1 : 2, s);
  else if (v == M4RI_B200_MMA_FP4_LDG_SINGLE || v == M4RI_B200_MMA_FP4_LDG_PAIR)
    k_mul_abt_mma5(C, A, Bt, clear, v == M4RI_B200_MMA_FP4_LDG_SINGLE ? 1 : 2, s);
  else if (v == M4RI_B200_MMA_FP4_SINGLE || v == M4RI_B200_MMA_FP4_PAIR)
    k_mul_abt_mma4(C, A, Bt, clear, v == M4RI_B200_MMA_FP4_SINGLE ? 1 : 2, nullptr, s);
  else k_mul_abt_mma2(C, A, Bt, clear, v == M4RI_B200_MMA_V2_SINGLE ? 1 : 2, s);
}

// ---- device matrices -------------------------------------------------------------
// Keep freed blocks of the CURRENT device's stream-ordered pool cached instead of returning them
// to the driver at every synchronize (a fresh 128 MiB cudaMallocAsync costs ~0.5 ms, more with
// peer mappings). Per device: a process may select its GPU after the library was first touched
// (one rank per GPU under torchrun), so this cannot live in the one-time probe().
static void keep_pool_cached() {
  static std::atomic<unsigned long long> done{0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
  const unsigned long long bit = 1ull << dev;
  if (done.load(std::memory_order_relaxed) & bit) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
    unsigned long long keep = ~0ull;
    (void)cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  (void)cudaGetLastError();
  done.fetch_or(bit);
}

DMat *dmat_new(rci_t nrows, rci_t ncols, cudaStream_t s) {
  keep_pool_cached();
  if (nrows < 0 || ncols < 0) die("m4ri_b200: negative matrix dimension\n");
  DMat *M = (DMat *)calloc(1, sizeof(DMat));
  if (!M) die("m4ri_b200: out of memory\n");
  M->nrows = nrows;
  M->ncols = ncols;
  M->width = width_of(ncols);
  M->rowstride = dev_rowstride(ncols);
  size_t bytes = (size_t)(nrows > 0 ? nrows : 1) * M->rowstride * 8;
  M4B_CUDA(cudaMallocAsync(&M->base, bytes, s));
  M4B_CUDA(cudaMemsetAsync(M->base, 0, bytes, s));
  M->d = (word *)M->base;
  return M;
}

void dmat_delete(DMat *M, cudaStream_t s) {
  if (!M) return;
  if (M->base) M4B_CUDA(cudaFreeAsync(M->base, s));
  free(M);
}

// Rows [r0, r1) of H -> the same rows of D. rows[] is the only thing the reference
// guarantees ("multi-block"): maximal runs of evenly strided rows go as one
// 2-D copy each.
// One run of evenly strided rows between host and device. When the rows are back to back on both
// sides (stride == width) it is ONE linear copy: a 2-D copy of 64-byte rows runs at a quarter of
// the PCIe rate (2^20 x 512 bits: 5.6 ms as 2-D, 1.3 ms linear).
static void copy_rows(void *dst, size_t dpitch, const void *src, size_t spitch, size_t wbytes, size_t nrows,
                      cudaMemcpyKind kind, cudaStream_t s) {
  if (dpitch == wbytes && spitch == wbytes) M4B_CUDA(cudaMemcpyAsync(dst, src, wbytes * nrows, kind, s));
  else M4B_CUDA(cudaMemcpy2DAsync(dst, dpitch, src, spitch, wbytes, nrows, kind, s));
}

static void upload_rows(DMat *D, const mzd_t *H, rci_t r0, rci_t r1, cudaStream_t s) {
  if (D->nrows != H->nrows || D->ncols != H->ncols)
    die("m4ri_b200: upload: shape mismatch (%d x %d vs %d x %d)\n", D->nrows, D->ncols, H->nrows,
        H->ncols);
  if (r1 <= r0 || H->width == 0) return;
  rci_t i = r0;
  while (i < r1) {
    rci_t j = i + 1;
    while (j < r1 && H->rows[j] == H->rows[i] + (size_t)(j - i) * H->rowstride) j++;
    copy_rows(D->d + (size_t)i * D->rowstride, (size_t)D->rowstride * 8, H->rows[i], (size_t)H->rowstride * 8,
              (size_t)H->width * 8, j - i, cudaMemcpyHostToDevice, s);
    i = j;
  }
  if (D->ncols % 64) {  // never trust excess bits of a caller's matrix
    DMat W = *D;
    W.d = D->d + (size_t)r0 * D->rowstride;
    W.nrows = r1 - r0;
    W.base = nullptr;
    k_mask_excess(&W, s);
  }
}
void upload(DMat *D, const mzd_t *H, cudaStream_t s) { upload_rows(D, H, 0, H->nrows, s); }

// A result written into a WINDOW whose last word is shared with the parent's other columns
// (ncols % 64 != 0): the bits beyond ncols in that word belong to the parent and must survive
// the download. The host's last-word column of rows [r0, r1) is fetched and merged into the
// device copy before the rows travel back (all on the stream: no extra synchronize).
static bool shares_last_word(const mzd_t *H) {
  return (H->ncols % 64) && (H->flags & (mzd_flag_windowed_zerooffset | mzd_flag_windowed_zeroexcess));
}
static void merge_foreign_excess(const mzd_t *H, const DMat *D, rci_t r0, rci_t r1, cudaStream_t s) {
  if (r1 <= r0 || !H->width) return;
  word *col;
  M4B_CUDA(cudaMallocAsync(&col, (size_t)(r1 - r0) * 8, s));
  rci_t i = r0;
  while (i < r1) {
    rci_t j = i + 1;
    while (j < r1 && H->rows[j] == H->rows[i] + (size_t)(j - i) * H->rowstride) j++;
    M4B_CUDA(cudaMemcpy2DAsync(col + (i - r0), 8, H->rows[i] + H->width - 1, (size_t)H->rowstride * 8, 8,
                               j - i, cudaMemcpyHostToDevice, s));
    i = j;
  }
  k_merge_excess(D->d + (size_t)r0 * D->rowstride + D->width - 1, D->rowstride, col, r1 - r0,
                 high_mask(H->ncols), s);
  M4B_CUDA(cudaFreeAsync(col, s));
}

static void download_rows(mzd_t *H, const DMat *D, rci_t r0, rci_t r1, cudaStream_t s) {
  if (D->nrows != H->nrows || D->ncols != H->ncols)
    die("m4ri_b200: download: shape mismatch\n");
  if (r1 <= r0 || H->width == 0) return;
  if (shares_last_word(H)) merge_foreign_excess(H, D, r0, r1, s);
  rci_t i = r0;
  while (i < r1) {
    rci_t j = i + 1;
    while (j < r1 && H->rows[j] == H->rows[i] + (size_t)(j - i) * H->rowstride) j++;
    copy_rows(H->rows[i], (size_t)H->rowstride * 8, D->d + (size_t)i * D->rowstride, (size_t)D->rowstride * 8,
              (size_t)H->width * 8, j - i, cudaMemcpyDeviceToHost, s);
    i = j;
  }
}
void download(mzd_t *H, const DMat *D, cudaStream_t s) { download_rows(H, D, 0, H->nrows, s); }

// ---- algorithm choice ---------------------------------------------------------------
static int resolve_algo(int algo, int m, int k, int n) {
  std::call_once(g_once, probe);
  if (algo == M4RI_B200_ALGO_AUTO) algo = g_algo.load();
  if (algo == M4RI_B200_ALGO_AUTO) {
    // Chosen on the evidence of profiles/ (DESIGN.md 5): the tcgen05 path wins at every
    // size from n = 2048 up (4.43 vs 2.92 Pbit-op/s at n = 32768). Tiny products are
    // launch-latency bound: the AND+popc kernel has the smallest fixed cost.
    long long work = (long long)m * n;
    if (work <= 64LL * 64 || k <= 64) algo = M4RI_B200_ALGO_SIMPLE;
    else algo = mma_available() ? M4RI_B200_ALGO_MMA : M4RI_B200_ALGO_M4RM;
  }
  return algo;
}

// ---- Strassen-Winograd outer recursion (what mzd_mul does above its cutoff,
// m4ri-sys/src/strassen.rs:5-18) ---------------------------------------------------
// form 0: C = A*B (B is k x n). form 1: C = A*Bt^T (Bt is n x k). 7 leaf products
// and 14 XOR passes + 1 copy on quadrant windows per level; the XORs are HBM-bound
// elementwise kernels (k_add) on quarter-size matrices.
DMat dwin(const DMat *M, int r0, int c0, int nr, int nc) {
  DMat W;
  W.d = M->d + (size_t)r0 * M->rowstride + c0 / 64;
  W.nrows = nr;
  W.ncols = nc;
  W.width = width_of(nc);
  W.rowstride = M->rowstride;
  W.base = nullptr;
  return W;
}

static void leaf_mul(DMat *C, const DMat *A, const DMat *B, int clear, int algo, int form,
                     cudaStream_t s) {
  if (form == 0) k_mul_m4rm(C, A, B, clear, s);
  else if (algo == M4RI_B200_ALGO_MMA) k_mul_abt_mma(C, A, B, clear, s);
  else k_mul_abt_simple(C, A, B, clear, s);
}

static DMat *dmat_new_overwritten(rci_t nrows, rci_t ncols, cudaStream_t s);
// ---- the last Strassen level, fused (tensor path, default kernel) --------------------------
// Classic Strassen over GF(2): every operand of the seven products is a quadrant or the sum of
// two, and every product goes into one or two quadrants of C:
//   M1 = (A11+A22)(B11+B22) -> C11, C22     M5 = (A11+A12) B22       -> C11, C12
//   M2 = (A21+A22) B11      -> C21, C22     M6 = (A21+A11)(B11+B12)  -> C22
//   M3 = A11 (B12+B22)      -> C12, C22     M7 = (A12+A22)(B21+B22)  -> C11
//   M4 = A22 (B21+B11)      -> C11, C21
// The ten operand sums are two launches (k_quad_sums: 4 quadrants read once, 5 sums written) and
// the product kernel's epilogue writes or accumulates into both destinations itself, so the level
// costs 2 add passes instead of the 15 of the Winograd schedule below (0.44 -> 0.15 ms at
// n = 32768) and C (^)= needs no temporary. Bt is n x k: quadrant (i, j) of B is rows j, columns
// i of Bt.
static bool fused_level_ok(const DMat *C, const DMat *A, int algo, int form) {
  const int v = mma_variant();
  return algo == M4RI_B200_ALGO_MMA && form == 1 && !getenv("M4RI_B200_NO_FUSED_STRASSEN") &&
         (v == M4RI_B200_MMA_FP4_S256_PAIR || v == M4RI_B200_MMA_FP4_S256_SINGLE) && C->nrows % 512 == 0 &&
         C->ncols % 512 == 0 && A->ncols % 512 == 0 && !(C->rowstride & 1) && !((uintptr_t)C->d & 15);
}
static void strassen_fused_level(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s) {
  const int m2 = C->nrows / 2, n2 = C->ncols / 2, k2 = A->ncols / 2;
  const int cg = mma_variant() == M4RI_B200_MMA_FP4_S256_SINGLE ? 1 : 2;
  DMat Aq[4] = {dwin(A, 0, 0, m2, k2), dwin(A, 0, k2, m2, k2), dwin(A, m2, 0, m2, k2), dwin(A, m2, k2, m2, k2)};
  // B11, B12, B21, B22 as windows of Bt
  DMat Bq[4] = {dwin(Bt, 0, 0, n2, k2), dwin(Bt, n2, 0, n2, k2), dwin(Bt, 0, k2, n2, k2), dwin(Bt, n2, k2, n2, k2)};
  DMat Cq[4] = {dwin(C, 0, 0, m2, n2), dwin(C, 0, n2, m2, n2), dwin(C, m2, 0, m2, n2), dwin(C, m2, n2, m2, n2)};
  DMat *SA[5], *SB[5];
  for (int i = 0; i < 5; i++) SA[i] = dmat_new_overwritten(m2, k2, s), SB[i] = dmat_new_overwritten(n2, k2, s);
  // A11+A22, A21+A22, A11+A12, A21+A11, A12+A22 / B11+B22, B12+B22, B21+B11, B11+B12, B21+B22
  static const int pa[5][2] = {{0, 3}, {2, 3}, {0, 1}, {2, 0}, {1, 3}};
  static const int pb[5][2] = {{0, 3}, {1, 3}, {2, 0}, {0, 1}, {2, 3}};
  k_quad_sums(SA, Aq, pa, s);
  k_quad_sums(SB, Bq, pb, s);
  if (!getenv("M4RI_B200_NO_GROUPED_STRASSEN")) {
    // all seven products as ONE persistent launch (no launch gaps, one partial last wave instead of
    // seven): destinations are shared between concurrent products, so C is zeroed first when it is
    // to be overwritten and every product accumulates with atomic XORs
    if (clear) k_zero(C, s);
    DMat *C1[7] = {&Cq[0], &Cq[2], &Cq[1], &Cq[0], &Cq[0], &Cq[3], &Cq[0]};
    DMat *C2[7] = {&Cq[3], &Cq[3], &Cq[3], &Cq[2], &Cq[1], nullptr, nullptr};
    const DMat *X[7] = {SA[0], SA[1], &Aq[0], &Aq[3], SA[2], SA[3], SA[4]};
    const DMat *Y[7] = {SB[0], &Bq[0], SB[1], SB[2], &Bq[3], SB[3], SB[4]};
    k_mul_abt_mma6_group(7, C1, C2, X, Y, cg, s);
  } else {
    bool fresh[4] = {clear != 0, clear != 0, clear != 0, clear != 0};  // the next write overwrites the quadrant
    auto prod = [&](const DMat *X, const DMat *Y, int t1, int t2) {
      k_mul_abt_mma6(&Cq[t1], X, Y, fresh[t1], cg, s, t2 >= 0 ? &Cq[t2] : nullptr, t2 >= 0 ? fresh[t2] : 0);
      fresh[t1] = false;
      if (t2 >= 0) fresh[t2] = false;
    };
    prod(SA[0], SB[0], 0, 3);   // M1
    prod(SA[1], &Bq[0], 2, 3);  // M2
    prod(&Aq[0], SB[1], 1, 3);  // M3
    prod(&Aq[3], SB[2], 0, 2);  // M4
    prod(SA[2], &Bq[3], 0, 1);  // M5
    prod(SA[3], SB[3], 3, -1);  // M6
    prod(SA[4], SB[4], 0, -1);  // M7
  }
  for (int i = 0; i < 5; i++) dmat_delete(SA[i], s), dmat_delete(SB[i], s);
}

// C = A (*) B, overwriting C. Quadrant (i,j) of the logical k x n matrix B.
static DMat *dmat_new_overwritten(rci_t nrows, rci_t ncols, cudaStream_t s);
static void strassen_mul(DMat *C, const DMat *A, const DMat *B, int algo, int form, int levels,
                         cudaStream_t s) {
  const int m = C->nrows, n = C->ncols, k = A->ncols;
  // halves must stay 128-bit aligned column windows and whole 256-row tiles
  if (levels <= 0 || (m % 512) || (n % 512) || (k % 512)) {
    leaf_mul(C, A, B, 1, algo, form, s);
    return;
  }
  if (levels == 1 && fused_level_ok(C, A, algo, form)) {
    strassen_fused_level(C, A, B, 1, s);
    return;
  }
  const int m2 = m / 2, n2 = n / 2, k2 = k / 2;
  auto Bq = [&](int i, int j) {
    return form == 0 ? dwin(B, i * k2, j * n2, k2, n2) : dwin(B, j * n2, i * k2, n2, k2);
  };
  DMat A11 = dwin(A, 0, 0, m2, k2), A12 = dwin(A, 0, k2, m2, k2), A21 = dwin(A, m2, 0, m2, k2),
       A22 = dwin(A, m2, k2, m2, k2);
  DMat B11 = Bq(0, 0), B12 = Bq(0, 1), B21 = Bq(1, 0), B22 = Bq(1, 1);
  DMat C11 = dwin(C, 0, 0, m2, n2), C12 = dwin(C, 0, n2, m2, n2), C21 = dwin(C, m2, 0, m2, n2),
       C22 = dwin(C, m2, n2, m2, n2);
  // (every word of the three temporaries is written before it is read: no memset)
  DMat *X = dmat_new_overwritten(m2, k2, s), *Y = dmat_new_overwritten(B11.nrows, B11.ncols, s),
       *Z = dmat_new_overwritten(m2, n2, s);
  const int L = levels - 1;
  k_add(X, &A21, &A22, s);                       // S1
  k_add(Y, &B11, &B12, s);                       // T1
  strassen_mul(&C22, X, Y, algo, form, L, s);    // P5
  k_copy(&C12, &C22, s);
  k_add(X, X, &A11, s);                          // S2
  k_add(Y, Y, &B22, s);                          // T2
  strassen_mul(&C21, X, Y, algo, form, L, s);    // P6
  strassen_mul(&C11, &A11, &B11, algo, form, L, s);  // P1
  k_add(&C21, &C21, &C11, s);                    // U2 = P1 + P6
  k_add(&C12, &C12, &C21, s);                    // U2 + P5
  k_add(&C22, &C22, &C21, s);                    // U2 + P5
  k_add(X, X, &A12, s);                          // S4
  strassen_mul(Z, X, &B22, algo, form, L, s);    // P3
  k_add(&C12, &C12, Z, s);                       // C12 done
  k_add(Y, Y, &B21, s);                          // T4
  strassen_mul(Z, &A22, Y, algo, form, L, s);    // P4
  k_add(&C21, &C21, Z, s);
  strassen_mul(Z, &A12, &B21, algo, form, L, s); // P2
  k_add(&C11, &C11, Z, s);                       // C11 done
  k_add(X, &A11, &A21, s);                       // S3
  k_add(Y, &B22, &B12, s);                       // T3
  strassen_mul(Z, X, Y, algo, form, L, s);       // P7
  k_add(&C21, &C21, Z, s);                       // C21 done
  k_add(&C22, &C22, Z, s);                       // C22 done
  dmat_delete(X, s);
  dmat_delete(Y, s);
  dmat_delete(Z, s);
}

// ---- burst or sustained? ----------------------------------------------------------------------
// The level rule below was tuned on short bursts. Under a long run of large products the board's
// power management lowers the SM clock (sw_power_cap: n = 32768 drifts 7.6 -> 8.0 ms within twenty
// products), and a schedule with one more Strassen level — 12.5 % fewer tensor operations on
// shorter, less efficient leaves — then wins and holds its time (profiles/r3e_levels_*.json: 7.685 ms
// flat against 7.74 and climbing). The library therefore counts large tensor-path products that
// arrive back to back (issued within 250 ms of the previous one); from the fifth on it takes the
// extra level where the leaves stay >= 8192. A lone call, or calls with pauses between them, keep
// the burst schedule. M4RI_B200_SUSTAINED=0 never switches, =1 always does. The result is the same
// bits either way.
static std::atomic<long long> g_last_large_ns{0};
static std::atomic<int> g_large_streak{0};
static int sustained_setting() {
  static const int v = [] {
    const char *e = getenv("M4RI_B200_SUSTAINED");
    return e && *e ? atoi(e) : -1;
  }();
  return v;
}
static void note_large_product(int m, int k, int n) {  // called once per executed device product
  if ((double)m * (double)k * (double)n < 16384.0 * 16384.0 * 16384.0) return;
  const long long now = std::chrono::duration_cast<std::chrono::nanoseconds>(
                            std::chrono::steady_clock::now().time_since_epoch()).count();
  const long long last = g_last_large_ns.exchange(now);
  if (last && now - last < 250000000LL) g_large_streak.fetch_add(1);
  else g_large_streak.store(0);
}
static bool sustained_load() {
  const int set = sustained_setting();
  return set >= 0 ? set != 0 : g_large_streak.load() >= 4;
}

// Levels of recursion for an m x k x n product: the explicit setting, else as many
// as keep every leaf dimension >= the cutoff (tuned on B200).
static int strassen_levels(int m, int k, int n, int algo, int cutoff) {
  if (cutoff < 0) return 0;  // the explicit Four-Russians entry points never recurse
  int set = g_strassen.load();
  if (set >= 0) return set;
  // tuned on B200 (profiles/r1g_size_sweep_fp4.jsonl): one level pays from n = 32768 up on the
  // tensor path (leaves of 16384), never on the Four-Russians path below n = 65536
  if (cutoff <= 0 && algo == M4RI_B200_ALGO_MMA) {
    // Measured rule (profiles/r2p_rect_sweep*.jsonl, r2p_size_sweep.jsonl). What decides whether a
    // level pays is the LEAF the recursion ends in: the tensor kernel's per-tile cost is fixed, so
    // its efficiency follows the leaf's contraction length (0.81 / 0.87 / 0.90 of the pipe at
    // k = 8192 / 16384 / 32768) and it needs whole 256-row tiles over all SMs.
    //   one level (the fused one: operand sums in two launches, the seven products one persistent
    //   launch): leaf k >= 8192, sides >= 4096, volume >= 8192^3 — 16384^3 1.079 -> 1.038 ms,
    //   8192 x 32768 x 16384 1.038 -> 0.998; no gain at 8192 x 32768 x 8192 or leaf k = 4096;
    //   more levels (unfused Winograd schedule above the fused one): leaf k >= 16384, sides >= 4096,
    //   volume >= 4096 x 16384 x 16384 — 65536^3 63.6 -> 57.9 ms, the 2-GPU row-block
    //   32768 x 65536 x 65536 31.5 -> 29.3, the 4-GPU one 15.9 -> 15.1; a second level with leaves
    //   of k = 8192 loses (32768^3: 7.62 -> 7.81).
    const bool fused = !getenv("M4RI_B200_NO_FUSED_STRASSEN") &&
                       (mma_variant() == M4RI_B200_MMA_FP4_S256_PAIR || mma_variant() == M4RI_B200_MMA_FP4_S256_SINGLE);
    auto leaf_ok = [&](int L) {
      if ((m % (256 << L)) || (k % (256 << L)) || (n % (256 << L))) return false;
      const long long lm = m >> L, lk = k >> L, ln = n >> L;
      if (lm < 4096 || ln < 4096) return false;
      const double vol = (double)lm * (double)lk * (double)ln;
      if (!fused) return lk >= 4096 && vol >= 16384.0 * 16384.0 * 16384.0;  // the r1 rule (other variants)
      if (L == 1) return lk >= 8192 && vol >= 8192.0 * 8192.0 * 8192.0;
      return lk >= 16384 && vol >= 4096.0 * 16384.0 * 16384.0;
    };
    int L = 0;
    while (L < 4 && leaf_ok(L + 1)) L++;
    if (fused && L >= 1 && L < 4 && sustained_load()) {
      // one more level than the burst rule, while the leaves keep k >= 8192, 4096-row sides and
      // the volume the fused level needs
      const int X = L + 1;
      const long long lm = m >> X, lk = k >> X, ln = n >> X;
      if (!(m % (256 << X)) && !(k % (256 << X)) && !(n % (256 << X)) && lm >= 4096 && ln >= 4096 && lk >= 8192 &&
          (double)lm * (double)lk * (double)ln >= 8192.0 * 8192.0 * 8192.0)
        L = X;
    }
    return L;
  }
  if (cutoff <= 0) cutoff = algo == M4RI_B200_ALGO_MMA ? 16384 : 32768;
  if (cutoff < 2048) cutoff = 2048;
  int L = 0, d = m < k ? m : k;
  d = d < n ? d : n;
  while (L < 4 && d / 2 >= cutoff) {
    d /= 2;
    L++;
  }
  return L;
}

static void product_with_strassen(DMat *C, const DMat *A, const DMat *B, int clear, int algo,
                                  int form, int cutoff, cudaStream_t s) {
  if (algo == M4RI_B200_ALGO_MMA) note_large_product(C->nrows, A->ncols, C->ncols);
  const int L = strassen_levels(C->nrows, A->ncols, C->ncols, algo, cutoff);
  if (L <= 0 || (C->nrows % 512) || (C->ncols % 512) || (A->ncols % 512)) {
    leaf_mul(C, A, B, clear, algo, form, s);
  } else if (L == 1 && fused_level_ok(C, A, algo, form)) {
    strassen_fused_level(C, A, B, clear, s);  // accumulates in the epilogue: no temporary
  } else if (clear) {
    strassen_mul(C, A, B, algo, form, L, s);
  } else {
    DMat *T = dmat_new(C->nrows, C->ncols, s);
    strassen_mul(T, A, B, algo, form, L, s);
    k_add(C, C, T, s);
    dmat_delete(T, s);
  }
}

// true when neither the call nor the process-wide setting forces an algorithm: the thin
// (HBM-bound) shapes may then skip the product kernels
static bool thin_shape_auto(int algo) {
  std::call_once(g_once, probe);
  return algo == M4RI_B200_ALGO_AUTO && g_algo.load() == M4RI_B200_ALGO_AUTO;
}
static bool aligned16(const DMat *M) { return !((uintptr_t)M->d & 15) && !(M->rowstride & 1); }

// C (^)= A*B on device
void dmul(DMat *C, const DMat *A, const DMat *B, int clear, int algo, cudaStream_t s, int cutoff) {
  if (A->ncols != B->nrows || C->nrows != A->nrows || C->ncols != B->ncols)
    die("m4ri_b200: mul: shapes (%d x %d) * (%d x %d) -> (%d x %d) do not match\n", A->nrows,
        A->ncols, B->nrows, B->ncols, C->nrows, C->ncols);
  if (thin_shape_auto(algo) && C->nrows > 0 && C->ncols > 0 && A->ncols > 0) {
    // HBM-bound shapes (no product kernel): a few rows of A = v*B (mul_va); a few
    // columns of B = a batched mat-vec over A (one pass), through two thin transposes
    if (C->nrows <= 4) {
      k_mul_va(C, A, B, clear, s);
      return;
    }
    if (C->ncols <= 32 && C->nrows >= 256 && aligned16(A) && aligned16(C)) {
      DMat *Bt = dmat_new(B->ncols, B->nrows, s), *Y = dmat_new(C->ncols, C->nrows, s);
      k_transpose(Bt, B, s);
      k_mul_v_batched(Y, A, Bt, 1, s);
      if (clear) {
        k_transpose(C, Y, s);
      } else {
        DMat *T = dmat_new(C->nrows, C->ncols, s);
        k_transpose(T, Y, s);
        k_add(C, C, T, s);
        dmat_delete(T, s);
      }
      dmat_delete(Bt, s);
      dmat_delete(Y, s);
      return;
    }
  }
  algo = resolve_algo(algo, C->nrows, A->ncols, C->ncols);
  if (algo == M4RI_B200_ALGO_M4RM) {
    product_with_strassen(C, A, B, clear, algo, 0, cutoff, s);
    return;
  }
  DMat *Bt = dmat_new(B->ncols, B->nrows, s);
  k_transpose(Bt, B, s);
  if (algo == M4RI_B200_ALGO_MMA) product_with_strassen(C, A, Bt, clear, algo, 1, cutoff, s);
  else k_mul_abt_simple(C, A, Bt, clear, s);
  dmat_delete(Bt, s);
}

// C (^)= A*Bt^T on device
static void dmul_abt(DMat *C, const DMat *A, const DMat *Bt, int clear, int algo, cudaStream_t s) {
  if (A->ncols != Bt->ncols || C->nrows != A->nrows || C->ncols != Bt->nrows)
    die("m4ri_b200: mul (pre-transposed): shapes (%d x %d) * (%d x %d)^T -> (%d x %d) do not match\n",
        A->nrows, A->ncols, Bt->nrows, Bt->ncols, C->nrows, C->ncols);
  if (thin_shape_auto(algo) && C->nrows > 0 && C->nrows <= 32 && C->ncols >= 256 && A->ncols > 0 &&
      aligned16(A) && aligned16(Bt)) {
    k_mul_v_batched(C, Bt, A, clear, s);  // rows of A are a few packed vectors against Bt: one pass over Bt
    return;
  }
  algo = resolve_algo(algo, C->nrows, A->ncols, C->ncols);
  if (algo == M4RI_B200_ALGO_MMA) {
    product_with_strassen(C, A, Bt, clear, algo, 1, 0, s);
  } else if (algo == M4RI_B200_ALGO_M4RM) {
    DMat *B = dmat_new(Bt->ncols, Bt->nrows, s);
    k_transpose(B, Bt, s);
    product_with_strassen(C, A, B, clear, algo, 0, 0, s);
    dmat_delete(B, s);
  } else {
    k_mul_abt_simple(C, A, Bt, clear, s);
  }
}

// Large host-ABI products on the tensor path: the PCIe copies are hidden behind the
// kernels. B goes up first (and is bit-transposed once); then A travels in row-chunks on a
// copy stream while the tensor cores multiply the previous chunk and a third stream brings
// finished rows of C back — H2D, compute and D2H overlap (separate copy engines), so the
// call costs about upload(B) + kernels instead of upload(A,B) + kernels + download(C).
static thread_local cudaStream_t t_h2d = nullptr, t_d2h = nullptr;
static int pipeline_parts(const char *env, int dflt) {  // experiment hook: chunk counts of the host pipeline
  const char *e = getenv(env);
  int v = e ? atoi(e) : 0;
  return v >= 1 && v <= 64 ? v : dflt;
}
static const size_t PIPELINE_MIN_BYTES = [] {
  const char *e = getenv("M4RI_B200_PIPE_MIN_MB");  // operands (A + B) from this size on take the copy/compute pipeline
  return (size_t)(e ? atoi(e) : 24) << 20;
}();

// Time factor of a pipeline piece for the Strassen levels it takes (profiles/r2p_rect_sweep*.jsonl):
// the small fused level saves ~4 %, a level on big leaves ~9.5 %, two levels ~15 %.
static double piece_level_factor(int pm, int k, int pn) {
  const int lv = (pm % 512 || pn % 512 || k % 512) ? 0 : strassen_levels(pm, k, pn, M4RI_B200_ALGO_MMA, 0);
  if (lv <= 0) return 1.0;
  const double leaf = (double)(pm >> lv) * (double)(k >> lv) * (double)(pn >> lv);
  return lv >= 2 ? 0.85 * pow(0.93, lv - 2) : (leaf >= 16384.0 * 16384.0 * 16384.0 ? 0.905 : 0.96);
}

// Row chunks of a GPU's block when its Bt is already complete on the device (the N-GPU single
// call): chunks of A go up, (chunk x k x n) products run, chunks of C go home. Few chunks keep the
// Strassen levels of a big product, many chunks hide more of the first upload and the last download.
static int plan_row_chunks(rci_t m, rci_t k, rci_t n, int clear) {
  static const int cand[] = {1, 2, 3, 4, 6, 8, 12, 16};
  // (contractions longer than the kernel's 32768-bit chunk take accumulating passes: +7 %)
  const double pcie = 50e9 * 1e-3, wave = k * 1.14e-6 * (k > 32768 ? 1.07 : 1.0), launch = 0.015;
  double best = 1e300;
  int best_na = 1;
  for (int ca : cand) {
    const rci_t per = ((m + ca - 1) / ca + 255) / 256 * 256;
    const int na = (m + per - 1) / per;
    double up = 0, t = 0, last = 0;
    for (int i = 0; i < na; i++) {
      const double rows = (double)((i + 1) * (long long)per < m ? per : m - i * (long long)per);
      up += rows * (k + (clear ? 0 : n)) / 8 / pcie;
      const double tiles = ceil(rows / 256) * ceil((double)n / 256);
      t = (t > up ? t : up) + ceil(tiles / 74) * wave * piece_level_factor((int)rows, k, n) + launch;
      last = rows;
    }
    t += last * n / 8 / pcie;
    if (t < best) best = t, best_na = na;
  }
  return best_na;
}

// Block plan of the pipeline below: (row-chunks of A) x (column blocks of B), chosen by replaying
// the schedule on a cost model — PCIe at 55 GB/s each way, a product = ceil(tiles / 74 CTA pairs)
// waves of 256 x 256 x k tiles at the measured 7.4 Pop/s (1.31 ns per wave per bit of k) + 8 us
// per launch. Few blocks start late (all of B's first block and A's first chunk must arrive);
// many blocks waste partial waves. The model reproduces the measured sweep at n = 32768 within
// 2 % (profiles/r1k_e2e_sweep.jsonl: 8x1 12.85, 8x4 11.15, 4x4 10.98, 8x8 13.2, 16x8 14.0 ms).
static void plan_pipeline(rci_t m, rci_t k, rci_t n, int clear, int *best_na, int *best_nb) {
  static const int cand[] = {1, 2, 3, 4, 6, 8, 12, 16};
  // constants fitted to profiles/r2p_e2e_sweep*.jsonl (n = 12288 ... 65536, rms error 4 %)
  const double pcie = 50e9 * 1e-3 /* bytes per ms */, wave = k * 1.14e-6 /* mma6 */ * (k > 32768 ? 1.07 : 1.0), launch = 0.015;
  double best = 1e300;
  *best_na = *best_nb = 1;
  for (int ca : cand)
    for (int cb : cand) {
      const rci_t per = ((m + ca - 1) / ca + 255) / 256 * 256, cper = ((n + cb - 1) / cb + 255) / 256 * 256;
      const int na = (m + per - 1) / per, nb = (n + cper - 1) / cper;
      auto rows_of = [&](int i) { return (double)((i + 1) * (long long)per < m ? per : m - i * (long long)per); };
      auto cols_of = [&](int j) { return (double)((j + 1) * (long long)cper < n ? cper : n - j * (long long)cper); };
      std::vector<double> arr_a(na), arr_b(nb);
      double up = cols_of(0) * k / 8 / pcie;
      arr_b[0] = up;
      for (int i = 0; i < na; i++) {
        up += rows_of(i) * (k + (clear ? 0 : n)) / 8 / pcie;
        arr_a[i] = up;
      }
      for (int j = 1; j < nb; j++) {
        up += cols_of(j) * k / 8 / pcie;
        arr_b[j] = up;
      }
      double t = 0;
      for (int j = 0; j < nb; j++)
        for (int i = 0; i < na; i++) {
          const double ready = arr_a[i] > arr_b[j] ? arr_a[i] : arr_b[j];
          const double tiles = ceil(rows_of(i) / 256) * ceil(cols_of(j) / 256);
          const double f = piece_level_factor((int)rows_of(i), k, (int)cols_of(j));
          t = (t > ready ? t : ready) + ceil(tiles / 74) * wave * f + launch;
        }
      t += rows_of(na - 1) * cols_of(nb - 1) / 8 / pcie;  // the last block's way home
      if (t < best) best = t, *best_na = na, *best_nb = nb;
    }
}

// Words [w0, w1) of every row of H -> the same words of D (a column block), and words [w0, w1)
// of rows [r0, r1) of D -> H: the 2-D copies of upload_rows/download_rows on a column window.
static void upload_cols(DMat *D, const mzd_t *H, wi_t w0, wi_t w1, cudaStream_t s) {
  if (w1 <= w0 || H->nrows == 0) return;
  rci_t i = 0;
  while (i < H->nrows) {
    rci_t j = i + 1;
    while (j < H->nrows && H->rows[j] == H->rows[i] + (size_t)(j - i) * H->rowstride) j++;
    M4B_CUDA(cudaMemcpy2DAsync(D->d + (size_t)i * D->rowstride + w0, (size_t)D->rowstride * 8,
                               H->rows[i] + w0, (size_t)H->rowstride * 8, (size_t)(w1 - w0) * 8,
                               j - i, cudaMemcpyHostToDevice, s));
    i = j;
  }
  if (w1 == H->width && D->ncols % 64) k_mask_excess(D, s);  // touches the last word of each row only
}
static void download_block(mzd_t *H, const DMat *D, rci_t r0, rci_t r1, wi_t w0, wi_t w1,
                           cudaStream_t s) {
  if (r1 <= r0 || w1 <= w0) return;
  if (w1 == H->width && shares_last_word(H)) merge_foreign_excess(H, D, r0, r1, s);
  rci_t i = r0;
  while (i < r1) {
    rci_t j = i + 1;
    while (j < r1 && H->rows[j] == H->rows[i] + (size_t)(j - i) * H->rowstride) j++;
    M4B_CUDA(cudaMemcpy2DAsync(H->rows[i] + w0, (size_t)H->rowstride * 8,
                               D->d + (size_t)i * D->rowstride + w0, (size_t)D->rowstride * 8,
                               (size_t)(w1 - w0) * 8, j - i, cudaMemcpyDeviceToHost, s));
    i = j;
  }
}

// Device buffer every word of which is about to be overwritten: skips dmat_new's memset when
// the rows have no padding word (stride == width).
static DMat *dmat_new_overwritten(rci_t nrows, rci_t ncols, cudaStream_t s) {
  if (dev_rowstride(ncols) != width_of(ncols) || nrows <= 0) return dmat_new(nrows, ncols, s);
  DMat *M = (DMat *)calloc(1, sizeof(DMat));
  if (!M) die("m4ri_b200: out of memory\n");
  M->nrows = nrows;
  M->ncols = ncols;
  M->width = M->rowstride = width_of(ncols);
  keep_pool_cached();
  M4B_CUDA(cudaMallocAsync(&M->base, (size_t)nrows * M->rowstride * 8, s));
  M->d = (word *)M->base;
  return M;
}

static void pipelined_host_product(mzd_t *C, const mzd_t *A, const mzd_t *B, int clear, int form,
                                   cudaStream_t s) {
  if (!t_h2d) M4B_CUDA(cudaStreamCreateWithFlags(&t_h2d, cudaStreamNonBlocking));
  if (!t_d2h) M4B_CUDA(cudaStreamCreateWithFlags(&t_d2h, cudaStreamNonBlocking));
  const rci_t m = A->nrows, n = C->ncols, k = A->ncols;
  DMat *dA = dmat_new_overwritten(m, k, s), *dB = dmat_new_overwritten(B->nrows, B->ncols, s),
       *dC = dmat_new_overwritten(m, n, s);
  DMat *dBt = form == 0 ? dmat_new_overwritten(n, k, s) : dB;
  // C is computed in (row-chunk of A) x (column block of B) pieces so that the tensor cores
  // start after one column block of B and one row-chunk of A have arrived (1/NB + 1/NA of the
  // upload), not after all of B. Blocks are multiples of the 256 x 256 tile.
  int na, nb;
  plan_pipeline(m, k, n, clear, &na, &nb);
  na = pipeline_parts("M4RI_B200_PIPE_ROWS", na);
  nb = pipeline_parts("M4RI_B200_PIPE_COLS", nb);
  rci_t per = ((m + na - 1) / na + 255) / 256 * 256;
  na = (m + per - 1) / per;
  rci_t cper = ((n + nb - 1) / nb + 255) / 256 * 256;
  nb = (n + cper - 1) / cper;
  std::vector<cudaEvent_t> ev(na + nb + na * nb + 1);
  for (auto &e : ev) M4B_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  cudaEvent_t *ev_a = ev.data(), *ev_b = ev.data() + na, *ev_c = ev.data() + na + nb;
  cudaEvent_t e_alloc = ev[na + nb + na * nb];
  M4B_CUDA(cudaEventRecord(e_alloc, s));  // buffers exist
  M4B_CUDA(cudaStreamWaitEvent(t_h2d, e_alloc, 0));
  auto upload_b = [&](int j) {
    const rci_t c0 = j * cper, c1 = c0 + cper < n ? c0 + cper : n;
    if (form == 0) upload_cols(dB, B, c0 / 64, (c1 + 63) / 64, t_h2d);  // k x n: a column block
    else upload_rows(dB, B, c0, c1, t_h2d);                             // n x k: a row block
    M4B_CUDA(cudaEventRecord(ev_b[j], t_h2d));
  };
  // upload order: B block 0, every row-chunk of A (the first column block's products follow
  // them one by one), then the other blocks of B
  upload_b(0);
  for (int i = 0; i < na; i++) {
    const rci_t r0 = i * per, r1 = r0 + per < m ? r0 + per : m;
    upload_rows(dA, A, r0, r1, t_h2d);
    if (!clear) upload_rows(dC, C, r0, r1, t_h2d);
    M4B_CUDA(cudaEventRecord(ev_a[i], t_h2d));
  }
  for (int j = 1; j < nb; j++) upload_b(j);
  for (int j = 0; j < nb; j++) {
    const rci_t c0 = j * cper, c1 = c0 + cper < n ? c0 + cper : n;
    M4B_CUDA(cudaStreamWaitEvent(s, ev_b[j], 0));
    DMat Btj = dwin(dBt, c0, 0, c1 - c0, k);
    if (form == 0) {
      DMat Bj = dwin(dB, 0, c0, k, c1 - c0);
      k_transpose(&Btj, &Bj, s);
    }
    for (int i = 0; i < na; i++) {
      const rci_t r0 = i * per, r1 = r0 + per < m ? r0 + per : m;
      DMat Ai = dwin(dA, r0, 0, r1 - r0, k), Cij = dwin(dC, r0, c0, r1 - r0, c1 - c0);
      if (j == 0) M4B_CUDA(cudaStreamWaitEvent(s, ev_a[i], 0));
      product_with_strassen(&Cij, &Ai, &Btj, clear, M4RI_B200_ALGO_MMA, 1, 0, s);  // a level where the block is big enough
      cudaEvent_t e = ev_c[j * na + i];
      M4B_CUDA(cudaEventRecord(e, s));
      M4B_CUDA(cudaStreamWaitEvent(t_d2h, e, 0));
      download_block(C, dC, r0, r1, c0 / 64, (c1 + 63) / 64, t_d2h);
    }
  }
  M4B_CUDA(cudaEventRecord(e_alloc, t_d2h));  // reuse: everything has left the device
  M4B_CUDA(cudaStreamWaitEvent(s, e_alloc, 0));
  if (dBt != dB) dmat_delete(dBt, s);
  dmat_delete(dA, s);
  dmat_delete(dB, s);
  dmat_delete(dC, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  for (auto &e : ev) M4B_CUDA(cudaEventDestroy(e));
}

// ---- several GPUs behind one host-ABI call ("single process ... what a library
// behind a C ABI should do; no launcher") ---------------------------------------------------
// m4ri_b200_set_devices(G) / M4RI_B200_DEVICES=G: a large host product is cut into G row-blocks
// of A and C; a resident worker thread per GPU runs the copy/compute pipeline above on its
// block — its rows of A over its own PCIe link — and writes its rows of C. B travels once when
// every pair of the GPUs is peer-capable (slab g over GPU g's link, then a transposing scatter
// over NVLink into every GPU's copy of B^T, multi_worker below); otherwise every GPU uploads all
// of B itself and the blocks are fully independent. One multi-GPU call at a time.
constexpr int MULTI_MAX = 8;            // GPUs one call may span (MultiJob's per-GPU arrays, scatter fan-out)
static std::atomic<int> g_devices{-1};  // -1: not decided yet (environment)

static int multi_devices() {
  int g = g_devices.load();
  if (g < 0) {
    const char *e = getenv("M4RI_B200_DEVICES");
    g = e ? atoi(e) : 1;
    if (g < 1) g = 1;
    g_devices.store(g);
  }
  std::call_once(g_once, probe);
  if (g > MULTI_MAX) g = MULTI_MAX;  // a box with more visible GPUs still uses the first eight
  return g < g_ndev ? g : (g_ndev > 0 ? g_ndev : 1);
}

struct DeviceWorker {
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  std::function<void()> job;
  bool pending = false;
  int dev = 0;
  void loop() {
    M4B_CUDA(cudaSetDevice(dev));
    std::unique_lock<std::mutex> lk(mu);
    for (;;) {
      cv.wait(lk, [&] { return pending; });
      std::function<void()> j = std::move(job);
      lk.unlock();
      j();  // dies loudly on any CUDA error, like every other entry point
      lk.lock();
      pending = false;
      cv.notify_all();
    }
  }
};
static std::mutex g_workers_mu;  // also serialises multi-GPU calls
static std::vector<DeviceWorker *> g_workers;

// ---- one product over the G GPUs of the process, B travelling ONCE over PCIe -----------------
// GPU g uploads only ITS k-slab of B (|B|/G over its own link), bit-transposes it and stores the
// slab of Bt into every GPU's Bt buffer with peer stores over NVLink (transpose_wide_kernel, the
// exchange of the torchrun path, sharded.py) — the pre-transposed form peer-copies row slabs of Bt
// instead. Its rows of A go up in chunks behind the slab; every chunk is multiplied as soon as it
// and the whole Bt have landed, and its rows of C go home while the next chunk is multiplied.
// Host traffic: |A| + |B| up, |C| down in total (round 1: |A| + G |B| up).
// Cross-device ordering is by events; an event must have been RECORDED (on the host timeline)
// before another thread makes a stream wait for it, hence the two host barriers.
struct HostBarrier {
  std::mutex mu;
  std::condition_variable cv;
  int n = 0, waiting = 0;
  unsigned long long gen = 0;
  void wait() {
    std::unique_lock<std::mutex> lk(mu);
    const unsigned long long g = gen;
    if (++waiting == n) {
      waiting = 0;
      gen++;
      cv.notify_all();
    } else {
      cv.wait(lk, [&] { return gen != g; });
    }
  }
};
struct MultiJob {
  int G = 0, used = 0, clear = 0, form = 0;
  mzd_t *C = nullptr;
  const mzd_t *A = nullptr, *B = nullptr;
  rci_t per = 0, slab = 0;  // rows of A/C per GPU; rows of B (form 0) or of Bt (form 1) per GPU
  DMat *dBt[MULTI_MAX] = {};
  cudaEvent_t e_alloc[MULTI_MAX] = {}, e_xchg[MULTI_MAX] = {};
  HostBarrier bar;
};

static std::atomic<int> g_peer_ok{-1};  // -1 unknown, 0 no, 1 every pair of the first G devices is peer-capable
static int g_peer_G = 0;
static bool peers_ready(int G) {  // under g_workers_mu
  if (g_peer_ok.load() >= 0 && g_peer_G >= G) return g_peer_ok.load() == 1;
  bool ok = true;
  for (int a = 0; a < G && ok; a++)
    for (int b = 0; b < G && ok; b++) {
      int can = 0;
      if (a != b && (cudaDeviceCanAccessPeer(&can, a, b) != cudaSuccess || !can)) ok = false;
    }
  (void)cudaGetLastError();
  if (getenv("M4RI_B200_NO_PEER")) ok = false;  // experiment/test hook: the independent-upload plan
  g_peer_ok.store(ok ? 1 : 0);
  g_peer_G = G;
  return ok;
}
// called by worker `dev` (its device current): peers may read/write this device's pool memory
static void open_pool_to_peers(int dev, int G) {
  static std::atomic<int> opened[64];  // number of leading devices this one is already open to
  if (dev >= 64 || opened[dev].load() >= G) return;  // (set_devices may grow G between calls)
  cudaMemPool_t pool;
  M4B_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
  for (int p = 0; p < G; p++) {
    if (p == dev) continue;
    cudaMemAccessDesc d = {};
    d.location.type = cudaMemLocationTypeDevice;
    d.location.id = p;
    d.flags = cudaMemAccessFlagsProtReadWrite;
    M4B_CUDA(cudaMemPoolSetAccess(pool, &d, 1));
    cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);  // kernels of this device may address p's memory
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) M4B_CUDA(e);
    (void)cudaGetLastError();
  }
  opened[dev].store(G);
}

static void multi_worker(MultiJob *J, int g) {
  if (!t_h2d) M4B_CUDA(cudaStreamCreateWithFlags(&t_h2d, cudaStreamNonBlocking));
  if (!t_d2h) M4B_CUDA(cudaStreamCreateWithFlags(&t_d2h, cudaStreamNonBlocking));
  cudaStream_t s = current_stream();
  open_pool_to_peers(g, J->G);
  const mzd_t *A = J->A, *B = J->B;
  mzd_t *C = J->C;
  const int clear = J->clear, form = J->form;
  const rci_t M = A->nrows, k = A->ncols, n = C->ncols;
  const rci_t r0 = (rci_t)g * J->per < M ? (rci_t)g * J->per : M, r1 = r0 + J->per < M ? r0 + J->per : M, m = r1 - r0;
  const rci_t slab_total = form == 0 ? k : n;  // rows of the operand B as the caller holds it
  const rci_t s0 = (rci_t)g * J->slab < slab_total ? (rci_t)g * J->slab : slab_total,
              s1 = s0 + J->slab < slab_total ? s0 + J->slab : slab_total;
  // shallow views: rows [r0, r1) of A and C, rows [s0, s1) of B, through the callers' rows[]
  mzd_t Av = *A, Cv = *C, Bv = *B;
  Av.nrows = Cv.nrows = m;
  Av.rows = A->rows + r0;
  Cv.rows = C->rows + r0;
  Bv.nrows = s1 - s0;
  Bv.rows = B->rows + s0;
  DMat *dA = dmat_new_overwritten(m, k, s), *dC = dmat_new_overwritten(m, n, s);
  DMat *dBt = dmat_new_overwritten(n, k, s);
  DMat *dBs = form == 0 ? dmat_new_overwritten(s1 - s0, n, s) : nullptr;
  J->dBt[g] = dBt;
  M4B_CUDA(cudaEventCreateWithFlags(&J->e_alloc[g], cudaEventDisableTiming));
  M4B_CUDA(cudaEventCreateWithFlags(&J->e_xchg[g], cudaEventDisableTiming));
  M4B_CUDA(cudaEventRecord(J->e_alloc[g], s));
  M4B_CUDA(cudaStreamWaitEvent(t_h2d, J->e_alloc[g], 0));
  // ---- uploads: this GPU's slab of B first, then its rows of A in chunks ----
  cudaEvent_t e_slab;
  M4B_CUDA(cudaEventCreateWithFlags(&e_slab, cudaEventDisableTiming));
  if (s1 > s0) {
    if (form == 0) {
      upload_rows(dBs, &Bv, 0, s1 - s0, t_h2d);
    } else {  // rows [s0, s1) of Bt go straight into place
      DMat W = dwin(dBt, s0, 0, s1 - s0, k);
      upload_rows(&W, &Bv, 0, s1 - s0, t_h2d);
    }
  }
  M4B_CUDA(cudaEventRecord(e_slab, t_h2d));
  int na = m > 0 ? plan_row_chunks(m, k, n, clear) : 1;
  const bool forced = getenv("M4RI_B200_PIPE_ROWS") != nullptr;
  na = pipeline_parts("M4RI_B200_PIPE_ROWS", na);
  rci_t per = m > 0 ? ((m + na - 1) / na + 255) / 256 * 256 : 256;
  na = m > 0 ? (m + per - 1) / per : 0;
  // chunk i = rows [cut[i], cut[i + 1])
  std::vector<rci_t> cut;
  if (!forced && na <= 6 && m >= 12288 && m % 1024 == 0 && k >= 32768 && !getenv("M4RI_B200_UNIFORM_CHUNKS")) {
    // Few big chunks keep their Strassen levels but expose the first upload and the last download:
    // a small chunk at either end (2048 rows: nothing to wait for after the exchange, a short way
    // home at the end) around chunks as big as the levels like them (12288 rows take one level,
    // 16384 two); a chunk's upload hides behind its predecessor's product while it has at most
    // six times its rows (PCIe 0.16 us per 64 Ki-bit row, tensor cores ~1 us).
    cut.push_back(0);
    cut.push_back(2048);
    rci_t left = m - 4096, piece = left < 12288 ? left : 12288;
    while (left > 0) {
      cut.push_back(cut.back() + piece);
      left -= piece;
      piece = left < 16384 ? left : 16384;
    }
    cut.push_back(m);
  } else {
    for (int i = 0; i < na; i++) cut.push_back(i * per);
    cut.push_back(m);
  }
  na = (int)cut.size() - 1;
  std::vector<cudaEvent_t> ev(2 * na + 1);
  for (auto &e : ev) M4B_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (int i = 0; i < na; i++) {
    const rci_t a0 = cut[i], a1 = cut[i + 1];
    upload_rows(dA, &Av, a0, a1, t_h2d);
    if (!clear) upload_rows(dC, &Cv, a0, a1, t_h2d);
    M4B_CUDA(cudaEventRecord(ev[i], t_h2d));
  }
  J->bar.wait();  // every GPU's Bt buffer exists and its e_alloc is recorded
  // ---- exchange: my slab -> every GPU's Bt ----
  M4B_CUDA(cudaStreamWaitEvent(s, e_slab, 0));
  for (int d = 0; d < J->used; d++)
    if (d != g) M4B_CUDA(cudaStreamWaitEvent(s, J->e_alloc[d], 0));
  if (s1 > s0) {
    if (form == 0) {
      DMat wins[MULTI_MAX];
      DMat *dst[MULTI_MAX];
      for (int d = 0; d < J->used; d++) {
        wins[d] = dwin(J->dBt[d], 0, s0, n, s1 - s0);  // columns [s0, s1) of Bt: 128-byte aligned (slabs of 1024 rows)
        dst[d] = &wins[d];
      }
      k_transpose_scatter(dst, J->used, dBs, s);
    } else {
      for (int d = 0; d < J->used; d++)
        if (d != g)
          M4B_CUDA(cudaMemcpyPeerAsync(J->dBt[d]->d + (size_t)s0 * J->dBt[d]->rowstride, d,
                                       dBt->d + (size_t)s0 * dBt->rowstride, g,
                                       (size_t)(s1 - s0) * dBt->rowstride * 8, s));
    }
  }
  M4B_CUDA(cudaEventRecord(J->e_xchg[g], s));
  J->bar.wait();  // every exchange is recorded
  for (int d = 0; d < J->used; d++)
    if (d != g) M4B_CUDA(cudaStreamWaitEvent(s, J->e_xchg[d], 0));
  // ---- products, chunk by chunk, rows of C going home behind them ----
  for (int i = 0; i < na; i++) {
    const rci_t a0 = cut[i], a1 = cut[i + 1];
    DMat Ai = dwin(dA, a0, 0, a1 - a0, k), Ci = dwin(dC, a0, 0, a1 - a0, n);
    M4B_CUDA(cudaStreamWaitEvent(s, ev[i], 0));
    product_with_strassen(&Ci, &Ai, dBt, clear, M4RI_B200_ALGO_MMA, 1, 0, s);
    M4B_CUDA(cudaEventRecord(ev[na + i], s));
    M4B_CUDA(cudaStreamWaitEvent(t_d2h, ev[na + i], 0));
    download_rows(&Cv, dC, a0, a1, t_d2h);
  }
  M4B_CUDA(cudaEventRecord(ev[2 * na], t_d2h));
  M4B_CUDA(cudaStreamWaitEvent(s, ev[2 * na], 0));
  M4B_CUDA(cudaStreamSynchronize(s));
  J->bar.wait();  // nobody frees a Bt buffer a peer may still be writing or reading
  dmat_delete(dA, s);
  dmat_delete(dC, s);
  dmat_delete(dBt, s);
  if (dBs) dmat_delete(dBs, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  for (auto &e : ev) M4B_CUDA(cudaEventDestroy(e));
  M4B_CUDA(cudaEventDestroy(e_slab));
}

static void run_on_workers(int used, const std::function<void(int)> &fn) {  // under g_workers_mu
  for (int i = 0; i < used; i++) {
    DeviceWorker *w = g_workers[i];
    {
      std::lock_guard<std::mutex> lk(w->mu);
      w->job = [=] { fn(i); };
      w->pending = true;
    }
    w->cv.notify_all();
  }
  for (int i = 0; i < used; i++) {
    DeviceWorker *w = g_workers[i];
    std::unique_lock<std::mutex> lk(w->mu);
    w->cv.wait(lk, [&] { return !w->pending; });
  }
}

static void multi_device_host_product(mzd_t *C, const mzd_t *A, const mzd_t *B, int clear, int form, int G) {
  std::lock_guard<std::mutex> g(g_workers_mu);
  while ((int)g_workers.size() < G) {  // workers live for the rest of the process
    DeviceWorker *w = new DeviceWorker;
    w->dev = (int)g_workers.size();
    w->th = std::thread([w] { w->loop(); });
    w->th.detach();
    g_workers.push_back(w);
  }
  const rci_t m = A->nrows;
  const rci_t per = ((m + G - 1) / G + 255) / 256 * 256;
  const int used = (m + per - 1) / per;
  if (used > 1 && peers_ready(used)) {
    // B travels once: slab g of it over GPU g's link, then NVLink
    MultiJob J;
    J.G = J.used = used;
    J.clear = clear;
    J.form = form;
    J.C = C;
    J.A = A;
    J.B = B;
    J.per = per;
    const rci_t rows_b = B->nrows;  // k (form 0) or n (form 1)
    J.slab = ((rows_b + used - 1) / used + 1023) / 1024 * 1024;
    J.bar.n = used;
    run_on_workers(used, [&](int i) { multi_worker(&J, i); });
    for (int i = 0; i < used; i++) {
      M4B_CUDA(cudaEventDestroy(J.e_alloc[i]));
      M4B_CUDA(cudaEventDestroy(J.e_xchg[i]));
    }
    return;
  }
  // no peer access between the GPUs: independent row-blocks, every GPU uploads all of B
  std::vector<mzd_t> Av(used), Cv(used);
  for (int i = 0; i < used; i++) {
    const rci_t r0 = (rci_t)i * per, r1 = r0 + per < m ? r0 + per : m;
    Av[i] = *A;
    Av[i].nrows = r1 - r0;
    Av[i].rows = A->rows + r0;
    Cv[i] = *C;
    Cv[i].nrows = r1 - r0;
    Cv[i].rows = C->rows + r0;
  }
  run_on_workers(used, [&](int i) { pipelined_host_product(&Cv[i], &Av[i], B, clear, form, current_stream()); });
}

// y (^)= A*x for the shapes the reference's mat-vec calls have: BinMatrix::mul_slice passes x as an
// n x 1 column matrix to mzd_mul_naive (m4ri-rust/src/friendly/binary_matrix.rs:426-427) and gets
// an m x 1 column back (one word per row). The column is packed on the host (O(n)), A is read
// once by the HBM-bound mat-vec kernels (kernels_aux.cu: mul_v_*), the packed y is spread over
// the m one-word rows (O(m)) — no transpose, no 64x64 product tile for one useful column.
// x_is_row: x is already a packed 1 x n row (the pre-transposed form, _mzd_mul_naive with a
// one-row Bt).
static mzd_t *host_matvec(mzd_t *y, const mzd_t *A, const mzd_t *x, int x_is_row, int clear) {
  const rci_t m = A->nrows, n = A->ncols;
  cudaStream_t s = current_stream();
  DMat *dA = dmat_new(m, n, s), *dx = dmat_new(1, n, s), *dy = dmat_new(1, m, s);
  upload(dA, A, s);
  const int xw = width_of(n), yw = width_of(m);
  word *hx = (word *)calloc((size_t)(xw > 0 ? xw : 1), 8), *hy = (word *)calloc((size_t)(yw > 0 ? yw : 1), 8);
  if (!hx || !hy) die("m4ri_b200: out of memory\n");
  if (x_is_row) {
    memcpy(hx, x->rows[0], (size_t)xw * 8);
    if (xw) hx[xw - 1] &= high_mask(n);
  } else {
    for (rci_t j = 0; j < n; j++) hx[j / 64] |= (x->rows[j][0] & 1) << (j % 64);
  }
  if (xw) M4B_CUDA(cudaMemcpyAsync(dx->d, hx, (size_t)xw * 8, cudaMemcpyHostToDevice, s));
  k_mul_v(dy, dA, dx, 1, s);
  if (yw) M4B_CUDA(cudaMemcpyAsync(hy, dy->d, (size_t)yw * 8, cudaMemcpyDeviceToHost, s));
  dmat_delete(dA, s);
  dmat_delete(dx, s);
  dmat_delete(dy, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  // (a window's word also holds the parent's other columns: keep them)
  const word keep = shares_last_word(y) ? ~(word)1 : 0;
  if (m > 0 && !keep && clear && y->rows[m - 1] == y->rows[0] + (size_t)(m - 1) * y->rowstride) {
    // the usual case (a fresh m x 1 result of this library): evenly strided rows, no indirection
    word *p = y->rows[0];
    const size_t st = (size_t)y->rowstride;
    for (rci_t i = 0; i < m; i++) p[i * st] = (hy[i >> 6] >> (i & 63)) & 1;
  } else {
    for (rci_t i = 0; i < m; i++) {
      const word old = y->rows[i][0], bit = (hy[i / 64] >> (i % 64)) & 1;
      y->rows[i][0] = (old & keep) | ((clear ? bit : (old ^ bit)) & 1);
    }
  }
  free(hx);
  free(hy);
  return y;
}

// ---- small products: one copy each way ----------------------------------------------------
// Below 256 KiB per operand a product through the ABI is mostly fixed costs: three pool
// allocations with their memsets, two or three copies out of pageable memory (small matrices are
// malloc'ed: the driver stages each of them), the excess-bit kernels, the frees. The small path
// packs A and B (and C when accumulating) into ONE pinned staging block in the device layout
// (excess bits masked on the way), moves it with one H2D copy into a per-thread device arena,
// runs the usual kernels on views of the arena, brings C back with one D2H copy and unpacks it
// into the caller's rows. Same kernels, same bits: 59-72 us -> see tools/entry_point_time.py.
struct TinyArena {
  int dev = -1;
  word *d = nullptr, *h = nullptr;  // device arena / pinned staging, TINY_WORDS each
};
constexpr size_t TINY_OPERAND_MAX = (size_t)256 << 10;   // arena: three operands of this size
static const size_t TINY_OPERAND_BYTES = [] {             // per operand, in the device layout
  const char *e = getenv("M4RI_B200_TINY_KB");
  size_t kb = e ? (size_t)atoi(e) : 256;  // = the size from which mzd_init hands out pinned memory
  return (kb > 256 ? 256 : kb) << 10;
}();
constexpr size_t TINY_WORDS = 3 * TINY_OPERAND_MAX / 8;
static thread_local TinyArena t_tiny;

static size_t tiny_words(rci_t nrows, rci_t ncols) { return (size_t)nrows * dev_rowstride(ncols); }
// rows of H -> dst in the device layout (stride words per row), bits beyond ncols cleared
static void tiny_pack(word *dst, const mzd_t *H) {
  const int stride = dev_rowstride(H->ncols);
  const word hm = H->high_bitmask;
  for (rci_t i = 0; i < H->nrows; i++) {
    word *d = dst + (size_t)i * stride;
    const word *src = H->rows[i];
    for (wi_t w = 0; w + 1 < H->width; w++) d[w] = src[w];
    if (H->width) d[H->width - 1] = src[H->width - 1] & hm;
    for (int w = H->width; w < stride; w++) d[w] = 0;
  }
}
static void tiny_unpack(mzd_t *H, const word *src) {
  const int stride = dev_rowstride(H->ncols);
  const word hm = H->high_bitmask;
  const bool keep = shares_last_word(H);
  for (rci_t i = 0; i < H->nrows; i++) {
    const word *s_ = src + (size_t)i * stride;
    word *d = H->rows[i];
    for (wi_t w = 0; w + 1 < H->width; w++) d[w] = s_[w];
    if (H->width)  // a window keeps its parent's bits; a whole matrix has zero excess bits
      d[H->width - 1] = keep ? ((d[H->width - 1] & ~hm) | (s_[H->width - 1] & hm)) : (s_[H->width - 1] & hm);
  }
}
static DMat tiny_view(word *d, rci_t nrows, rci_t ncols) {
  DMat M;
  M.d = d;
  M.nrows = nrows;
  M.ncols = ncols;
  M.width = width_of(ncols);
  M.rowstride = dev_rowstride(ncols);
  M.base = nullptr;
  return M;
}
static void dmul_abt(DMat *C, const DMat *A, const DMat *Bt, int clear, int algo, cudaStream_t s);
// true when the product was done here
static bool tiny_host_product(mzd_t *C, const mzd_t *A, const mzd_t *B, int clear, int algo, int form, int cutoff,
                              cudaStream_t s) {
  static const bool off = getenv("M4RI_B200_NO_TINY_PATH") != nullptr;
  if (off || form > 1) return false;
  const size_t wa = tiny_words(A->nrows, A->ncols), wb = tiny_words(B->nrows, B->ncols),
               wc = tiny_words(C->nrows, C->ncols);
  if (wa == 0 || wb == 0 || wc == 0) return false;
  if (wa * 8 > TINY_OPERAND_BYTES || wb * 8 > TINY_OPERAND_BYTES || wc * 8 > TINY_OPERAND_BYTES) return false;
  int dev = 0;
  M4B_CUDA(cudaGetDevice(&dev));
  TinyArena &T = t_tiny;
  if (T.dev != dev) {  // first use on this thread (or the thread moved to another GPU)
    if (T.d) {
      (void)cudaFree(T.d);
      (void)cudaFreeHost(T.h);
    }
    M4B_CUDA(cudaMalloc(&T.d, TINY_WORDS * 8));
    M4B_CUDA(cudaMallocHost(&T.h, TINY_WORDS * 8));
    T.dev = dev;
  }
  // layout: [A | B | C], every region 16-byte aligned (strides are even)
  const size_t oa = 0, ob = wa, oc = wa + wb;
  tiny_pack(T.h + oa, A);
  tiny_pack(T.h + ob, B);
  if (!clear) tiny_pack(T.h + oc, C);
  M4B_CUDA(cudaMemcpyAsync(T.d, T.h, (clear ? oc : oc + wc) * 8, cudaMemcpyHostToDevice, s));
  DMat dA = tiny_view(T.d + oa, A->nrows, A->ncols), dB = tiny_view(T.d + ob, B->nrows, B->ncols),
       dC = tiny_view(T.d + oc, C->nrows, C->ncols);
  if (form == 0) dmul(&dC, &dA, &dB, clear, algo, s, cutoff);
  else dmul_abt(&dC, &dA, &dB, clear, algo, s);
  M4B_CUDA(cudaMemcpyAsync(T.h + oc, T.d + oc, wc * 8, cudaMemcpyDeviceToHost, s));
  M4B_CUDA(cudaStreamSynchronize(s));
  tiny_unpack(C, T.h + oc);
  return true;
}

// Shared body of the host-ABI products. form: 0 = A*B, 1 = A*Bt^T, 2 = v*A.
static mzd_t *host_product(mzd_t *C, const mzd_t *A, const mzd_t *B, int clear, int algo, int form,
                           const char *who, int cutoff = 0) {
  rci_t m = A->nrows, k = A->ncols;
  rci_t n = form == 1 ? B->nrows : B->ncols;
  rci_t kb = form == 1 ? B->ncols : B->nrows;
  if (k != kb) die("m4ri_b200: %s: A ncols (%d) need to match B nrows (%d)\n", who, k, kb);
  if (!C) {
    if (!clear) die("m4ri_b200: %s: C must not be NULL when accumulating\n", who);
    C = mzd_init_impl(m, n, 0);  // every valid word is downloaded over
  } else if (C->nrows != m || C->ncols != n) {
    die("m4ri_b200: %s: C (%d x %d) has wrong dimensions, expected (%d x %d)\n", who, C->nrows,
        C->ncols, m, n);
  }
  if (C == A || C == B) die("m4ri_b200: %s: C must not alias an input\n", who);
  // The *_naive entry points name M4RI's cubic algorithm; the result does not depend on the
  // algorithm, and from 2^37 bit-operations up (n = 4096: 0.37 vs 0.21 ms; n = 8192: 1.86 vs
  // 0.68 ms; 12x at n = 32768) the AND+popcount kernel only costs time: large products behind
  // these names take the library's own choice, like the mat-vec shapes below. (The kernel stays
  // reachable at any size through m4ri_b200_set_algo / m4ri_b200_dmul with ALGO_SIMPLE.)
  if (algo == M4RI_B200_ALGO_SIMPLE && 2.0 * (double)m * (double)k * (double)n >= 137438953472.0 &&
      !getenv("M4RI_B200_NAIVE_IS_NAIVE"))
    algo = M4RI_B200_ALGO_AUTO;
  cudaStream_t s = current_stream();
  // The reference's mat-vec shapes never need a product kernel (whatever algorithm the entry
  // point names — over GF(2) the result is the same):
  //   A * x, x an n x 1 column (mul_slice, binary_matrix.rs:426-427) or a one-row Bt  -> mul_v_*
  //   v * A with a few rows of v (Mul<&BinMatrix> for &BinVector, binary_matrix.rs:552-563) -> mul_va
  if (m > 0 && n == 1 && k > 0 && (form == 0 || form == 1)) return host_matvec(C, A, B, form == 1, clear);
  if (form == 0 && m > 0 && m <= 4 && n > 0 && k > 0) form = 2;
  if (m > 0 && n > 1 && k > 0 && tiny_host_product(C, A, B, clear, algo, form, cutoff, s)) return C;
  if (form != 2 && m > 0 && n > 0 && k > 0 && resolve_algo(algo, m, k, n) == M4RI_B200_ALGO_MMA &&
      ((size_t)m * A->width + (size_t)B->nrows * B->width) * 8 >= PIPELINE_MIN_BYTES) {
    const int G = multi_devices();
    if (G > 1 && m >= 512) {  // row-blocks of A and C over the GPUs of this process
      multi_device_host_product(C, A, B, clear, form, G);
      return C;
    }
    // (every block of the pipeline takes a Strassen level where its shape pays for one)
    pipelined_host_product(C, A, B, clear, form, s);
    return C;
  }
  DMat *dA = dmat_new(A->nrows, A->ncols, s), *dB = dmat_new(B->nrows, B->ncols, s),
       *dC = dmat_new(m, n, s);
  upload(dA, A, s);
  upload(dB, B, s);
  if (!clear) upload(dC, C, s);
  if (form == 0) dmul(dC, dA, dB, clear, algo, s, cutoff);
  else if (form == 1) dmul_abt(dC, dA, dB, clear, algo, s);
  else k_mul_va(dC, dA, dB, clear, s);
  download(C, dC, s);
  dmat_delete(dA, s);
  dmat_delete(dB, s);
  dmat_delete(dC, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return C;
}

}  // namespace m4b

using namespace m4b;

extern "C" {

// ---- hot-path entry points (reference declarations cited in include/m4ri_b200.h) ----
mzd_t *mzd_mul(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff) {
  // cutoff = minimal dimension for the Strassen recursion (0 = tuned default); the
  // product does not depend on it
  return host_product(C, A, B, 1, M4RI_B200_ALGO_AUTO, 0, "mzd_mul", cutoff < 0 ? 0 : cutoff);
}
mzd_t *mzd_addmul(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff) {
  return host_product(C, A, B, 0, M4RI_B200_ALGO_AUTO, 0, "mzd_addmul", cutoff < 0 ? 0 : cutoff);
}
// mp.rs:6,8 — M4RI's OpenMP products: here every product already runs on all SMs (and, with
// m4ri_b200_set_devices, on all GPUs of the process), so they are the same calls
mzd_t *mzd_mul_mp(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff) { return mzd_mul(C, A, B, cutoff); }
mzd_t *mzd_addmul_mp(mzd_t *C, const mzd_t *A, const mzd_t *B, int cutoff) { return mzd_addmul(C, A, B, cutoff); }
mzd_t *mzd_mul_m4rm(mzd_t *C, const mzd_t *A, const mzd_t *B, int k) {
  (void)k;  // table width is fixed by the shared-memory budget
  return host_product(C, A, B, 1, M4RI_B200_ALGO_M4RM, 0, "mzd_mul_m4rm", -1);
}
mzd_t *mzd_addmul_m4rm(mzd_t *C, const mzd_t *A, const mzd_t *B, int k) {
  (void)k;
  return host_product(C, A, B, 0, M4RI_B200_ALGO_M4RM, 0, "mzd_addmul_m4rm", -1);
}
mzd_t *_mzd_mul_m4rm(mzd_t *C, const mzd_t *A, const mzd_t *B, int k, int clear) {
  (void)k;
  return host_product(C, A, B, clear != 0, M4RI_B200_ALGO_M4RM, 0, "_mzd_mul_m4rm", -1);
}
mzd_t *mzd_mul_naive(mzd_t *C, const mzd_t *A, const mzd_t *B) {
  return host_product(C, A, B, 1, M4RI_B200_ALGO_SIMPLE, 0, "mzd_mul_naive");
}
mzd_t *mzd_addmul_naive(mzd_t *C, const mzd_t *A, const mzd_t *B) {
  return host_product(C, A, B, 0, M4RI_B200_ALGO_SIMPLE, 0, "mzd_addmul_naive");
}
mzd_t *_mzd_mul_naive(mzd_t *C, const mzd_t *A, const mzd_t *Bt, int clear) {
  return host_product(C, A, Bt, clear != 0, M4RI_B200_ALGO_SIMPLE, 1, "_mzd_mul_naive");
}
mzd_t *_mzd_mul_va(mzd_t *C, const mzd_t *v, const mzd_t *A, int clear) {
  return host_product(C, v, A, clear != 0, M4RI_B200_ALGO_AUTO, 2, "_mzd_mul_va");
}

// y = A*x with x (n x 1) and y (m x 1) column matrices, one word per row
// (the shape BinMatrix::mul_slice builds, binary_matrix.rs:416-431). The column
// <-> packed-row marshalling is O(m+n) on the host; A never leaves packed form.
mzd_t *mzd_mul_v(mzd_t *y, const mzd_t *A, const mzd_t *x) {
  rci_t m = A->nrows, n = A->ncols;
  if (x->nrows != n || x->ncols != 1)
    die("m4ri_b200: mzd_mul_v: x must be %d x 1 (got %d x %d)\n", n, x->nrows, x->ncols);
  if (!y) y = mzd_init_impl(m, 1, 0);
  else if (y->nrows != m || y->ncols != 1)
    die("m4ri_b200: mzd_mul_v: y must be %d x 1\n", m);
  return host_matvec(y, A, x, 0, 1);
}

// C = A ^ B on the GPU (the Strassen / addmul "add"; reference mzd_add mzd.rs:223).
// C may be NULL (allocate) or alias A or B (binary_matrix.rs:497-523).
mzd_t *mzd_add(mzd_t *C, const mzd_t *A, const mzd_t *B) {
  if (A->nrows != B->nrows || A->ncols != B->ncols)
    die("m4ri_b200: mzd_add: rows and columns must match\n");
  if (!C) C = mzd_init_impl(A->nrows, A->ncols, 0);
  else if (C->nrows != A->nrows || C->ncols != A->ncols)
    die("m4ri_b200: mzd_add: rows and columns of returned matrix must match\n");
  if (A->nrows == 0 || A->width == 0) return C;
  cudaStream_t s = current_stream();
  DMat *dA = dmat_new(A->nrows, A->ncols, s), *dB = (B == A) ? dA : dmat_new(B->nrows, B->ncols, s);
  upload(dA, A, s);
  if (dB != dA) upload(dB, B, s);
  k_add(dA, dA, dB, s);
  download(C, dA, s);
  if (dB != dA) dmat_delete(dB, s);
  dmat_delete(dA, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return C;
}
mzd_t *mzd_sub(mzd_t *C, const mzd_t *A, const mzd_t *B) { return mzd_add(C, A, B); }

// DST = A^T on the GPU (reference mzd_transpose mzd.rs:148; on the mat-vec path
// twice per call, binary_matrix.rs:272-279,335,426).
mzd_t *mzd_transpose(mzd_t *DST, const mzd_t *A) {
  if (!DST) DST = mzd_init_impl(A->ncols, A->nrows, 0);
  else if (DST->nrows != A->ncols || DST->ncols != A->nrows)
    die("m4ri_b200: mzd_transpose: wrong size for return matrix\n");
  if (DST == A) die("m4ri_b200: mzd_transpose: in-place transpose is not supported\n");
  if (A->nrows == 0 || A->ncols == 0) return DST;
  cudaStream_t s = current_stream();
  DMat *dA = dmat_new(A->nrows, A->ncols, s), *dT = dmat_new(A->ncols, A->nrows, s);
  upload(dA, A, s);
  k_transpose(dT, dA, s);
  download(DST, dT, s);
  dmat_delete(dA, s);
  dmat_delete(dT, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return DST;
}

// ---- additive API -------------------------------------------------------------------
void m4ri_b200_set_algo(int algo) {
  std::call_once(g_once, probe);
  if (algo < 0 || algo > 3) die("m4ri_b200: unknown algorithm %d\n", algo);
  g_algo.store(algo);
}
int m4ri_b200_get_algo(void) {
  std::call_once(g_once, probe);
  return g_algo.load();
}
void m4ri_b200_set_strassen_levels(int levels) { g_strassen.store(levels); }
void m4ri_b200_pipeline_plan(int m, int k, int n, int clear, int *row_chunks, int *col_blocks) {
  plan_pipeline(m, k, n, clear, row_chunks, col_blocks);
}
void m4ri_b200_set_devices(int n) { g_devices.store(n < 1 ? 1 : n); }
int m4ri_b200_get_devices(void) { return multi_devices(); }
void m4ri_b200_set_sm_limit(int sms) { g_sm_limit.store(sms < 0 ? 0 : (sms & ~1)); }
void m4ri_b200_set_mma_variant(int variant) {
  if (variant < 0 || variant > 9) die("m4ri_b200: unknown MMA variant %d\n", variant);
  g_mma_variant.store(variant);
}

int m4ri_b200_device_count(void) {
  std::call_once(g_once, probe);
  return g_ndev;
}
int m4ri_b200_device_info(char *name, int name_len, int *sms, int *cc_major, int *cc_minor) {
  std::call_once(g_once, probe);
  if (g_ndev <= 0) return 0;
  if (name && name_len > 0) {
    strncpy(name, g_name, (size_t)name_len - 1);
    name[name_len - 1] = 0;
  }
  if (sms) *sms = g_sms;
  if (cc_major) *cc_major = g_cc_major;
  if (cc_minor) *cc_minor = g_cc_minor;
  return 1;
}
void m4ri_b200_set_stream(void *stream) {
  t_user_stream = (cudaStream_t)stream;
  t_use_user_stream = true;
}
void m4ri_b200_use_own_stream(void) {
  t_user_stream = nullptr;
  t_use_user_stream = false;
}
void m4ri_b200_synchronize(void) { M4B_CUDA(cudaStreamSynchronize(current_stream())); }
unsigned long long m4ri_b200_launch_count(void) { return g_launches.load(); }
const char *m4ri_b200_version(void) { return "m4ri_b200 0.1 (sm_100a)"; }

m4ri_b200_dmat_t *m4ri_b200_dmat_init(rci_t nrows, rci_t ncols) {
  return dmat_new(nrows, ncols, current_stream());
}
m4ri_b200_dmat_t *m4ri_b200_dmat_wrap(void *dptr, rci_t nrows, rci_t ncols, wi_t rowstride) {
  if (rowstride < width_of(ncols) || (rowstride & 1) || ((uintptr_t)dptr & 15))
    die("m4ri_b200: dmat_wrap: rowstride must be even and >= width, pointer 16-byte aligned\n");
  DMat *M = (DMat *)calloc(1, sizeof(DMat));
  if (!M) die("m4ri_b200: out of memory\n");
  M->d = (word *)dptr;
  M->nrows = nrows;
  M->ncols = ncols;
  M->width = width_of(ncols);
  M->rowstride = rowstride;
  M->base = nullptr;
  return M;
}
void m4ri_b200_dmat_free(m4ri_b200_dmat_t *M) {
  if (M) dmat_delete(M, M->base ? current_stream() : nullptr);
}
rci_t m4ri_b200_dmat_nrows(const m4ri_b200_dmat_t *M) { return M->nrows; }
rci_t m4ri_b200_dmat_ncols(const m4ri_b200_dmat_t *M) { return M->ncols; }
wi_t m4ri_b200_dmat_rowstride(const m4ri_b200_dmat_t *M) { return M->rowstride; }
void *m4ri_b200_dmat_ptr(const m4ri_b200_dmat_t *M) { return M->d; }

void m4ri_b200_upload(m4ri_b200_dmat_t *D, const mzd_t *H) { upload(D, H, current_stream()); }
void m4ri_b200_download(mzd_t *H, const m4ri_b200_dmat_t *D) { download(H, D, current_stream()); }

void m4ri_b200_dmul(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *B,
                    int clear, int algo) {
  dmul(C, A, B, clear, algo, current_stream());
}
void m4ri_b200_dmul_abt(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A,
                        const m4ri_b200_dmat_t *Bt, int clear, int algo) {
  dmul_abt(C, A, Bt, clear, algo, current_stream());
}
void m4ri_b200_dmul_v(m4ri_b200_dmat_t *y, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *x,
                      int clear) {
  if (x->nrows != 1 || x->ncols != A->ncols || y->nrows != 1 || y->ncols != A->nrows)
    die("m4ri_b200: dmul_v: x must be 1 x %d and y 1 x %d\n", A->ncols, A->nrows);
  k_mul_v(y, A, x, clear, current_stream());
}
void m4ri_b200_dmul_v_batched(m4ri_b200_dmat_t *Y, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *X,
                              int clear) {
  if (X->ncols != A->ncols || Y->nrows != X->nrows || Y->ncols != A->nrows)
    die("m4ri_b200: dmul_v_batched: X must be b x %d and Y b x %d\n", A->ncols, A->nrows);
  // Y = X * A^T: beyond 32 vectors the contraction is a product for the tensor cores
  if (X->nrows > 32) dmul_abt(Y, X, A, clear, M4RI_B200_ALGO_AUTO, current_stream());
  else k_mul_v_batched(Y, A, X, clear, current_stream());
}
void m4ri_b200_dmul_va(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *v, const m4ri_b200_dmat_t *A,
                       int clear) {
  if (v->ncols != A->nrows || C->nrows != v->nrows || C->ncols != A->ncols)
    die("m4ri_b200: dmul_va: shapes do not match\n");
  k_mul_va(C, v, A, clear, current_stream());
}
void m4ri_b200_dtranspose(m4ri_b200_dmat_t *DST, const m4ri_b200_dmat_t *A) {
  if (DST->nrows != A->ncols || DST->ncols != A->nrows)
    die("m4ri_b200: dtranspose: wrong size for return matrix\n");
  k_transpose(DST, A, current_stream());
}
void m4ri_b200_dtranspose_scatter(m4ri_b200_dmat_t *const *DST, int ndst, const m4ri_b200_dmat_t *A) {
  k_transpose_scatter(DST, ndst, A, current_stream());
}
void m4ri_b200_dadd(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A, const m4ri_b200_dmat_t *B) {
  if (A->nrows != B->nrows || A->ncols != B->ncols || C->nrows != A->nrows || C->ncols != A->ncols)
    die("m4ri_b200: dadd: rows and columns must match\n");
  k_add(C, A, B, current_stream());
}
void m4ri_b200_dcopy(m4ri_b200_dmat_t *DST, const m4ri_b200_dmat_t *A) {
  if (DST->nrows != A->nrows || DST->ncols != A->ncols) die("m4ri_b200: dcopy: shape mismatch\n");
  k_copy(DST, A, current_stream());
}
void m4ri_b200_dfill_random(m4ri_b200_dmat_t *M, uint64_t seed) {
  k_fill_random(M, seed, 0, current_stream());
}
void m4ri_b200_dfill_random_at(m4ri_b200_dmat_t *M, uint64_t seed, uint64_t first_index) {
  k_fill_random(M, seed, first_index, current_stream());
}
void m4ri_b200_timing_reset(void) {
  std::lock_guard<std::mutex> g(g_tmu);
  for (auto &p : g_tev) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
  g_tev.clear();
  g_tops = 0;
  g_timing_thread = std::this_thread::get_id();
  g_timing_on = true;
}
double m4ri_b200_timing_ms(void) {
  std::lock_guard<std::mutex> g(g_tmu);
  g_timing_on = false;  // the window ends here: later launches are not recorded (nor slowed down)
  double tot = 0;
  for (auto &p : g_tev) {
    M4B_CUDA(cudaEventSynchronize(p.second));
    float ms = 0;
    M4B_CUDA(cudaEventElapsedTime(&ms, p.first, p.second));
    tot += ms;
  }
  return tot;
}
int m4ri_b200_timing_count(void) {
  std::lock_guard<std::mutex> g(g_tmu);
  return (int)g_tev.size();
}
const char *m4ri_b200_timing_name(void) { return g_tname; }
double m4ri_b200_timing_ops(void) {
  std::lock_guard<std::mutex> g(g_tmu);
  return g_tops;
}
void m4ri_b200_dmul_abt_fp4_debug(m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A,
                                  const m4ri_b200_dmat_t *Bt, void *acc, int cg) {
  if (A->ncols != Bt->ncols || C->nrows != A->nrows || C->ncols != Bt->nrows)
    die("m4ri_b200: dmul_abt_fp4_debug: shapes do not match\n");
  k_mul_abt_mma4(C, A, Bt, 1, cg == 1 ? 1 : 2, (float *)acc, current_stream());
}
uint64_t m4ri_b200_dchecksum(const m4ri_b200_dmat_t *M) { return k_checksum(M, current_stream()); }

int m4ri_b200_dis_zero(const m4ri_b200_dmat_t *M) { return k_is_zero(M, current_stream()) ? 1 : 0; }
void m4ri_b200_dflip_bit(m4ri_b200_dmat_t *M, rci_t r, rci_t c) { k_flip_bit(M, r, c, current_stream()); }
int m4ri_b200_algo_available(int algo) {
  std::call_once(g_once, probe);
  if (g_ndev <= 0 || g_cc_major < 10) return 0;
  if (algo == M4RI_B200_ALGO_MMA) return mma_available() ? 1 : 0;
  return algo >= 0 && algo <= 3;
}

// Freivalds with 64 packed vectors; the three thin products use the AND+popc
// kernel so the check shares no code with the M4RM / MMA kernels it verifies.
int m4ri_b200_dfreivalds(const m4ri_b200_dmat_t *C, const m4ri_b200_dmat_t *A,
                         const m4ri_b200_dmat_t *B, uint64_t seed) {
  cudaStream_t s = current_stream();
  rci_t m = A->nrows, k = A->ncols, n = B->ncols;
  if (B->nrows != k || C->nrows != m || C->ncols != n) die("m4ri_b200: dfreivalds: shape mismatch\n");
  // X^T is 64 x n with row j of X = splitmix64 word j: generate X (n x 64) then transpose.
  DMat *X = dmat_new(n, 64, s), *Xt = dmat_new(64, n, s);
  k_fill_random(X, seed, 0, s);
  k_transpose(Xt, X, s);
  DMat *BX = dmat_new(k, 64, s), *BXt = dmat_new(64, k, s), *ABX = dmat_new(m, 64, s),
       *CX = dmat_new(m, 64, s);
  k_mul_abt_simple(BX, B, Xt, 1, s);
  k_transpose(BXt, BX, s);
  k_mul_abt_simple(ABX, A, BXt, 1, s);
  k_mul_abt_simple(CX, C, Xt, 1, s);
  k_add(CX, CX, ABX, s);
  bool chk_zero = k_is_zero(CX, s);
  dmat_delete(X, s); dmat_delete(Xt, s); dmat_delete(BX, s); dmat_delete(BXt, s);
  dmat_delete(ABX, s); dmat_delete(CX, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return (int)chk_zero;
}

}  // extern "C"
