// mzd_host.cpp — the allocator set of the C-ABI: host-side mzd_t management.
//
// These are the O(size) host-memory utilities m4ri-rust's BinMatrix needs around
// the multiplication path (SURVEY 8b "must also export — because one allocator
// must own every mzd_t the path touches"): construction, windows, copy, compare,
// identity and random fill of HOST matrices — descriptor and memcpy work, no GF(2)
// arithmetic. Everything that computes (the products, mzd_add, mzd_transpose:
// SURVEY 8a) is in api.cu and runs on the GPU.
//
// Layout contract (m4ri-sys/src/mzd.rs:24-79, checked by the reference tests at
// mzd.rs:381-422): 64-byte descriptor, rows[i] == blocks[0].begin + offset_vector
// + i*rowstride, rowstride = (width < 3 || width even) ? width : width + 1,
// excess bits and padding words are zero.
#include <string.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "common.cuh"

static_assert(sizeof(mzd_t) == 64, "mzd_t must be 64 bytes (m4ri-sys/src/mzd.rs:385)");
static_assert(offsetof(mzd_t, high_bitmask) == 40 && offsetof(mzd_t, blocks) == 48 &&
                  offsetof(mzd_t, rows) == 56 && offsetof(mzd_t, flags) == 24,
              "mzd_t field offsets (m4ri-sys/src/mzd.rs:24-79)");

namespace m4b {
// how a matrix's words were allocated, kept in mzd_t::padding[0]
enum : uint8_t { ALLOC_NONE = 0, ALLOC_MALLOC = 0xA1, ALLOC_PINNED = 0xA2 };
// padding[9] of every descriptor this library creates: mzd_free refuses anything else (a matrix
// allocated by another M4RI in the same process must be freed by that library — mixing
// allocators is unsupported)
constexpr uint8_t DESC_MAGIC = 0x5B;
// Matrices at least this large get page-locked memory so the H2D/D2H copies of a
// product run at full PCIe rate; small ones (the reference's tests create
// thousands of 10x10 matrices) stay in ordinary memory.
constexpr size_t PINNED_MIN_BYTES = 256 * 1024;

// ---- cache of freed blocks ---------------------------------------------------------
// m4ri-rust allocates the result of every product (`mzd_mul(NULL, a, b, 0)`,
// binary_matrix.rs:465) and frees it on drop: for a 128 MiB result cudaHostAlloc + cudaFreeHost
// cost several times the product itself. Freed blocks of PINNED_MIN_BYTES or more are kept
// (up to M4RI_B200_HOST_CACHE_MB, default 4096) and handed out again to requests of the same
// or a slightly smaller size. Applies to both kinds of block so that the logic is exercised on
// a CPU-only box as well.
struct CachedBlock {
  word *p;
  size_t cap;
  uint8_t kind;
};
static std::mutex g_cache_mu;
static std::vector<CachedBlock> g_cache;
static size_t g_cache_bytes = 0;
static std::atomic<unsigned long long> g_cache_hits{0}, g_cache_misses{0};

static size_t cache_limit() {
  static const size_t lim = [] {
    const char *e = getenv("M4RI_B200_HOST_CACHE_MB");
    long long mb = e ? atoll(e) : 4096;
    return (size_t)(mb < 0 ? 0 : mb) << 20;
  }();
  return lim;
}

static void release_block(word *p, uint8_t kind) {
  if (kind == ALLOC_PINNED) cudaFreeHost(p);
  else free(p);
}

// smallest cached block of `kind` with bytes <= cap <= bytes * 5/4, or nullptr
static word *cache_take(size_t bytes, uint8_t kind, size_t *cap) {
  std::lock_guard<std::mutex> g(g_cache_mu);
  int best = -1;
  for (int i = 0; i < (int)g_cache.size(); i++) {
    const CachedBlock &b = g_cache[i];
    if (b.kind == kind && b.cap >= bytes && b.cap <= bytes + bytes / 4 &&
        (best < 0 || b.cap < g_cache[best].cap))
      best = i;
  }
  if (best < 0) return nullptr;
  word *p = g_cache[best].p;
  *cap = g_cache[best].cap;
  g_cache_bytes -= g_cache[best].cap;
  g_cache[best] = g_cache.back();
  g_cache.pop_back();
  return p;
}

static bool cache_put(word *p, size_t cap, uint8_t kind) {
  if (cap < PINNED_MIN_BYTES) return false;
  std::lock_guard<std::mutex> g(g_cache_mu);
  if (g_cache_bytes + cap > cache_limit()) return false;
  g_cache.push_back({p, cap, kind});
  g_cache_bytes += cap;
  return true;
}

// zero == 0: the caller overwrites every valid word (a product's result); only the padding
// word of each row (rowstride > width) is cleared.
mzd_t *mzd_init_impl(rci_t r, rci_t c, int zero);
}  // namespace m4b

using namespace m4b;

extern "C" {

mzd_t *mzd_init(rci_t r, rci_t c) { return m4b::mzd_init_impl(r, c, 1); }

void m4ri_b200_host_cache_trim(void) {
  std::vector<CachedBlock> drop;
  {
    std::lock_guard<std::mutex> g(g_cache_mu);
    drop.swap(g_cache);
    g_cache_bytes = 0;
  }
  for (auto &b : drop) release_block(b.p, b.kind);
}

void m4ri_b200_host_cache_stats(unsigned long long *hits, unsigned long long *misses,
                                unsigned long long *bytes_held) {
  if (hits) *hits = g_cache_hits.load();
  if (misses) *misses = g_cache_misses.load();
  std::lock_guard<std::mutex> g(g_cache_mu);
  if (bytes_held) *bytes_held = g_cache_bytes;
}

}  // extern "C"

mzd_t *m4b::mzd_init_impl(rci_t r, rci_t c, int zero) {
  if (r < 0 || c < 0) die("m4ri_b200: mzd_init(%d, %d): negative dimension\n", r, c);
  mzd_t *A = (mzd_t *)calloc(1, sizeof(mzd_t));
  if (!A) die("m4ri_b200: out of memory\n");
  A->nrows = r;
  A->ncols = c;
  A->width = width_of(c);
  A->rowstride = (A->width < 3 || (A->width & 1) == 0) ? A->width : A->width + 1;  // mzd.rs:35-37
  A->high_bitmask = high_mask(c);
  A->flags = (c % 64) ? mzd_flag_nonzero_excess : 0;
  A->offset_vector = 0;
  A->row_offset = 0;
  int lg = 0;  // single block: blockrows = next power of two >= nrows
  while (((rci_t)1 << lg) < (r > 0 ? r : 1)) lg++;
  A->blockrows_log = (uint8_t)lg;

  size_t words = (size_t)r * (size_t)A->rowstride;
  size_t bytes = words * sizeof(word);
  A->blocks = (mzd_block_t *)calloc(2, sizeof(mzd_block_t));  // [1] is the {0,NULL,NULL} terminator
  A->rows = (word **)calloc((size_t)(r > 0 ? r : 1), sizeof(word *));
  if (!A->blocks || !A->rows) die("m4ri_b200: out of memory\n");
  word *data = nullptr;
  uint8_t kind = ALLOC_MALLOC;
  size_t cap = bytes ? ((bytes + 63) & ~(size_t)63) : 64;  // capacity of the block actually held
  bool fresh = true;
  const bool big = bytes >= PINNED_MIN_BYTES;
  if (big && m4ri_b200_device_count() > 0) {
    size_t got = 0;
    if ((data = cache_take(bytes, ALLOC_PINNED, &got))) {
      kind = ALLOC_PINNED, cap = got, fresh = false;
    } else if (cudaHostAlloc((void **)&data, cap, cudaHostAllocPortable) == cudaSuccess) {
      kind = ALLOC_PINNED;
    } else {
      (void)cudaGetLastError();
      data = nullptr;
    }
  }
  if (!data) {
    size_t got = 0;
    if (big && (data = cache_take(bytes, ALLOC_MALLOC, &got))) {
      cap = got, fresh = false;
    } else {
      data = (word *)aligned_alloc(64, cap);
      if (!data) die("m4ri_b200: out of memory (%zu bytes)\n", bytes);
    }
  }
  if (big) (fresh ? g_cache_misses : g_cache_hits).fetch_add(1);
  if (zero) {
    memset(data, 0, bytes ? bytes : cap);
  } else if (A->rowstride > A->width) {
    for (rci_t i = 0; i < r; i++) data[(size_t)i * A->rowstride + A->width] = 0;
  }
  A->padding[0] = kind;
  // capacity of the block (>= the bytes in use) in padding[1..8]: mzd_free returns it to the cache
  memcpy(&A->padding[1], &cap, sizeof(cap));
  A->padding[9] = DESC_MAGIC;
  A->blocks[0].size = bytes;
  A->blocks[0].begin = data;
  A->blocks[0].end = data + words;
  for (rci_t i = 0; i < r; i++) A->rows[i] = data + (size_t)i * A->rowstride;
  return A;
}

extern "C" {

void mzd_free(mzd_t *A) {
  if (!A) return;
  if (A->padding[9] != DESC_MAGIC ||
      (A->padding[0] != ALLOC_NONE && A->padding[0] != ALLOC_MALLOC && A->padding[0] != ALLOC_PINNED))
    die("m4ri_b200: mzd_free: this mzd_t was not created by libm4ri_b200 (mixing M4RI allocators in "
        "one process is unsupported)\n");
  if (A->blocks && A->padding[0] != ALLOC_NONE) {
    size_t cap = 0;
    memcpy(&cap, &A->padding[1], sizeof(cap));
    if (!cache_put(A->blocks[0].begin, cap, A->padding[0])) release_block(A->blocks[0].begin, A->padding[0]);
    free(A->blocks);
  }
  free(A->rows);
  free(A);
}

// Window = view sharing the parent's words (mzd.rs:121-127); lowc % 64 == 0.
mzd_t *mzd_init_window(mzd_t *M, rci_t lowr, rci_t lowc, rci_t highr, rci_t highc) {
  if (lowc % 64) die("m4ri_b200: mzd_init_window: lowc must be a multiple of 64\n");
  if (lowr < 0 || highr > M->nrows || lowr > highr || lowc < 0 || highc > M->ncols || lowc > highc)
    die("m4ri_b200: mzd_init_window: window out of range\n");
  mzd_t *W = (mzd_t *)calloc(1, sizeof(mzd_t));
  if (!W) die("m4ri_b200: out of memory\n");
  W->nrows = highr - lowr;
  W->ncols = highc - lowc;
  W->width = width_of(W->ncols);
  W->rowstride = M->rowstride;
  W->high_bitmask = high_mask(W->ncols);
  W->flags = mzd_flag_windowed_zerooffset |
             ((W->ncols % 64) ? mzd_flag_nonzero_excess : mzd_flag_windowed_zeroexcess);
  W->blockrows_log = M->blockrows_log;
  W->row_offset = M->row_offset + lowr;
  W->offset_vector = M->offset_vector + lowr * M->rowstride + lowc / 64;
  W->blocks = M->blocks;  // shared, not owned
  W->padding[0] = ALLOC_NONE;
  W->padding[9] = DESC_MAGIC;
  W->rows = (word **)calloc((size_t)(W->nrows > 0 ? W->nrows : 1), sizeof(word *));
  if (!W->rows) die("m4ri_b200: out of memory\n");
  for (rci_t i = 0; i < W->nrows; i++) W->rows[i] = M->rows[lowr + i] + lowc / 64;
  return W;
}

// bits [0, ncols) of src appended to dst starting at bit column c0 (dst bits there are zero)
static void append_bits(word *dst, rci_t c0, const word *src, rci_t ncols, word src_mask) {
  const wi_t sw = (ncols + 63) / 64;
  const int sh = c0 % 64;
  word *d = dst + c0 / 64;
  for (wi_t w = 0; w < sw; w++) {
    const word v = (w == sw - 1) ? (src[w] & src_mask) : src[w];
    d[w] |= v << sh;
    if (sh && (w < sw - 1 || (ncols - 1) % 64 + sh >= 64)) d[w + 1] |= v >> (64 - sh);
  }
}

// C = [A B] (mzd.rs:195; BinMatrix::augmented, binary_matrix.rs:221-231). O(size) host copy.
mzd_t *mzd_concat(mzd_t *C, const mzd_t *A, const mzd_t *B) {
  if (A->nrows != B->nrows) die("m4ri_b200: mzd_concat: bad arguments to concat\n");
  if (!C) C = mzd_init(A->nrows, A->ncols + B->ncols);
  else if (C->nrows != A->nrows || C->ncols != A->ncols + B->ncols)
    die("m4ri_b200: mzd_concat: C has wrong dimension\n");
  for (rci_t i = 0; i < A->nrows; i++) {
    word *dst = C->rows[i];
    if (!C->width) continue;
    // C may be a window whose last word also holds columns of its parent: those bits stay
    const word foreign = dst[C->width - 1] & ~C->high_bitmask;
    memset(dst, 0, sizeof(word) * (size_t)C->width);
    if (A->width) {
      memcpy(dst, A->rows[i], sizeof(word) * (size_t)A->width);
      dst[A->width - 1] &= A->high_bitmask;
    }
    if (B->width) append_bits(dst, A->ncols, B->rows[i], B->ncols, B->high_bitmask);
    dst[C->width - 1] |= foreign;
  }
  return C;
}

// C = [A; B] (mzd.rs:201; BinMatrix::stacked, binary_matrix.rs:234-244). O(size) host copy.
mzd_t *mzd_stack(mzd_t *C, const mzd_t *A, const mzd_t *B) {
  if (A->ncols != B->ncols) die("m4ri_b200: mzd_stack: A->ncols != B->ncols!\n");
  if (!C) C = mzd_init(A->nrows + B->nrows, A->ncols);
  else if (C->nrows != A->nrows + B->nrows || C->ncols != A->ncols)
    die("m4ri_b200: mzd_stack: C has wrong dimension!\n");
  for (rci_t i = 0; i < C->nrows; i++) {
    const mzd_t *S = i < A->nrows ? A : B;
    const word *src = S->rows[i < A->nrows ? i : i - A->nrows];
    if (!C->width) continue;
    word *dst = C->rows[i];
    const wi_t l = C->width - 1;
    const word last = (dst[l] & ~C->high_bitmask) | (src[l] & C->high_bitmask);  // a window's parent keeps its bits
    memcpy(dst, src, sizeof(word) * (size_t)l);
    dst[l] = last;
  }
  return C;
}

mzd_t *mzd_copy(mzd_t *N, const mzd_t *P) {
  if (N == P) return N;
  if (!N) N = mzd_init(P->nrows, P->ncols);
  else if (N->nrows < P->nrows || N->ncols < P->ncols)
    die("m4ri_b200: mzd_copy: target matrix is too small\n");
  if (P->width == 0) return N;
  word hm = P->high_bitmask;
  for (rci_t i = 0; i < P->nrows; i++) {
    const word *src = P->rows[i];
    word *dst = N->rows[i];
    memcpy(dst, src, sizeof(word) * (size_t)(P->width - 1));
    // keep the target's bits beyond P's last column (it may be wider)
    dst[P->width - 1] = (dst[P->width - 1] & ~hm) | (src[P->width - 1] & hm);
  }
  return N;
}

int mzd_equal(const mzd_t *A, const mzd_t *B) {
  if (A->nrows != B->nrows || A->ncols != B->ncols) return 0;
  if (A == B || A->width == 0) return 1;
  word hm = A->high_bitmask;
  for (rci_t i = 0; i < A->nrows; i++) {
    const word *a = A->rows[i], *b = B->rows[i];
    if (memcmp(a, b, sizeof(word) * (size_t)(A->width - 1)) != 0) return 0;
    if ((a[A->width - 1] ^ b[A->width - 1]) & hm) return 0;
  }
  return 1;
}

// value odd: identity on the main diagonal; value even: zero matrix (mzd.rs:198).
void mzd_set_ui(mzd_t *A, unsigned int value) {
  for (rci_t i = 0; i < A->nrows; i++) memset(A->rows[i], 0, sizeof(word) * (size_t)A->width);
  if (value & 1) {
    rci_t d = A->nrows < A->ncols ? A->nrows : A->ncols;
    for (rci_t i = 0; i < d; i++) A->rows[i][i / 64] |= m4ri_one << (i % 64);
  }
}

// Uniform bits, NOT stream-compatible with M4RI's generator (which the reference
// never relies on: SURVEY 8c). splitmix64 over a process-wide counter.
static std::atomic<uint64_t> g_rand_state{0x4D345249B200ULL};
void mzd_randomize(mzd_t *A) {
  if (A->width == 0) return;
  uint64_t words = (uint64_t)A->nrows * (uint64_t)A->width;
  uint64_t base = g_rand_state.fetch_add(words * 0x9E3779B97F4A7C15ULL);
  for (rci_t i = 0; i < A->nrows; i++) {
    word *row = A->rows[i];
    for (wi_t w = 0; w < A->width; w++) {
      uint64_t z = (base += 0x9E3779B97F4A7C15ULL);
      z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
      z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
      row[w] = z ^ (z >> 31);
    }
    row[A->width - 1] &= A->high_bitmask;
  }
}

// ---- permutations (m4ri-sys/src/mzp.rs:14-51): M4RI's published mzp_t, host bookkeeping ----------
// A window is its own small descriptor sharing the parent's values: it is released with
// mzp_free_window, never mzp_free.
mzp_t *mzp_init(rci_t length) {
  if (length < 0) die("m4ri_b200: mzp_init: negative length\n");
  mzp_t *P = (mzp_t *)malloc(sizeof(mzp_t));
  if (!P) die("m4ri_b200: mzp_init: out of memory\n");
  P->values = (rci_t *)malloc(sizeof(rci_t) * (size_t)(length > 0 ? length : 1));
  if (!P->values) die("m4ri_b200: mzp_init: out of memory\n");
  P->length = length;
  for (rci_t i = 0; i < length; i++) P->values[i] = i;
  return P;
}
void mzp_free(mzp_t *P) {
  if (!P) return;
  free(P->values);
  free(P);
}
mzp_t *mzp_init_window(mzp_t *P, rci_t begin, rci_t end) {
  if (begin < 0 || end < begin || end > P->length) die("m4ri_b200: mzp_init_window: bad range\n");
  mzp_t *W = (mzp_t *)malloc(sizeof(mzp_t));
  if (!W) die("m4ri_b200: mzp_init_window: out of memory\n");
  W->values = P->values + begin;
  W->length = end - begin;
  return W;
}
void mzp_free_window(mzp_t *condemned) { free(condemned); }
void Mzp_free_window(mzp_t *condemned) { free(condemned); }  // the spelling mzp.rs:29 declares
mzp_t *mzp_copy(mzp_t *P, const mzp_t *Q) {
  if (!P) P = mzp_init(Q->length);
  else if (P->length < Q->length) die("m4ri_b200: mzp_copy: target is too short\n");
  for (rci_t i = 0; i < Q->length; i++) P->values[i] = Q->values[i];
  return P;
}
void mzp_set_ui(mzp_t *P, unsigned int value) {
  (void)value;
  for (rci_t i = 0; i < P->length; i++) P->values[i] = i;
}
void mzp_print(const mzp_t *P) {
  printf("[ ");
  for (rci_t i = 0; i < P->length; i++) printf("%d ", (int)P->values[i]);
  printf("]");
}

// ---- row / column utilities m4ri-sys also binds (mzd.rs:130,141,144,205,233,240): host word work ----
void mzd_row_swap(mzd_t *M, rci_t rowa, rci_t rowb) {
  if (rowa == rowb || M->width == 0) return;
  word *a = M->rows[rowa], *b = M->rows[rowb];
  const word hm = M->high_bitmask;
  for (wi_t w = 0; w + 1 < M->width; w++) {
    const word t = a[w];
    a[w] = b[w];
    b[w] = t;
  }
  // the last word may be shared with a parent's other columns (windows): swap the valid bits only
  const wi_t l = M->width - 1;
  const word t = (a[l] ^ b[l]) & hm;
  a[l] ^= t;
  b[l] ^= t;
}
void mzd_copy_row(mzd_t *B, rci_t i, const mzd_t *A, rci_t j) {
  if (A->ncols > B->ncols) die("m4ri_b200: mzd_copy_row: source has more columns than the target\n");
  if (A->width == 0) return;
  const word *src = A->rows[j];
  word *dst = B->rows[i];
  for (wi_t w = 0; w + 1 < A->width; w++) dst[w] = src[w];
  const wi_t l = A->width - 1;
  dst[l] = (dst[l] & ~A->high_bitmask) | (src[l] & A->high_bitmask);
}
void mzd_col_swap(mzd_t *M, rci_t cola, rci_t colb) {
  if (cola == colb) return;
  const wi_t wa = cola / 64, wb = colb / 64;
  const int ba = cola % 64, bb = colb % 64;
  for (rci_t i = 0; i < M->nrows; i++) {
    word *row = M->rows[i];
    const word x = ((row[wa] >> ba) ^ (row[wb] >> bb)) & 1;
    row[wa] ^= x << ba;
    row[wb] ^= x << bb;
  }
}
int mzd_is_zero(const mzd_t *A) {
  if (A->width == 0) return 1;
  for (rci_t i = 0; i < A->nrows; i++) {
    const word *row = A->rows[i];
    word acc = 0;
    for (wi_t w = 0; w + 1 < A->width; w++) acc |= row[w];
    acc |= row[A->width - 1] & A->high_bitmask;
    if (acc) return 0;
  }
  return 1;
}
void mzd_row_clear_offset(mzd_t *M, rci_t row, rci_t coloffset) {
  if (coloffset >= M->ncols || M->width == 0) return;
  word *r = M->rows[row];
  const wi_t w0 = coloffset / 64;
  const word keep_low = (coloffset % 64) ? ((m4ri_one << (coloffset % 64)) - 1) : 0;
  for (wi_t w = w0; w < M->width; w++) {
    word clear = m4ri_ffff;
    if (w == w0) clear &= ~keep_low;
    if (w == M->width - 1) clear &= M->high_bitmask;
    r[w] &= ~clear;
  }
}
// S = M[lowr .. highr) x [lowc .. highc) as a copy (S NULL: allocate); any lowc.
mzd_t *mzd_submatrix(mzd_t *S, const mzd_t *M, rci_t lowr, rci_t lowc, rci_t highr, rci_t highc) {
  if (lowr < 0 || lowc < 0 || highr < lowr || highc < lowc || highr > M->nrows || highc > M->ncols)
    die("m4ri_b200: mzd_submatrix: window out of range\n");
  const rci_t nr = highr - lowr, nc = highc - lowc;
  if (!S) S = mzd_init(nr, nc);
  else if (S->nrows < nr || S->ncols < nc) die("m4ri_b200: mzd_submatrix: got S with dimension %d x %d but expected %d x %d\n", S->nrows, S->ncols, nr, nc);
  if (nc == 0) return S;
  const wi_t w0 = lowc / 64, nw = width_of(nc);
  const int sh = lowc % 64;
  const word hm = high_mask(nc);
  for (rci_t i = 0; i < nr; i++) {
    const word *src = M->rows[lowr + i];
    word *dst = S->rows[i];
    for (wi_t w = 0; w < nw; w++) {
      word v = src[w0 + w] >> sh;
      if (sh && w0 + w + 1 < M->width) v |= src[w0 + w + 1] << (64 - sh);
      if (w == nw - 1) dst[w] = (dst[w] & ~hm) | (v & hm);
      else dst[w] = v;
    }
  }
  return S;
}

}  // extern "C"
