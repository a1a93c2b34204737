/*
 * gf2_oracle.c — CPU restatement of the GF(2) multiplication path of M4RI as
 * m4ri-rust uses it. TEST INFRASTRUCTURE ONLY: it may be imported, linked or
 * executed only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs, and only as the checker or the timed CPU baseline.
 * The product (libm4ri_b200.so) never links or calls anything in this file.
 *
 * PARITY STATUS: "parity unpinned" for non-identity products. The arithmetic the
 * reference runs lives in a third-party C library that is absent from
 * /root/reference: M4RI, fork bitbucket.org/thomwiggers/m4ri pinned at commit
 * 5cd604d29068651c23265e94ccd785a821c3dfe6 (.gitmodules:1-3; the vendor directory
 * m4ri-sys/vendor/m4ri is empty), and the reference's own tests hold only
 * identity products (m4ri-rust/src/friendly/binary_matrix.rs:661-686), which
 * tests/test_oracle.py replays against this file. What makes parity well defined
 * anyway: a product over GF(2) is a unique mathematical object (no rounding, no
 * order dependence, independent of M4RI's `cutoff`/`k`), so this file restates
 * the published semantics twice — from the definition (oracle_mul_naive) and as
 * the published Method of the Four Russians (oracle_mul_m4rm: Gray-code tables
 * of 2^8 row combinations per 8-row slab of B, one table lookup + XOR per row of
 * A; mechanism documented at m4ri-sys/src/brilliantrussian.rs:17-193) — and the
 * tests prove the two agree. oracle_mul_strassen adds the strategy of the
 * reference's default entry point (Strassen-Winograd above a cutoff, Four Russians
 * at the leaves; m4ri-sys/src/strassen.rs:5-18), checked against both.
 *
 * Semantics followed (all file:line relative to /root/reference):
 *   bit order  (r,c) = (row[c/64] >> (c%64)) & 1      m4ri-sys/src/mzd.rs:263-269
 *   width / rowstride / high_bitmask                    m4ri-sys/src/mzd.rs:26-39,73-74
 *   C = A*B, C ^= A*B                                   m4ri-sys/src/strassen.rs:18,31
 *                                                       m4ri-sys/src/brilliantrussian.rs:216,224
 *   C (^)= A*Bt^T with `clear`                          m4ri-sys/src/mzd.rs:154-168
 *   C (^)= v*A with `clear`                             m4ri-sys/src/mzd.rs:175-181
 *   transpose, add                                      m4ri-sys/src/mzd.rs:148,223
 *   PLE / PLUQ decomposition (oracle_ple)               m4ri-sys/src/ple.rs:35,65
 *   echelon form / rank, inverse, solve (oracle_rref)   m4ri-sys/src/echelonform.rs:16,
 *                                                       brilliantrussian.rs:208, solve.rs:23
 *
 * Matrices are plain arrays of 64-bit words, row-major, `stride` words per row,
 * `ncols` valid bits per row; bits beyond ncols are kept zero on output.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef uint64_t word;

static inline int width_of(int ncols) { return (ncols + 63) / 64; }
static inline word high_mask(int ncols) {
  int r = ncols % 64;
  return r ? (((word)1 << r) - 1) : ~(word)0;
}

/* ---- deterministic input generator ---------- */
static inline word splitmix64_at(uint64_t seed, uint64_t index) {
  /* the index-th output (0-based) of splitmix64 started at `seed` */
  word z = seed + (index + 1) * 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

/* word (r,w) of the matrix = splitmix64 output number r*width + w; last word masked */
void oracle_fill_random(word *M, int nrows, int ncols, int stride, uint64_t seed) {
  int width = width_of(ncols);
  word hm = high_mask(ncols);
  for (int r = 0; r < nrows; r++) {
    word *row = M + (size_t)r * stride;
    for (int w = 0; w < width; w++) row[w] = splitmix64_at(seed, (uint64_t)r * width + w);
    if (width) row[width - 1] &= hm;
    for (int w = width; w < stride; w++) row[w] = 0;
  }
}

/* position-mixed XOR fold over the valid words; padding words are ignored */
uint64_t oracle_checksum(const word *M, int nrows, int ncols, int stride) {
  int width = width_of(ncols);
  uint64_t acc = 0;
  for (int r = 0; r < nrows; r++)
    for (int w = 0; w < width; w++) {
      word z = M[(size_t)r * stride + w] + ((uint64_t)r * width + w + 1) * 0x9E3779B97F4A7C15ULL;
      z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
      z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
      acc ^= z ^ (z >> 31);
    }
  return acc;
}

static inline int get_bit(const word *M, int stride, int r, int c) {
  return (int)((M[(size_t)r * stride + c / 64] >> (c % 64)) & 1);
}

static void clear_or_mask(word *C, int cs, int m, int n, int clear) {
  int width = width_of(n);
  word hm = high_mask(n);
  for (int i = 0; i < m; i++) {
    word *row = C + (size_t)i * cs;
    if (clear) memset(row, 0, sizeof(word) * (size_t)width);
    else if (width) row[width - 1] &= hm;
  }
}

/* ---- from the definition: C[i,j] (^)= XOR_l A[i,l] & B[l,j] ------------------ */
void oracle_mul_naive(word *C, int cs, const word *A, int as, const word *B, int bs, int m, int k,
                      int n, int clear) {
  clear_or_mask(C, cs, m, n, clear);
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      int acc = 0;
      for (int l = 0; l < k; l++) acc ^= get_bit(A, as, i, l) & get_bit(B, bs, l, j);
      if (acc) C[(size_t)i * cs + j / 64] ^= (word)1 << (j % 64);
    }
}

/* ---- Method of the Four Russians, t = 8, four tables (32 bits of k) per pass --
 * Column blocks of C are independent, so threads split the columns: every
 * thread builds the tables of its own block and nothing is computed twice.   */
#define ORACLE_CB 16 /* words per column block: 1024 columns */
void oracle_mul_m4rm(word *C, int cs, const word *A, int as, const word *B, int bs, int m, int k,
                     int n, int clear, int nthreads) {
  clear_or_mask(C, cs, m, n, clear);
  int width = width_of(n);
  int nblocks = (width + ORACLE_CB - 1) / ORACLE_CB;
  word hm = high_mask(n);
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
  nthreads = 1;
#endif
#pragma omp parallel num_threads(nthreads)
  {
    word(*T)[256][ORACLE_CB] = aligned_alloc(64, sizeof(word) * 4 * 256 * ORACLE_CB);
#pragma omp for schedule(dynamic, 1)
    for (int cb = 0; cb < nblocks; cb++) {
      int w0 = cb * ORACLE_CB;
      int wn = width - w0 < ORACLE_CB ? width - w0 : ORACLE_CB;
      for (int k0 = 0; k0 < k; k0 += 32) {
        int kb = k - k0 < 32 ? k - k0 : 32;
        int ntab = (kb + 7) / 8;
        for (int t = 0; t < ntab; t++) {
          /* entry e = XOR of the rows k0+8t+b of B for the set bits b of e, built
           * from the entry with the lowest set bit removed (one XOR per entry) */
          memset(T[t][0], 0, sizeof(word) * ORACLE_CB);
          for (int e = 1; e < 256; e++) {
            int b = __builtin_ctz(e);
            int row = k0 + 8 * t + b;
            const word *prev = T[t][e & (e - 1)];
            if (row < k) {
              const word *brow = B + (size_t)row * bs + w0;
              for (int w = 0; w < wn; w++) T[t][e][w] = prev[w] ^ brow[w];
            } else {
              for (int w = 0; w < wn; w++) T[t][e][w] = prev[w];
            }
          }
        }
        word kmask = kb == 32 ? 0xffffffffULL : (((word)1 << kb) - 1);
        for (int i = 0; i < m; i++) {
          word a = (A[(size_t)i * as + k0 / 64] >> (k0 % 64)) & kmask;
          if (!a) continue;
          word *crow = C + (size_t)i * cs + w0;
          const word *t0 = T[0][a & 255];
          if (ntab == 4) {
            const word *t1 = T[1][(a >> 8) & 255], *t2 = T[2][(a >> 16) & 255],
                       *t3 = T[3][(a >> 24) & 255];
            for (int w = 0; w < wn; w++) crow[w] ^= t0[w] ^ t1[w] ^ t2[w] ^ t3[w];
          } else {
            for (int w = 0; w < wn; w++) crow[w] ^= t0[w];
            for (int t = 1; t < ntab; t++) {
              const word *tt = T[t][(a >> (8 * t)) & 255];
              for (int w = 0; w < wn; w++) crow[w] ^= tt[w];
            }
          }
        }
      }
      if (w0 + wn == width)
        for (int i = 0; i < m; i++) C[(size_t)i * cs + width - 1] &= hm;
    }
    free(T);
  }
}

/* ---- Strassen-Winograd above a cutoff, Four Russians at the leaves ------------
 * What mzd_mul (the reference's default strategy, m4ri-rust/src/friendly/
 * binary_matrix.rs:63-74 -> m4ri-sys/src/strassen.rs:5-18) is documented to do:
 * "Strassen-Winograd ... cutoff: minimal dimension for Strassen recursion". Seven
 * half-size products and fifteen half-size XORs per level (over GF(2) every
 * subtraction is an XOR):
 *   S1=A21+A22 S2=S1+A11 S3=A11+A21 S4=A12+S2   T1=B12+B11 T2=B22+T1 T3=B22+B12 T4=T2+B21
 *   P1=A11*B11 P2=A12*B21 P3=S4*B22 P4=A22*T4 P5=S1*T1 P6=S2*T2 P7=S3*T3
 *   C11=P1+P2  U2=P1+P6  U3=U2+P7  U4=U2+P5  C12=U4+P3  C21=U3+P4  C22=U3+P5
 * Halves are multiples of 64 columns (whole words); what does not fit the even core
 * (odd tails) is peeled off and multiplied by the Four-Russians routine. The seven
 * products of a level are OpenMP tasks; leaves run single-threaded.               */
static void xor_rows(word *D, int ds, const word *X, int xs, const word *Y, int ys, int rows, int words) {
  for (int i = 0; i < rows; i++) {
    word *d = D + (size_t)i * ds;
    const word *x = X + (size_t)i * xs, *y = Y + (size_t)i * ys;
    for (int w = 0; w < words; w++) d[w] = x[w] ^ y[w];
  }
}

/* C = A*B, every dimension a multiple of 64; C is overwritten */
static void strassen_rec(word *C, int cs, const word *A, int as, const word *B, int bs, int m, int k,
                         int n, int cutoff) {
  if (m < 2 * cutoff || k < 2 * cutoff || n < 2 * cutoff || (m % 128) || (k % 128) || (n % 128)) {
    oracle_mul_m4rm(C, cs, A, as, B, bs, m, k, n, 1, 1);
    return;
  }
  const int m2 = m / 2, k2 = k / 2, n2 = n / 2, kw = k2 / 64, nw = n2 / 64;
  const word *A11 = A, *A12 = A + kw, *A21 = A + (size_t)m2 * as, *A22 = A21 + kw;
  const word *B11 = B, *B12 = B + nw, *B21 = B + (size_t)k2 * bs, *B22 = B21 + nw;
  word *C11 = C, *C12 = C + nw, *C21 = C + (size_t)m2 * cs, *C22 = C21 + nw;
  word *S = malloc(sizeof(word) * 4 * (size_t)m2 * kw), *T = malloc(sizeof(word) * 4 * (size_t)k2 * nw),
       *P = malloc(sizeof(word) * 5 * (size_t)m2 * nw);
  if (!S || !T || !P) abort();
  word *S1 = S, *S2 = S1 + (size_t)m2 * kw, *S3 = S2 + (size_t)m2 * kw, *S4 = S3 + (size_t)m2 * kw;
  word *T1 = T, *T2 = T1 + (size_t)k2 * nw, *T3 = T2 + (size_t)k2 * nw, *T4 = T3 + (size_t)k2 * nw;
  word *P3 = P, *P4 = P3 + (size_t)m2 * nw, *P5 = P4 + (size_t)m2 * nw, *P6 = P5 + (size_t)m2 * nw,
       *P7 = P6 + (size_t)m2 * nw;
  xor_rows(S1, kw, A21, as, A22, as, m2, kw);
  xor_rows(S2, kw, S1, kw, A11, as, m2, kw);
  xor_rows(S3, kw, A11, as, A21, as, m2, kw);
  xor_rows(S4, kw, A12, as, S2, kw, m2, kw);
  xor_rows(T1, nw, B12, bs, B11, bs, k2, nw);
  xor_rows(T2, nw, B22, bs, T1, nw, k2, nw);
  xor_rows(T3, nw, B22, bs, B12, bs, k2, nw);
  xor_rows(T4, nw, T2, nw, B21, bs, k2, nw);
  /* P1 -> C11, P2 -> C12 (scratch until C12 is assembled), the rest into P */
#pragma omp task
  strassen_rec(C11, cs, A11, as, B11, bs, m2, k2, n2, cutoff);
#pragma omp task
  strassen_rec(C12, cs, A12, as, B21, bs, m2, k2, n2, cutoff);
#pragma omp task
  strassen_rec(P3, nw, S4, kw, B22, bs, m2, k2, n2, cutoff);
#pragma omp task
  strassen_rec(P4, nw, A22, as, T4, nw, m2, k2, n2, cutoff);
#pragma omp task
  strassen_rec(P5, nw, S1, kw, T1, nw, m2, k2, n2, cutoff);
#pragma omp task
  strassen_rec(P6, nw, S2, kw, T2, nw, m2, k2, n2, cutoff);
#pragma omp task
  strassen_rec(P7, nw, S3, kw, T3, nw, m2, k2, n2, cutoff);
#pragma omp taskwait
  xor_rows(P6, nw, P6, nw, C11, cs, m2, nw);  /* U2 = P1 + P6 */
  xor_rows(C11, cs, C11, cs, C12, cs, m2, nw); /* C11 = P1 + P2 */
  xor_rows(P7, nw, P7, nw, P6, nw, m2, nw);   /* U3 = U2 + P7 */
  xor_rows(P6, nw, P6, nw, P5, nw, m2, nw);   /* U4 = U2 + P5 */
  xor_rows(C12, cs, P6, nw, P3, nw, m2, nw);  /* C12 = U4 + P3 */
  xor_rows(C21, cs, P7, nw, P4, nw, m2, nw);  /* C21 = U3 + P4 */
  xor_rows(C22, cs, P7, nw, P5, nw, m2, nw);  /* C22 = U3 + P5 */
  free(S);
  free(T);
  free(P);
}

void oracle_mul_strassen(word *C, int cs, const word *A, int as, const word *B, int bs, int m, int k,
                         int n, int clear, int cutoff, int nthreads) {
  if (cutoff <= 0) cutoff = 2048; /* fastest leaf size measured on the build host (1024-4096 within 10 %) */
  if (cutoff < 64) cutoff = 64;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
  nthreads = 1;
#endif
  /* even core: the largest multiples of 128 that fit; tails go to the Four Russians */
  const int mc = m / 128 * 128, kc = k / 128 * 128, nc = n / 128 * 128;
  if (mc < 2 * cutoff || kc < 2 * cutoff || nc < 2 * cutoff) {
    oracle_mul_m4rm(C, cs, A, as, B, bs, m, k, n, clear, nthreads);
    return;
  }
  const int ncw = nc / 64;
  word *D = C;
  int ds = cs;
  if (!clear) { /* C ^= A*B: core product into a temporary, XORed in afterwards */
    D = malloc(sizeof(word) * (size_t)mc * ncw);
    if (!D) abort();
    ds = ncw;
  }
#pragma omp parallel num_threads(nthreads)
#pragma omp single
  strassen_rec(D, ds, A, as, B, bs, mc, kc, nc, cutoff);
  if (!clear) {
    xor_rows(C, cs, C, cs, D, ds, mc, ncw);
    free(D);
  }
  /* peel: k tail into the core block, then the column tail (all rows, all of k) and the
   * row tail (core columns only); column windows start on whole words */
  if (k > kc) oracle_mul_m4rm(C, cs, A + kc / 64, as, B + (size_t)kc * bs, bs, mc, k - kc, nc, 0, nthreads);
  if (n > nc) oracle_mul_m4rm(C + ncw, cs, A, as, B + ncw, bs, m, k, n - nc, clear, nthreads);
  if (m > mc) oracle_mul_m4rm(C + (size_t)mc * cs, cs, A + (size_t)mc * as, as, B, bs, m - mc, k, nc, clear, nthreads);
}

/* ---- C (^)= A * Bt^T, both operands k-contiguous (mzd.rs:154-168) ------------ */
void oracle_mul_abt(word *C, int cs, const word *A, int as, const word *Bt, int bts, int m, int k,
                    int n, int clear) {
  clear_or_mask(C, cs, m, n, clear);
  int kw = width_of(k);
  word km = high_mask(k);
  for (int i = 0; i < m; i++)
    for (int j = 0; j < n; j++) {
      word acc = 0;
      for (int w = 0; w < kw; w++) {
        word x = A[(size_t)i * as + w] & Bt[(size_t)j * bts + w];
        if (w == kw - 1) x &= km;
        acc ^= x;
      }
      if (__builtin_parityll(acc)) C[(size_t)i * cs + j / 64] ^= (word)1 << (j % 64);
    }
}

/* ---- C (^)= v * A, v is r x k with few rows (mzd.rs:175-181) ----------------- */
void oracle_mul_va(word *C, int cs, const word *v, int vs, const word *A, int as, int r, int k,
                   int n, int clear) {
  clear_or_mask(C, cs, r, n, clear);
  int width = width_of(n);
  word hm = high_mask(n);
  for (int i = 0; i < r; i++) {
    word *crow = C + (size_t)i * cs;
    for (int l = 0; l < k; l++)
      if (get_bit(v, vs, i, l)) {
        const word *arow = A + (size_t)l * as;
        for (int w = 0; w < width; w++) crow[w] ^= arow[w];
      }
    if (width) crow[width - 1] &= hm;
  }
}

/* ---- y (^)= A * x, x and y packed rows (x: n bits, y: m bits) ---------------- */
void oracle_mul_v(word *y, const word *A, int as, const word *x, int m, int n, int clear) {
  int nw = width_of(n), mw = width_of(m);
  word nm = high_mask(n);
  if (clear) memset(y, 0, sizeof(word) * (size_t)mw);
  for (int i = 0; i < m; i++) {
    word acc = 0;
    for (int w = 0; w < nw; w++) {
      word t = A[(size_t)i * as + w] & x[w];
      if (w == nw - 1) t &= nm;
      acc ^= t;
    }
    if (__builtin_parityll(acc)) y[i / 64] ^= (word)1 << (i % 64);
  }
}

/* ---- D = S^T (mzd.rs:148) and C = A ^ B (mzd.rs:223) ------------------------- */
void oracle_transpose(word *D, int ds, const word *S, int ss, int nrows, int ncols) {
  for (int j = 0; j < ncols; j++) memset(D + (size_t)j * ds, 0, sizeof(word) * (size_t)width_of(nrows));
  for (int i = 0; i < nrows; i++)
    for (int j = 0; j < ncols; j++)
      if (get_bit(S, ss, i, j)) D[(size_t)j * ds + i / 64] |= (word)1 << (i % 64);
}

void oracle_add(word *C, int cs, const word *A, int as, const word *B, int bs, int nrows,
                int ncols) {
  int width = width_of(ncols);
  word hm = high_mask(ncols);
  for (int i = 0; i < nrows; i++) {
    for (int w = 0; w < width; w++)
      C[(size_t)i * cs + w] = A[(size_t)i * as + w] ^ B[(size_t)i * bs + w];
    if (width) C[(size_t)i * cs + width - 1] &= hm;
  }
}

/* ---- Freivalds over GF(2): A*(B*X) == C*X for 64 random vectors packed in X ---
 * X is n x 64 (one word per row of X). Returns 1 when equal. O(mk + kn + mn). */
static void mat_times_x(word *out, const word *M, int ms, int rows, int cols, const word *X) {
  for (int i = 0; i < rows; i++) {
    word acc = 0;
    const word *row = M + (size_t)i * ms;
    for (int j = 0; j < cols; j++)
      if ((row[j / 64] >> (j % 64)) & 1) acc ^= X[j];
    out[i] = acc;
  }
}

int oracle_freivalds(const word *C, int cs, const word *A, int as, const word *B, int bs, int m,
                     int k, int n, uint64_t seed) {
  word *X = malloc(sizeof(word) * (size_t)n), *BX = malloc(sizeof(word) * (size_t)k),
       *ABX = malloc(sizeof(word) * (size_t)m), *CX = malloc(sizeof(word) * (size_t)m);
  for (int j = 0; j < n; j++) X[j] = splitmix64_at(seed, (uint64_t)j);
  mat_times_x(BX, B, bs, k, n, X);
  mat_times_x(ABX, A, as, m, k, BX);
  mat_times_x(CX, C, cs, m, n, X);
  int ok = memcmp(ABX, CX, sizeof(word) * (size_t)m) == 0;
  free(X); free(BX); free(ABX); free(CX);
  return ok;
}

/* ---- consumers of the product: Gauss-Jordan from the definition -----------
 * Reduced row echelon form in place, pivots searched in columns [0, pivot_cols) only, leftmost
 * column first, first row with the bit set swapped up (textbook elimination; what
 * mzd_echelonize(A, 1) returns is this form, m4ri-sys/src/echelonform.rs:9-16 — the reduced row
 * echelon form of a matrix is unique, so the row chosen as pivot does not matter). Returns the
 * rank; pivcols (rank entries, may be NULL) receives the pivot columns. */
int oracle_rref(word *M, int stride, int m, int n, int pivot_cols, int *pivcols) {
  const int width = width_of(n);
  int rank = 0;
  if (pivot_cols > n) pivot_cols = n;
  for (int c = 0; c < pivot_cols && rank < m; c++) {
    const int cw = c / 64;
    const word bit = (word)1 << (c % 64);
    int p = -1;
    for (int i = rank; i < m; i++)
      if (M[(size_t)i * stride + cw] & bit) { p = i; break; }
    if (p < 0) continue;
    if (p != rank)
      for (int w = 0; w < width; w++) {
        word t = M[(size_t)p * stride + w];
        M[(size_t)p * stride + w] = M[(size_t)rank * stride + w];
        M[(size_t)rank * stride + w] = t;
      }
    const word *pr = M + (size_t)rank * stride;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < m; i++) {
      if (i == rank || !(M[(size_t)i * stride + cw] & bit)) continue;
      word *row = M + (size_t)i * stride;
      for (int w = cw; w < width; w++) row[w] ^= pr[w];
      /* (words left of cw: the pivot row is zero there in every non-pivot column it could
       * still touch only if an earlier column had no pivot — handle those too) */
      for (int w = 0; w < cw; w++) row[w] ^= pr[w];
    }
    if (pivcols) pivcols[rank] = c;
    rank++;
  }
  return rank;
}

/* ---- PLE / PLUQ decomposition from the definition (m4ri-sys/src/ple.rs:35-65) --------------
 * Textbook elimination, column by column: the first row (at or below the current rank) with the
 * bit set is swapped up (P[rank] = that row, LAPACK-style transposition), Q[rank] = the column;
 * every lower row with the bit set gets the pivot row added on the columns RIGHT of the pivot
 * column and keeps the bit itself — its multiplier, the entry of L. Afterwards L is compressed:
 * column k of L moves from column Q[k] to column k (ple.rs:44-51: "L and S are stored in place
 * over A"). Storage on return (r = rank): row i < r = [L[i][0..i) | zeros | E[i][Q[i]..n)], row
 * i >= r = [L[i][0..r) | zeros]; P[i] = i and Q[i] = i beyond the rank. pluq != 0 additionally
 * swaps columns i and Q[i] of the echelon rows for i = 0..r-1 in turn (on rows 0..i: the rows
 * below hold L there), which makes the upper part U upper triangular (ple.rs:10-33). */
static int row_bit(const word *row, int c) { return (int)((row[c / 64] >> (c % 64)) & 1); }
static void row_put(word *row, int c, int v) {
  row[c / 64] = (row[c / 64] & ~((word)1 << (c % 64))) | ((word)(v & 1) << (c % 64));
}
int oracle_ple(word *M, int stride, int m, int n, int *P, int *Q, int pluq) {
  const int width = width_of(n);
  int rank = 0;
  for (int i = 0; i < m; i++) P[i] = i;
  for (int c = 0; c < n; c++) Q[c] = c;
  for (int c = 0; c < n && rank < m; c++) {
    const int cw = c / 64, cb = c % 64;
    const word bit = (word)1 << cb;
    int p = -1;
    for (int i = rank; i < m; i++)
      if (M[(size_t)i * stride + cw] & bit) { p = i; break; }
    if (p < 0) continue;
    if (p != rank)
      for (int w = 0; w < width; w++) {
        word t = M[(size_t)p * stride + w];
        M[(size_t)p * stride + w] = M[(size_t)rank * stride + w];
        M[(size_t)rank * stride + w] = t;
      }
    P[rank] = p;
    Q[rank] = c;
    const word *pr = M + (size_t)rank * stride;
    const word above = cb == 63 ? 0 : ~(word)0 << (cb + 1);
#pragma omp parallel for schedule(static)
    for (int i = rank + 1; i < m; i++) {
      word *row = M + (size_t)i * stride;
      if (!(row[cw] & bit)) continue;
      row[cw] ^= pr[cw] & above;
      for (int w = cw + 1; w < width; w++) row[w] ^= pr[w];
    }
    rank++;
  }
  /* compress L: ascending k never overwrites a multiplier that is still to be moved
   * (destinations so far are < k <= Q[k]) */
  for (int i = 0; i < m; i++) {
    word *row = M + (size_t)i * stride;
    const int kmax = i < rank ? i : rank;
    for (int k = 0; k < kmax; k++) {
      if (Q[k] == k) continue;
      const int v = row_bit(row, Q[k]);
      row_put(row, Q[k], 0);
      row_put(row, k, v);
    }
  }
  if (pluq)
    for (int i = 0; i < rank; i++) {
      if (Q[i] == i) continue;
      for (int r = 0; r <= i; r++) {
        word *row = M + (size_t)r * stride;
        const int a = row_bit(row, i), b = row_bit(row, Q[i]);
        row_put(row, i, b);
        row_put(row, Q[i], a);
      }
    }
  return rank;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
