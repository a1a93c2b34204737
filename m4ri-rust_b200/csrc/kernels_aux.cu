// kernels_aux.cu — the HBM-bound kernels around the product: XOR-add (the
// Strassen / addmul "add", reference mzd_add m4ri-sys/src/mzd.rs:223), bit-matrix
// transpose (mzd_transpose, mzd.rs:148), mat-vec (mul_slice path,
// m4ri-rust/src/friendly/binary_matrix.rs:416-431), vec-mat (_mzd_mul_va,
// mzd.rs:181), the AND+popcount product behind the *_naive entry points
// (mzd.rs:152-173), plus input generation and checksums for the harness.
#include <mutex>

#include "common.cuh"

namespace m4b {

// Programmatic dependent launch for the short HBM-bound kernels (a mat-vec over 64 MiB lasts ~14 us,
// of which ~1.5 us is launch latency and ramp): the kernel is launched with programmatic stream
// serialization and begins with griddepcontrol.wait — it reads and writes nothing before
// everything ahead of it in the stream has completed and flushed — so only its launch overlaps
// the tail of its predecessor. launch_dependents lets its own successor do the same.
#define M4B_PDL_ENTER() asm volatile("griddepcontrol.wait;\n\tgriddepcontrol.launch_dependents;" ::: "memory")
template <typename... KArgs, typename... Args>
static void launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t s, Args... args) {
  static const bool off = getenv("M4RI_B200_NO_PDL") != nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = off ? 0 : 1;
  M4B_CUDA(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
}

static inline unsigned cdiv(long long a, long long b) { return (unsigned)((a + b - 1) / b); }

// CTAs of `kernel` (256 threads, no dynamic shared memory) that are resident on the device at
// once: the grid of the persistent (grid-stride) kernels, so that there is no tail wave.
template <typename K>
static long long resident_ctas(K kernel) {
  static std::atomic<int> cache[64];  // per kernel instantiation and device (0 = not asked yet)
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
  int per_sm = cache[dev].load(std::memory_order_relaxed);
  if (per_sm == 0) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 256, 0) != cudaSuccess || per_sm < 1) {
      (void)cudaGetLastError();
      per_sm = 1;
    }
    cache[dev].store(per_sm, std::memory_order_relaxed);
  }
  return (long long)sm_count() * per_sm;
}

// ---------------------------------------------------------------------------
// zero / mask / copy / add
// ---------------------------------------------------------------------------
void k_zero(DMat *M, cudaStream_t s) {
  if (M->nrows == 0 || M->width == 0) return;
  M4B_CUDA(cudaMemset2DAsync(M->d, (size_t)M->rowstride * 8, 0, (size_t)M->width * 8, M->nrows, s));
}

__global__ void mask_excess_kernel(word *d, int nrows, int width, int rs, word hm) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < nrows) d[(size_t)r * rs + width - 1] &= hm;
}

void k_mask_excess(DMat *M, cudaStream_t s) {
  if (M->ncols % 64 == 0 || M->nrows == 0) return;
  mask_excess_kernel<<<cdiv(M->nrows, 256), 256, 0, s>>>(M->d, M->nrows, M->width, M->rowstride,
                                                          high_mask(M->ncols));
  M4B_LAUNCHED();
}

__global__ void merge_excess_kernel(word *last, int stride, const word *col, int nrows, word keep) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < nrows) last[(size_t)r * stride] = (last[(size_t)r * stride] & keep) | (col[r] & ~keep);
}
void k_merge_excess(word *last, int stride, const word *host_col, int nrows, word keep, cudaStream_t s) {
  if (nrows <= 0) return;
  merge_excess_kernel<<<cdiv(nrows, 256), 256, 0, s>>>(last, stride, host_col, nrows, keep);
  M4B_LAUNCHED();
}

void k_copy(DMat *D, const DMat *A, cudaStream_t s) {
  if (A->nrows == 0 || A->width == 0) return;
  M4B_CUDA(cudaMemcpy2DAsync(D->d, (size_t)D->rowstride * 8, A->d, (size_t)A->rowstride * 8,
                             (size_t)A->width * 8, A->nrows, cudaMemcpyDeviceToDevice, s));
}

// C = A ^ B over nrows x width words. 16-byte vectors: 3 x 16 B of traffic per
// thread-iteration, pure HBM streaming (algorithmic bytes = 3 * nrows * width * 8).
template <typename V>
__global__ void add_kernel(V *C, const V *A, const V *B, int nrows, int wv, int cs,
                           int as, int bs) {
  long long total = (long long)nrows * wv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int r = (int)(i / wv), w = (int)(i % wv);
    V a = A[(size_t)r * as + w], b = B[(size_t)r * bs + w], c;
    if constexpr (sizeof(V) == 16) {
      c.x = a.x ^ b.x; c.y = a.y ^ b.y;
    } else {
      c = a ^ b;
    }
    C[(size_t)r * cs + w] = c;
  }
}

void k_add(DMat *C, const DMat *A, const DMat *B, cudaStream_t s) {
  if (C->nrows == 0 || C->width == 0) return;
  bool vec = (C->width % 2 == 0) && (C->rowstride % 2 == 0) && (A->rowstride % 2 == 0) &&
             (B->rowstride % 2 == 0) && (((uintptr_t)C->d | (uintptr_t)A->d | (uintptr_t)B->d) % 16 == 0);
  long long total = (long long)C->nrows * (vec ? C->width / 2 : C->width);
  unsigned grid = (unsigned)min((long long)cdiv(total, 256), (long long)sm_count() * 16);
  if (vec)
    add_kernel<ulonglong2><<<grid, 256, 0, s>>>((ulonglong2 *)C->d, (const ulonglong2 *)A->d,
                                                (const ulonglong2 *)B->d, C->nrows, C->width / 2,
                                                C->rowstride / 2, A->rowstride / 2, B->rowstride / 2);
  else
    add_kernel<word><<<grid, 256, 0, s>>>(C->d, A->d, B->d, C->nrows, C->width, C->rowstride,
                                          A->rowstride, B->rowstride);
  M4B_LAUNCHED();
}

// The operand sums of one fused Strassen level (api.cu::strassen_fused_level): the four quadrants
// are read once, five pairwise sums are written. 16-byte vectors; 9 instead of 15 transfers.
struct QuadSums {
  const ulonglong2 *q[4];
  ulonglong2 *o[5];
  int pa[5], pb[5];
};
__global__ void quad_sums_kernel(QuadSums a, int nrows, int wv, int qs, int os) {
  const long long total = (long long)nrows * wv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / wv), w = (int)(i % wv);
    ulonglong2 v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) v[j] = a.q[j][(size_t)r * qs + w];
#pragma unroll
    for (int j = 0; j < 5; j++) {
      ulonglong2 x = v[0], y = v[0];
#pragma unroll
      for (int t = 1; t < 4; t++) {
        if (a.pa[j] == t) x = v[t];
        if (a.pb[j] == t) y = v[t];
      }
      a.o[j][(size_t)r * os + w] = make_ulonglong2(x.x ^ y.x, x.y ^ y.y);
    }
  }
}
void k_quad_sums(DMat *const out[5], const DMat q[4], const int pairs[5][2], cudaStream_t s) {
  QuadSums a;
  const int nrows = q[0].nrows, width = q[0].width;
  if (nrows == 0 || width == 0) return;
  for (int j = 0; j < 4; j++) {
    if (q[j].nrows != nrows || q[j].width != width || q[j].rowstride != q[0].rowstride || (q[j].rowstride & 1) ||
        ((uintptr_t)q[j].d & 15) || (width & 1))
      die("m4ri_b200: quad_sums needs four equal, 16-byte aligned quadrants\n");
    a.q[j] = (const ulonglong2 *)q[j].d;
  }
  for (int j = 0; j < 5; j++) {
    if (out[j]->nrows != nrows || out[j]->width != width || out[j]->rowstride != out[0]->rowstride ||
        (out[j]->rowstride & 1) || ((uintptr_t)out[j]->d & 15))
      die("m4ri_b200: quad_sums: bad output\n");
    a.o[j] = (ulonglong2 *)out[j]->d;
    a.pa[j] = pairs[j][0];
    a.pb[j] = pairs[j][1];
  }
  const long long total = (long long)nrows * (width / 2);
  unsigned grid = (unsigned)min((long long)cdiv(total, 256), (long long)sm_count() * 16);
  quad_sums_kernel<<<grid, 256, 0, s>>>(a, nrows, width / 2, q[0].rowstride / 2, out[0]->rowstride / 2);
  M4B_LAUNCHED();
}

// ---------------------------------------------------------------------------
// transpose: 256-row x 1024-bit input tiles. 32x32 bit blocks are turned in
// registers (five masked shuffle exchanges), collected in shared memory and
// written as 32 contiguous bytes per output row (whole sectors). HBM-bound:
// algorithmic bytes = 2 * nrows * width * 8.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t transpose32(uint32_t x, int lane) {
  uint32_t m = 0x0000FFFFu;
#pragma unroll
  for (int sft = 16; sft; sft >>= 1) {
    const uint32_t y = __shfl_xor_sync(0xffffffffu, x, sft);
    x = (lane & sft) ? ((x & ~m) | ((y >> sft) & m)) : ((x & m) | ((y << sft) & ~m));
    m ^= m << (sft >> 1);
  }
  return x;
}

// 32 x 32 bit block held by ONE thread (a[r] bit c -> a[c] bit r, LSB first): five stages of 16
// masked block swaps, ~480 integer operations and no shuffle — the warp-wide form above costs
// 160 shuffles and ~640 operations per block and was bound by the shuffle / shared-memory pipe
// (46-50 % of the HBM copy rate).
template <int J>
__device__ __forceinline__ void transpose_stage(uint32_t (&a)[32]) {
  constexpr uint32_t M = J == 16 ? 0x0000FFFFu : J == 8 ? 0x00FF00FFu : J == 4 ? 0x0F0F0F0Fu : J == 2 ? 0x33333333u : 0x55555555u;
#pragma unroll
  for (int k = 0; k < 32; k++) {
    if ((k & J) == 0) {  // (compile-time after unrolling: the array stays in registers)
      const uint32_t t = ((a[k] >> J) ^ a[k + J]) & M;
      a[k + J] ^= t;
      a[k] ^= t << J;
    }
  }
}
__device__ __forceinline__ void transpose32x32(uint32_t (&a)[32]) {
  transpose_stage<16>(a);
  transpose_stage<8>(a);
  transpose_stage<4>(a);
  transpose_stage<2>(a);
  transpose_stage<1>(a);
}

// Thread-local form of the same tile transpose (the default; M4RI_B200_TRANSPOSE_SHFL=1 keeps the
// shuffle form for A/B timing): thread t turns block (row-block t / 32, column word t % 32).
__global__ void __launch_bounds__(256)
transpose_local_kernel(uint32_t *__restrict__ D, const uint32_t *__restrict__ S, int nrows, int ncols,
                       int sw32, int ss32, int ds32, int dw32) {
  __shared__ uint32_t buf[256 * 33];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int r0 = blockIdx.y * 256, cw0 = blockIdx.x * 32;
#pragma unroll 8
  for (int i = 0; i < 32; i++) {
    const int row = i * 8 + w;
    uint32_t v = 0;
    if (r0 + row < nrows && cw0 + l < sw32) v = S[(size_t)(r0 + row) * ss32 + cw0 + l];
    buf[row * 33 + l] = v;
  }
  __syncthreads();
  // rows [32w, 32w + 32) of column word l: bank (i + l) % 32 — conflict-free
  uint32_t a[32];
#pragma unroll
  for (int i = 0; i < 32; i++) a[i] = buf[(32 * w + i) * 33 + l];
  transpose32x32(a);
  __syncthreads();
  // a[j] = word w of output row 32 l + j; out tile: 1024 rows x 8 words, every group of 32 rows
  // skewed by one word: for a fixed j the 32 lanes write rows 32 l + j — bank (8 j + w + l) % 32
#pragma unroll
  for (int j = 0; j < 32; j++) buf[(32 * l + j) * 8 + l + w] = a[j];
  __syncthreads();
#pragma unroll 8
  for (int i = 0; i < 32; i++) {
    const int idx = i * 256 + threadIdx.x;
    const int orl = idx >> 3, g = idx & 7;
    const int orow = cw0 * 32 + orl, ow = (r0 >> 5) + g;
    if (orow < ncols && ow < dw32) D[(size_t)orow * ds32 + ow] = buf[orl * 8 + (orl >> 5) + g];
  }
}

__global__ void __launch_bounds__(256)
transpose_kernel(uint32_t *__restrict__ D, const uint32_t *__restrict__ S, int nrows, int ncols,
                 int sw32, int ss32, int ds32, int dw32) {
  __shared__ uint32_t buf[256 * 33];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int r0 = blockIdx.y * 256, cw0 = blockIdx.x * 32;
#pragma unroll 8
  for (int i = 0; i < 32; i++) {
    const int row = i * 8 + w;
    uint32_t v = 0;
    if (r0 + row < nrows && cw0 + l < sw32) v = S[(size_t)(r0 + row) * ss32 + cw0 + l];
    buf[row * 33 + l] = v;
  }
  __syncthreads();
  // warp w turns the 32 blocks of input rows [32w, 32w+32): afterwards lane j of
  // res[c] is word w of output row (cw0 + c) * 32 + j
  uint32_t res[32];
#pragma unroll
  for (int c = 0; c < 32; c++) res[c] = transpose32(buf[(32 * w + l) * 33 + c], l);
  __syncthreads();
  // out tile: 1024 rows x 8 words, word index swizzled by the row (conflict-free both ways)
#pragma unroll
  for (int c = 0; c < 32; c++) buf[(c * 32 + l) * 8 + (w ^ ((l >> 2) & 7))] = res[c];
  __syncthreads();
#pragma unroll 8
  for (int i = 0; i < 32; i++) {
    const int idx = i * 256 + threadIdx.x;
    const int orl = idx >> 3, g = idx & 7;
    const int orow = cw0 * 32 + orl, ow = (r0 >> 5) + g;
    if (orow < ncols && ow < dw32)
      D[(size_t)orow * ds32 + ow] = buf[orl * 8 + (g ^ ((orl >> 2) & 7))];
  }
}

// The same transpose written to up to 8 destinations at once: the fused "broadcast + transpose"
// of the row-sharded product (sharded.py). The source slab may live in a peer GPU's memory
// (read once over NVLink) and every destination but one is a peer's Bt buffer (P2P stores):
// one pass over the data instead of an NCCL broadcast followed by a local transpose.
struct TransposeDsts {
  uint32_t *p[8];
  int n;
};
__global__ void __launch_bounds__(256)
transpose_scatter_kernel(TransposeDsts dsts, const uint32_t *__restrict__ S, int nrows, int ncols,
                         int sw32, int ss32, int ds32, int dw32) {
  __shared__ uint32_t buf[256 * 33];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int r0 = blockIdx.y * 256, cw0 = blockIdx.x * 32;
#pragma unroll 8
  for (int i = 0; i < 32; i++) {
    const int row = i * 8 + w;
    uint32_t v = 0;
    if (r0 + row < nrows && cw0 + l < sw32) v = S[(size_t)(r0 + row) * ss32 + cw0 + l];
    buf[row * 33 + l] = v;
  }
  __syncthreads();
  uint32_t res[32];
#pragma unroll
  for (int c = 0; c < 32; c++) res[c] = transpose32(buf[(32 * w + l) * 33 + c], l);
  __syncthreads();
#pragma unroll
  for (int c = 0; c < 32; c++) buf[(c * 32 + l) * 8 + (w ^ ((l >> 2) & 7))] = res[c];
  __syncthreads();
#pragma unroll 4
  for (int i = 0; i < 32; i++) {
    const int idx = i * 256 + threadIdx.x;
    const int orl = idx >> 3, g = idx & 7;
    const int orow = cw0 * 32 + orl, ow = (r0 >> 5) + g;
    if (orow < ncols && ow < dw32) {
      const uint32_t v = buf[orl * 8 + (g ^ ((orl >> 2) & 7))];
      for (int d = 0; d < dsts.n; d++) dsts.p[d][(size_t)orow * ds32 + ow] = v;
    }
  }
}

// Wide form of the same scatter: 1024-row x 1024-bit input tiles, so that every output row
// leaves the SM as 128 contiguous bytes (16-byte stores, a quarter-warp per row) instead of 32.
// Over NVLink the 32-byte peer stores of the kernel above reach ~190 GB/s per GPU; 128-byte
// segments are whole NVLink write packets (profiles/r1k_exchange_probe_*.json).
// 512 threads: each half-CTA turns two 256-row sub-tiles (own 33 KB input buffer, named
// barrier) into a shared 128 KB staging tile of 1024 output rows x 32 words; row r is stored
// rotated by r words, which makes both the transposed stores (32 rows, one word each) and the
// row-wise read-out (8 lanes x 4 words per row) conflict-free.
constexpr int TW_IN = 256 * 33;                       // words per input buffer
constexpr int TW_SMEM = (1024 * 32 + 2 * TW_IN) * 4;  // bytes
__global__ void __launch_bounds__(512)
transpose_wide_kernel(TransposeDsts dsts, const uint32_t *__restrict__ S, int nrows, int ncols,
                      int sw32, int ss32, int ds32, int dw32) {
  extern __shared__ __align__(16) uint32_t tw_smem[];
  uint32_t *out = tw_smem;
  const int half = threadIdx.x >> 8, t = threadIdx.x & 255;
  uint32_t *in = tw_smem + 1024 * 32 + half * TW_IN;
  const int w = t >> 5, l = t & 31;
  const int r0 = blockIdx.y * 1024, cw0 = blockIdx.x * 32;
  for (int sub = half; sub < 4; sub += 2) {
    const int rs = r0 + sub * 256;
#pragma unroll 8
    for (int i = 0; i < 32; i++) {
      const int row = i * 8 + w;
      uint32_t v = 0;
      if (rs + row < nrows && cw0 + l < sw32) v = S[(size_t)(rs + row) * ss32 + cw0 + l];
      in[row * 33 + l] = v;
    }
    asm volatile("bar.sync %0, 256;" ::"r"(1 + half) : "memory");
    // thread (w, l) turns the block of rows [32 w, 32 w + 32) x column word l in registers
    // (transpose32x32: no shuffles); a[j] = word (sub*8 + w) of output row 32 l + j
    uint32_t a[32];
#pragma unroll
    for (int i = 0; i < 32; i++) a[i] = in[(32 * w + i) * 33 + l];
    transpose32x32(a);
    asm volatile("bar.sync %0, 256;" ::"r"(1 + half) : "memory");  // the input buffer may be refilled
    // row r is stored rotated by r + r / 32 words: for a fixed j the 32 lanes (rows 32 l + j) hit 32 banks
#pragma unroll
    for (int j = 0; j < 32; j++) out[(32 * l + j) * 32 + ((sub * 8 + w + j + l) & 31)] = a[j];
  }
  __syncthreads();
#pragma unroll 4
  for (int it = 0; it < 16; it++) {
    const int idx = it * 512 + threadIdx.x;
    const int r = idx >> 3, g = idx & 7;
    const int orow = cw0 * 32 + r, ow = (r0 >> 5) + 4 * g;
    if (orow < ncols && ow < dw32) {
      uint4 v;
      const int rot = r + (r >> 5);
      v.x = out[r * 32 + ((4 * g + 0 + rot) & 31)];
      v.y = out[r * 32 + ((4 * g + 1 + rot) & 31)];
      v.z = out[r * 32 + ((4 * g + 2 + rot) & 31)];
      v.w = out[r * 32 + ((4 * g + 3 + rot) & 31)];
      const size_t off = (size_t)orow * ds32 + ow;
      if (ow + 4 <= dw32) {
        for (int d = 0; d < dsts.n; d++) *reinterpret_cast<uint4 *>(dsts.p[d] + off) = v;
      } else {  // ragged last chunk of a row: never write past the destination's width (windows)
        const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
        for (int d = 0; d < dsts.n; d++)
          for (int i = 0; i < 4 && ow + i < dw32; i++) dsts.p[d][off + i] = vv[i];
      }
    }
  }
}

static int transpose_variant() {  // M4RI_B200_TRANSPOSE=narrow|wide (experiments, tests); default: see callers
  const char *e = getenv("M4RI_B200_TRANSPOSE");
  return !e ? 0 : !strcmp(e, "narrow") ? 1 : !strcmp(e, "wide") ? 2 : 0;
}

static bool transpose_wide_ok(const uint32_t *const *dst, int ndst, int ds32) {
  if (ds32 % 4) return false;
  for (int i = 0; i < ndst; i++)
    if ((uintptr_t)dst[i] % 16) return false;
  return true;
}

static void launch_transpose_wide(const TransposeDsts &dsts, const DMat *A, int ds32, int dw32, cudaStream_t s) {
  static std::atomic<unsigned long long> attr_seen{0};
  if (first_use_on_this_device(attr_seen))
    M4B_CUDA(cudaFuncSetAttribute(transpose_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TW_SMEM));
  const int sw32 = A->width * 2;
  dim3 grid(cdiv(sw32, 32), cdiv(dw32, 32));
  transpose_wide_kernel<<<grid, 512, TW_SMEM, s>>>(dsts, (const uint32_t *)A->d, A->nrows, A->ncols, sw32,
                                                   A->rowstride * 2, ds32, dw32);
  M4B_LAUNCHED();
}

void k_transpose_scatter(DMat *const *D, int ndst, const DMat *A, cudaStream_t s) {
  if (ndst < 1 || ndst > 8) die("m4ri_b200: transpose_scatter takes 1 to 8 destinations\n");
  if (A->nrows == 0 || A->ncols == 0) return;
  TransposeDsts dsts;
  dsts.n = ndst;
  for (int i = 0; i < ndst; i++) {
    if (D[i]->nrows != A->ncols || D[i]->ncols != A->nrows || D[i]->rowstride != D[0]->rowstride)
      die("m4ri_b200: transpose_scatter: every destination must be ncols(A) x nrows(A) with one stride\n");
    dsts.p[i] = (uint32_t *)D[i]->d;
  }
  const int sw32 = A->width * 2, dw32 = D[0]->width * 2;
  // destinations of a scatter are peer or multicast mappings (stores over NVLink): the wide
  // kernel (whole 128-byte rows per store burst) unless told otherwise
  const int var = transpose_variant();
  if (var != 1 && transpose_wide_ok(dsts.p, ndst, D[0]->rowstride * 2)) {
    launch_transpose_wide(dsts, A, D[0]->rowstride * 2, dw32, s);
    return;
  }
  dim3 grid(cdiv(sw32, 32), cdiv(dw32, 8));
  transpose_scatter_kernel<<<grid, 256, 0, s>>>(dsts, (const uint32_t *)A->d, A->nrows, A->ncols, sw32,
                                                A->rowstride * 2, D[0]->rowstride * 2, dw32);
  M4B_LAUNCHED();
}

void k_transpose(DMat *D, const DMat *A, cudaStream_t s) {
  if (A->nrows == 0 || A->ncols == 0) return;
  const int sw32 = A->width * 2, dw32 = D->width * 2;
  if (transpose_variant() == 2) {
    TransposeDsts dsts;
    dsts.n = 1;
    dsts.p[0] = (uint32_t *)D->d;
    if (transpose_wide_ok(dsts.p, 1, D->rowstride * 2)) {
      launch_transpose_wide(dsts, A, D->rowstride * 2, dw32, s);
      return;
    }
  }
  // row blocks cover every 32-bit half of D's words, so stale bits of a reused
  // destination are overwritten with zeros (rows >= nrows read as zero)
  dim3 grid(cdiv(sw32, 32), cdiv(dw32, 8));
  static const bool shfl = getenv("M4RI_B200_TRANSPOSE_SHFL") != nullptr;
  if (shfl)
    transpose_kernel<<<grid, 256, 0, s>>>((uint32_t *)D->d, (const uint32_t *)A->d, A->nrows,
                                          A->ncols, sw32, A->rowstride * 2, D->rowstride * 2, dw32);
  else
    transpose_local_kernel<<<grid, 256, 0, s>>>((uint32_t *)D->d, (const uint32_t *)A->d, A->nrows,
                                                A->ncols, sw32, A->rowstride * 2, D->rowstride * 2, dw32);
  M4B_LAUNCHED();
}

// ---------------------------------------------------------------------------
// input generation and checksum (harness support)
// ---------------------------------------------------------------------------
__host__ __device__ static inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

__global__ void fill_random_kernel(word *d, int nrows, int width, int rs, uint64_t seed, word hm,
                                   uint64_t first) {
  long long total = (long long)nrows * width;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int r = (int)(i / width), w = (int)(i % width);
    word z = mix64(seed + (first + (uint64_t)i + 1) * 0x9E3779B97F4A7C15ULL);
    if (w == width - 1) z &= hm;
    d[(size_t)r * rs + w] = z;
  }
}

void k_fill_random(DMat *M, uint64_t seed, uint64_t first_index, cudaStream_t s) {
  if (M->nrows == 0 || M->width == 0) return;
  long long total = (long long)M->nrows * M->width;
  unsigned grid = (unsigned)min((long long)cdiv(total, 256), (long long)sm_count() * 16);
  fill_random_kernel<<<grid, 256, 0, s>>>(M->d, M->nrows, M->width, M->rowstride, seed,
                                          high_mask(M->ncols), first_index);
  M4B_LAUNCHED();
}

__global__ void checksum_kernel(const word *d, int nrows, int width, int rs,
                                unsigned long long *out) {
  long long total = (long long)nrows * width;
  uint64_t acc = 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    int r = (int)(i / width), w = (int)(i % width);
    acc ^= mix64(d[(size_t)r * rs + w] + (uint64_t)(i + 1) * 0x9E3779B97F4A7C15ULL);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicXor(out, (unsigned long long)acc);
}

uint64_t k_checksum(const DMat *M, cudaStream_t s) {
  unsigned long long *dout, h = 0;
  M4B_CUDA(cudaMallocAsync(&dout, 8, s));
  M4B_CUDA(cudaMemsetAsync(dout, 0, 8, s));
  long long total = (long long)M->nrows * M->width;
  if (total) {
    unsigned grid = (unsigned)min((long long)cdiv(total, 256), (long long)sm_count() * 16);
    checksum_kernel<<<grid, 256, 0, s>>>(M->d, M->nrows, M->width, M->rowstride, dout);
    M4B_LAUNCHED();
  }
  M4B_CUDA(cudaMemcpyAsync(&h, dout, 8, cudaMemcpyDeviceToHost, s));
  M4B_CUDA(cudaFreeAsync(dout, s));
  M4B_CUDA(cudaStreamSynchronize(s));
  return h;
}

__global__ void is_zero_kernel(const word *d, int nrows, int width, int rs, int *nonzero) {
  long long total = (long long)nrows * width;
  word acc = 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x)
    acc |= d[(size_t)(i / width) * rs + (i % width)];
  if (__any_sync(0xffffffffu, acc != 0) && (threadIdx.x & 31) == 0) atomicOr(nonzero, 1);
}

bool k_is_zero(const DMat *M, cudaStream_t s) {
  long long total = (long long)M->nrows * M->width;
  if (!total) return true;
  int *dflag, h = 0;
  M4B_CUDA(cudaMallocAsync(&dflag, 4, s));
  M4B_CUDA(cudaMemsetAsync(dflag, 0, 4, s));
  unsigned grid = (unsigned)min((long long)cdiv(total, 256), (long long)sm_count() * 16);
  is_zero_kernel<<<grid, 256, 0, s>>>(M->d, M->nrows, M->width, M->rowstride, dflag);
  M4B_LAUNCHED();
  M4B_CUDA(cudaMemcpyAsync(&h, dflag, 4, cudaMemcpyDeviceToHost, s));
  M4B_CUDA(cudaFreeAsync(dflag, s));
  M4B_CUDA(cudaStreamSynchronize(s));
  return h == 0;
}

__global__ void flip_bit_kernel(word *d, size_t idx, int bit) { d[idx] ^= (word)1 << bit; }
void k_flip_bit(DMat *M, int r, int c, cudaStream_t s) {
  if (r < 0 || r >= M->nrows || c < 0 || c >= M->ncols) die("m4ri_b200: flip_bit out of range\n");
  flip_bit_kernel<<<1, 1, 0, s>>>(M->d, (size_t)r * M->rowstride + c / 64, c % 64);
  M4B_LAUNCHED();
}

// ---------------------------------------------------------------------------
// mat-vec: y (^)= A*x, packed vector in, packed vector out. G lanes share a row
// (G = 32 for long rows, fewer for short ones so every lane still moves 16 B).
// HBM-bound: algorithmic bytes = 8 * (m*width + width) + ceil(m/64)*8.
// ---------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256)
mul_v_kernel(uint32_t *__restrict__ y, const ulonglong2 *__restrict__ A,
             const ulonglong2 *__restrict__ x, int m, int wv /*16B vectors per row*/,
             int as /*row stride in 16B vectors*/) {
  const int lane = threadIdx.x & 31;
  const int g = lane / G, gl = lane % G;
  constexpr int RPW = 32 / G;  // rows per warp
  long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  long long row = warp * RPW + g;
  uint64_t acc = 0;
  if (row < m) {
    const ulonglong2 *arow = A + (size_t)row * as;
#pragma unroll 4
    for (int w = gl; w < wv; w += G) {
      ulonglong2 a = arow[w];
      ulonglong2 b = __ldg(&x[w]);
      acc ^= (a.x & b.x) ^ (a.y & b.y);
    }
  }
#pragma unroll
  for (int o = G / 2; o; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  unsigned bit = (row < m) && gl == 0 && (__popcll(acc) & 1);
  unsigned ball = __ballot_sync(0xffffffffu, bit);
  if (lane == 0) {
    // compress the leader bits (lanes 0, G, 2G, ...) into RPW consecutive bits
    unsigned bits = 0;
#pragma unroll
    for (int i = 0; i < RPW; i++) bits |= ((ball >> (i * G)) & 1u) << i;
    long long r0 = warp * RPW;
    if (bits) atomicXor(&y[r0 >> 5], bits << (r0 & 31));
  }
}

// Short rows (width <= 16 words, the LPN shapes): one thread per row reads its whole row with
// 16-byte loads, a warp's 32 parities are one ballot = one aligned 32-bit word of y: no
// atomics, no memset of y, one launch.
template <int WV /*16-byte vectors per row*/>
__global__ void __launch_bounds__(256)
mul_v_rows_kernel(uint32_t *__restrict__ y, const ulonglong2 *__restrict__ A,
                  const ulonglong2 *__restrict__ x, int m, int as, int clear) {
  const long long row = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  ulonglong2 a[WV];
  uint64_t acc = 0;
  if (row < m) {
    const ulonglong2 *arow = A + (size_t)row * as;
#pragma unroll
    for (int w = 0; w < WV; w++)
      asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(a[w].x), "=l"(a[w].y) : "l"(arow + w));
#pragma unroll
    for (int w = 0; w < WV; w++) {
      const ulonglong2 b = __ldg(&x[w]);
      acc ^= (a[w].x & b.x) ^ (a[w].y & b.y);
    }
  }
  const unsigned bits = __ballot_sync(0xffffffffu, (row < m) && (__popcll(acc) & 1));
  if ((threadIdx.x & 31) == 0 && row < (((long long)m + 63) & ~63LL)) {  // whole 64-bit words of y
    uint32_t *dst = y + (row >> 5);
    *dst = clear ? bits : (*dst ^ bits);
  }
}

// Long rows (>= 32 16-byte vectors) and many of them: a CTA of 8 warps owns 32 consecutive rows
// = one aligned 32-bit word of y (no atomics, no memset of y). Warp w walks its 4 rows together:
// per step one 16-byte load of x feeds four independent 16-byte loads of A (two steps unrolled:
// 128 B in flight per lane); a row's parity is the parity of one ballot of the lane parities.
__global__ void __launch_bounds__(256)
mul_v_long_kernel(uint32_t *__restrict__ y, const ulonglong2 *__restrict__ A,
                  const ulonglong2 *__restrict__ x, int m, int wv, int as, int clear) {
  __shared__ unsigned sbits[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long row0 = blockIdx.x * 32LL + w * 4;
  const ulonglong2 *ar[4];
#pragma unroll
  for (int r = 0; r < 4; r++) {
    const long long row = row0 + r < m ? row0 + r : m - 1;  // rows past m: read a valid row, drop the bit
    ar[r] = A + (size_t)row * as;
  }
  uint64_t acc[4] = {0, 0, 0, 0};
#pragma unroll 2
  for (int v = lane; v < wv; v += 32) {
    const ulonglong2 b = __ldg(&x[v]);
#pragma unroll
    for (int r = 0; r < 4; r++) {
      ulonglong2 a;
      asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(a.x), "=l"(a.y) : "l"(ar[r] + v));
      acc[r] ^= (a.x & b.x) ^ (a.y & b.y);
    }
  }
  unsigned bits = 0;
#pragma unroll
  for (int r = 0; r < 4; r++) {
    const unsigned ball = __ballot_sync(0xffffffffu, __popcll(acc[r]) & 1);
    if (row0 + r < m) bits |= (unsigned)(__popc(ball) & 1) << r;
  }
  if (lane == 0) sbits[w] = bits;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned word32 = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) word32 |= sbits[i] << (4 * i);
    uint32_t *dst = y + blockIdx.x;
    *dst = clear ? word32 : (*dst ^ word32);
  }
}

static bool mul_v_force_rows() {  // M4RI_B200_MULV=rows: keep the thread-per-row kernel (A/B timing, tests)
  const char *e = getenv("M4RI_B200_MULV");
  return e && !strcmp(e, "rows");
}

// Short rows of 1, 2, 4 or 8 16-byte vectors stored back to back (row stride == row length):
// a warp owns 32 consecutive rows = 32*WV consecutive vectors and reads them with WV fully
// coalesced 512-byte requests (lane l of request j takes vector j*32 + l, i.e. row
// j*(32/WV) + l/WV, column l%WV — every 32-byte sector is fetched once, by one request).
// Lane parities are folded over the WV lanes of a row with shuffles; one more shuffle and one
// ballot put the 32 row parities in row order = one aligned 32-bit word of y.
template <int WV, int U /* 32-row groups per warp step: U*WV 16-byte loads in flight per lane */>
__global__ void __launch_bounds__(256)
mul_v_coal_kernel(uint32_t *__restrict__ y, const ulonglong2 *__restrict__ A,
                  const ulonglong2 *__restrict__ x, long long m, int clear) {
  constexpr int RPI = 32 / WV;  // rows per request
  M4B_PDL_ENTER();
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long mpad = (m + 63) & ~63LL;  // y is written in whole 64-bit words
  const long long nvec = m * WV;
  const ulonglong2 b = __ldg(&x[lane % WV]);
  // persistent: the grid is one wave of resident CTAs (no tail wave), every warp strides over
  // the 32*U-row steps
  for (long long row0 = warp * (32 * U); row0 < mpad; row0 += nwarps * (32 * U)) {
    const long long v0 = row0 * WV + lane;
    ulonglong2 a[U * WV];
#pragma unroll
    for (int j = 0; j < U * WV; j++) {
      a[j] = make_ulonglong2(0, 0);
      if (v0 + j * 32 < nvec)
        asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];"
                     : "=l"(a[j].x), "=l"(a[j].y) : "l"(A + v0 + j * 32));
    }
    unsigned v = 0;  // bit j = parity of this lane's vector of request j
#pragma unroll
    for (int j = 0; j < U * WV; j++)
      v |= (unsigned)(__popcll((a[j].x & b.x) ^ (a[j].y & b.y)) & 1) << j;
#pragma unroll
    for (int o = 1; o < WV; o <<= 1) v ^= __shfl_xor_sync(0xffffffffu, v, o);
    // row r (0..31) of group u: request u*WV + r / RPI, lanes (r % RPI)*WV ...
    const unsigned t = __shfl_sync(0xffffffffu, v, (lane % RPI) * WV) >> (lane / RPI);
#pragma unroll
    for (int u = 0; u < U; u++) {
      const unsigned bits = __ballot_sync(0xffffffffu, (t >> (u * WV)) & 1u);
      if (lane == 0 && row0 + u * 32 < mpad) {
        uint32_t *dst = y + ((row0 >> 5) + u);
        *dst = clear ? bits : (*dst ^ bits);
      }
    }
  }
}

// ---------------------------------------------------------------------------
// batched mat-vec: Y (^)= X * A^T, i.e. row q of Y (packed, nrows(A) bits) = A * x_q for the
// nb <= BB packed vectors x_q = rows of X — ONE pass over A for all of them (the LPN workload's
// "batched mzd_mul_v", BASELINE configs[4]). HBM-bound like the single mat-vec: algorithmic
// bytes = 8 * (m*width + nb*width) + nb * ceil(m/64)*8.
// ---------------------------------------------------------------------------
// acc ^= (a & b), folded to 32 bits: one LOP3 per 32-bit word
__device__ __forceinline__ unsigned and_xor_fold(unsigned acc, const ulonglong2 &a, const ulonglong2 &b) {
  acc ^= (unsigned)a.x & (unsigned)b.x;
  acc ^= (unsigned)(a.x >> 32) & (unsigned)(b.x >> 32);
  acc ^= (unsigned)a.y & (unsigned)b.y;
  acc ^= (unsigned)(a.y >> 32) & (unsigned)(b.y >> 32);
  return acc;
}

// Short rows stored back to back (row stride == row length = WV 16-byte vectors): one THREAD per
// row. The row is read with 32-byte loads (whole sectors: no sector is fetched twice although
// neighbouring threads are WV*16 bytes apart), the BB vectors sit in shared memory and are read
// as broadcasts; per 32-bit word of the row and vector one LOP3 (acc ^= a & x) — the integer
// pipe, not HBM, is what bounds eight vectors per pass (128 LOP3 per 64-byte row). BB ballots
// turn the 32 rows x BB parities of a warp into one 32-bit word of every Y row.
template <int WV, int BB>
__global__ void __launch_bounds__(256)
mul_vb_rows_kernel(uint32_t *__restrict__ y, long long ys32, const ulonglong2 *__restrict__ A,
                   const ulonglong2 *__restrict__ X, int xs, long long m, int nb, int clear) {
  constexpr int RPT = 2;  // rows per thread: a warp owns 64 rows (rows base + lane and base + 32 + lane);
                          // every broadcast read of a vector word feeds two rows
  __shared__ ulonglong2 xsm[BB][WV];
  M4B_PDL_ENTER();
  for (int i = threadIdx.x; i < BB * WV; i += 256) {
    const int q = i / WV, w = i % WV;
    xsm[q][w] = q < nb ? X[(size_t)q * xs + w] : make_ulonglong2(0, 0);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  // (no grid-stride loop: with one the compiler hoists the BB*WV vector loads out of it into
  // 128 registers and the occupancy collapses)
  const long long base = (blockIdx.x * 8LL + (threadIdx.x >> 5)) * (32 * RPT);  // first row of the warp
  if (base >= ((m + 63) & ~63LL)) return;
  ulonglong2 a[RPT][WV];
#pragma unroll
  for (int j = 0; j < RPT; j++) {
    const long long row = base + 32 * j + lane;
    if (row < m) {
      const ulonglong2 *ar = A + row * WV;
      if constexpr (WV % 2 == 0) {
#pragma unroll
        for (int w = 0; w < WV; w += 2)
          asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
                       : "=l"(a[j][w].x), "=l"(a[j][w].y), "=l"(a[j][w + 1].x), "=l"(a[j][w + 1].y) : "l"(ar + w));
      } else {
#pragma unroll
        for (int w = 0; w < WV; w++)
          asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(a[j][w].x), "=l"(a[j][w].y) : "l"(ar + w));
      }
    } else {
#pragma unroll
      for (int w = 0; w < WV; w++) a[j][w] = make_ulonglong2(0, 0);
    }
  }
  unsigned mine[RPT];
#pragma unroll
  for (int j = 0; j < RPT; j++) mine[j] = 0;
#pragma unroll
  for (int q = 0; q < BB; q++) {
    unsigned acc[RPT];
#pragma unroll
    for (int j = 0; j < RPT; j++) acc[j] = 0;
#pragma unroll
    for (int w = 0; w < WV; w++) {
      const ulonglong2 xv = xsm[q][w];
#pragma unroll
      for (int j = 0; j < RPT; j++) acc[j] = and_xor_fold(acc[j], a[j][w], xv);
    }
#pragma unroll
    for (int j = 0; j < RPT; j++) mine[j] |= (__popc(acc[j]) & 1u) << q;
  }
  // the warp's 64 rows are whole 64-bit words of y (m is padded to 64 for the store): lane q
  // stores the two 32-bit words of Y row q
#pragma unroll
  for (int j = 0; j < RPT; j++) {
    unsigned mybits = 0;
#pragma unroll
    for (int q = 0; q < BB; q++) {
      const unsigned bits = __ballot_sync(0xffffffffu, (mine[j] >> q) & 1u);
      if (lane == q) mybits = bits;
    }
    if (lane < nb) {
      uint32_t *dst = y + (size_t)lane * ys32 + ((base >> 5) + j);
      *dst = clear ? mybits : (*dst ^ mybits);
    }
  }
}

// Any row length and stride: a CTA of 4 warps owns 32 consecutive rows (one aligned 32-bit word of
// every Y row), a warp walks its 8 rows together: per step nb 16-byte loads of the vectors (L1
// resident) feed 8 independent 16-byte loads of A.
template <int BB>
__global__ void __launch_bounds__(128)
mul_vb_long_kernel(uint32_t *__restrict__ y, long long ys32, const ulonglong2 *__restrict__ A,
                   const ulonglong2 *__restrict__ X, int xs, int m, int wv, int as, int nb, int clear) {
  constexpr int R = 8;
  __shared__ unsigned sbits[4][BB];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long row0 = blockIdx.x * 32LL + w * R;
  const ulonglong2 *ar[R];
#pragma unroll
  for (int r = 0; r < R; r++) {
    const long long row = row0 + r < m ? row0 + r : m - 1;  // rows past m: read a valid row, drop the bits
    ar[r] = A + (size_t)row * as;
  }
  unsigned acc[R][BB];
#pragma unroll
  for (int r = 0; r < R; r++)
#pragma unroll
    for (int q = 0; q < BB; q++) acc[r][q] = 0;
  for (int v = lane; v < wv; v += 32) {
    ulonglong2 a[R];
#pragma unroll
    for (int r = 0; r < R; r++)
      asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(a[r].x), "=l"(a[r].y) : "l"(ar[r] + v));
#pragma unroll
    for (int q = 0; q < BB; q++) {
      if (q < nb) {
        const ulonglong2 b = __ldg(&X[(size_t)q * xs + v]);
#pragma unroll
        for (int r = 0; r < R; r++) acc[r][q] = and_xor_fold(acc[r][q], a[r], b);
      }
    }
  }
#pragma unroll
  for (int q = 0; q < BB; q++) {
    unsigned bits = 0;
#pragma unroll
    for (int r = 0; r < R; r++) {
      const unsigned ball = __ballot_sync(0xffffffffu, __popc(acc[r][q]) & 1);
      if (row0 + r < m) bits |= (unsigned)(__popc(ball) & 1) << r;
    }
    if (lane == 0) sbits[w][q] = bits;
  }
  __syncthreads();
  if (threadIdx.x < nb) {
    unsigned word32 = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) word32 |= sbits[i][threadIdx.x] << (R * i);
    uint32_t *dst = y + (size_t)threadIdx.x * ys32 + blockIdx.x;
    *dst = clear ? word32 : (*dst ^ word32);
  }
}

template <int BB>
static void launch_mul_vb(uint32_t *yy, long long ys32, const DMat *A, const ulonglong2 *xx, int xs, int nb,
                          int clear, cudaStream_t s) {
  const int wv = (A->width + 1) / 2, as = A->rowstride / 2, mm = A->nrows;
  const ulonglong2 *AA = (const ulonglong2 *)A->d;
  if (wv <= 8 && as == wv && (wv & (wv - 1)) == 0 && (wv == 1 || !((uintptr_t)A->d & 31))) {
    const unsigned grid = cdiv(((long long)mm + 63) / 64 * 64, 512);  // 8 warps x 64 rows
    switch (wv) {
      case 1: launch_pdl(mul_vb_rows_kernel<1, BB>, grid, 256, s, yy, ys32, AA, xx, xs, mm, nb, clear); break;
      case 2: launch_pdl(mul_vb_rows_kernel<2, BB>, grid, 256, s, yy, ys32, AA, xx, xs, mm, nb, clear); break;
      case 4: launch_pdl(mul_vb_rows_kernel<4, BB>, grid, 256, s, yy, ys32, AA, xx, xs, mm, nb, clear); break;
      default: launch_pdl(mul_vb_rows_kernel<8, BB>, grid, 256, s, yy, ys32, AA, xx, xs, mm, nb, clear); break;
    }
  } else {
    const unsigned grid = cdiv(((long long)mm + 63) / 64 * 64, 32);  // whole 64-bit words of y
    mul_vb_long_kernel<BB><<<grid, 128, 0, s>>>(yy, ys32, AA, xx, xs, mm, wv, as, nb, clear);
  }
  M4B_LAUNCHED();
}

// Y (b x m) (^)= X (b x n) * A (m x n)^T, b small: passes of up to 8 vectors, each one read of A
void k_mul_v_batched(DMat *Y, const DMat *A, const DMat *X, int clear, cudaStream_t s) {
  const int b = X->nrows;
  if (b == 0 || A->nrows == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(Y, s);
    return;
  }
  if (A->rowstride % 2 || ((uintptr_t)A->d % 16) || ((uintptr_t)X->d % 16) || X->rowstride % 2)
    die("m4ri_b200: batched mul_v needs 16-byte aligned rows\n");
  if (b == 1) {  // the single-vector kernels
    DMat y1 = *Y, x1 = *X;
    y1.base = x1.base = nullptr;
    k_mul_v(&y1, A, &x1, clear, s);
    return;
  }
  timing_begin("mul_vb_kernel", s, 2.0 * A->nrows * (double)A->ncols * b);
  for (int q0 = 0; q0 < b; q0 += 8) {
    const int nb = b - q0 < 8 ? b - q0 : 8;
    uint32_t *yy = (uint32_t *)(Y->d + (size_t)q0 * Y->rowstride);
    const ulonglong2 *xx = (const ulonglong2 *)(X->d + (size_t)q0 * X->rowstride);
    if (nb <= 2) launch_mul_vb<2>(yy, Y->rowstride * 2LL, A, xx, X->rowstride / 2, nb, clear, s);
    else if (nb <= 4) launch_mul_vb<4>(yy, Y->rowstride * 2LL, A, xx, X->rowstride / 2, nb, clear, s);
    else launch_mul_vb<8>(yy, Y->rowstride * 2LL, A, xx, X->rowstride / 2, nb, clear, s);
  }
  timing_end(s);
}

void k_mul_v(DMat *y, const DMat *A, const DMat *x, int clear, cudaStream_t s) {
  if (A->nrows == 0 || A->ncols == 0) {
    if (clear) k_zero(y, s);
    return;
  }
  if (A->rowstride % 2 || ((uintptr_t)A->d % 16) || ((uintptr_t)x->d % 16))
    die("m4ri_b200: mul_v needs 16-byte aligned rows\n");
  // x is read in 16-byte vectors up to ceil(width/2): its row is padded to an even stride.
  int wv = (A->width + 1) / 2, as = A->rowstride / 2;
  const double ops = 2.0 * A->nrows * (double)A->ncols;
  if (wv <= 8) {  // thread per row; the last 64-bit word of y (rows past m) is written whole
    unsigned grid = cdiv(((long long)A->nrows + 63) / 64 * 64, 256);
    uint32_t *yy = (uint32_t *)y->d;
    const ulonglong2 *AA = (const ulonglong2 *)A->d, *xx = (const ulonglong2 *)x->d;
    const int mm = A->nrows;
    if (as == wv && (wv & (wv - 1)) == 0 && !mul_v_force_rows()) {  // rows back to back: coalesced warp-per-32-rows kernel
      const long long groups = ((long long)mm + 63) / 64 * 2;  // 32-row groups
      // one wave of resident CTAs, grid-stride inside the kernel: no tail wave
      // M4RI_B200_MULV_WAVES=w: cap the grid at w waves of resident CTAs (grid-stride loop inside the
      // kernel); 0 = one CTA per 8 x 32U rows (every warp runs one step)
      static const int waves = getenv("M4RI_B200_MULV_WAVES") ? atoi(getenv("M4RI_B200_MULV_WAVES")) : 0;
      auto grid_of = [&](int U, long long cap) {
        const long long all = cdiv(cdiv(groups, U), 8);
        return (unsigned)(waves > 0 ? min(all, cap * waves) : all);
      };
      timing_begin("mul_v_coal_kernel", s, ops);
      switch (wv) {
        case 1: launch_pdl(mul_v_coal_kernel<1, 8>, grid_of(8, resident_ctas(mul_v_coal_kernel<1, 8>)), 256, s, yy, AA, xx, mm, clear); break;
        case 2: launch_pdl(mul_v_coal_kernel<2, 4>, grid_of(4, resident_ctas(mul_v_coal_kernel<2, 4>)), 256, s, yy, AA, xx, mm, clear); break;
        case 4: launch_pdl(mul_v_coal_kernel<4, 2>, grid_of(2, resident_ctas(mul_v_coal_kernel<4, 2>)), 256, s, yy, AA, xx, mm, clear); break;
        default: launch_pdl(mul_v_coal_kernel<8, 1>, grid_of(1, resident_ctas(mul_v_coal_kernel<8, 1>)), 256, s, yy, AA, xx, mm, clear); break;
      }
      M4B_LAUNCHED();
      timing_end(s);
      return;
    }
    timing_begin("mul_v_rows_kernel", s, ops);
    switch (wv) {
      case 1: mul_v_rows_kernel<1><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
      case 2: mul_v_rows_kernel<2><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
      case 3: mul_v_rows_kernel<3><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
      case 4: mul_v_rows_kernel<4><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
      case 5: mul_v_rows_kernel<5><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
      case 6: mul_v_rows_kernel<6><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
      case 7: mul_v_rows_kernel<7><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
      default: mul_v_rows_kernel<8><<<grid, 256, 0, s>>>(yy, AA, xx, mm, as, clear); break;
    }
    M4B_LAUNCHED();
    timing_end(s);
    return;
  }
  if (wv >= 32 && A->nrows >= 2048 && !mul_v_force_rows()) {
    // whole 64-bit words of y: an even number of 32-row CTAs
    unsigned grid = cdiv(((long long)A->nrows + 63) / 64 * 64, 32);
    timing_begin("mul_v_long_kernel", s, ops);
    mul_v_long_kernel<<<grid, 256, 0, s>>>((uint32_t *)y->d, (const ulonglong2 *)A->d,
                                           (const ulonglong2 *)x->d, A->nrows, wv, as, clear);
    M4B_LAUNCHED();
    timing_end(s);
    return;
  }
  if (clear) k_zero(y, s);
  int G = wv >= 32 ? 32 : wv >= 16 ? 16 : wv >= 8 ? 8 : wv >= 4 ? 4 : wv >= 2 ? 2 : 1;
  long long warps = cdiv(A->nrows, 32 / G);
  unsigned grid = cdiv(warps * 32, 256);
  uint32_t *yy = (uint32_t *)y->d;
  const ulonglong2 *AA = (const ulonglong2 *)A->d, *xx = (const ulonglong2 *)x->d;
  timing_begin("mul_v_kernel", s, ops);
  switch (G) {
    case 32: mul_v_kernel<32><<<grid, 256, 0, s>>>(yy, AA, xx, A->nrows, wv, as); break;
    case 16: mul_v_kernel<16><<<grid, 256, 0, s>>>(yy, AA, xx, A->nrows, wv, as); break;
    case 8: mul_v_kernel<8><<<grid, 256, 0, s>>>(yy, AA, xx, A->nrows, wv, as); break;
    case 4: mul_v_kernel<4><<<grid, 256, 0, s>>>(yy, AA, xx, A->nrows, wv, as); break;
    case 2: mul_v_kernel<2><<<grid, 256, 0, s>>>(yy, AA, xx, A->nrows, wv, as); break;
    default: mul_v_kernel<1><<<grid, 256, 0, s>>>(yy, AA, xx, A->nrows, wv, as); break;
  }
  M4B_LAUNCHED();
  timing_end(s);
}

// ---------------------------------------------------------------------------
// vec-mat: C (^)= v*A = XOR of the rows of A selected by the bits of v.
// grid = (column chunks, k chunks, rows of v); partial XORs merge with atomicXor.
// ---------------------------------------------------------------------------
constexpr int VA_ROWS = 256;
__global__ void __launch_bounds__(256)
mul_va_kernel(word *__restrict__ C, const word *__restrict__ v, const word *__restrict__ A, int k,
              int width, int cs, int vs, int as) {
  int w = blockIdx.x * blockDim.x + threadIdx.x;
  int l0 = blockIdx.y * VA_ROWS, l1 = min(k, l0 + VA_ROWS);
  int i = blockIdx.z;
  if (w >= width) return;
  word acc = 0;
  for (int lw = l0; lw < l1; lw += 64) {
    word bits = v[(size_t)i * vs + lw / 64];
    int cnt = min(64, l1 - lw);
    if (cnt < 64) bits &= (((word)1 << cnt) - 1);
    while (bits) {
      int b = __ffsll((long long)bits) - 1;
      bits &= bits - 1;
      acc ^= A[(size_t)(lw + b) * as + w];
    }
  }
  if (acc) atomicXor((unsigned long long *)&C[(size_t)i * cs + w], (unsigned long long)acc);
}

void k_mul_va(DMat *C, const DMat *v, const DMat *A, int clear, cudaStream_t s) {
  if (clear) k_zero(C, s);
  if (v->nrows == 0 || A->nrows == 0 || A->ncols == 0) return;
  dim3 grid(cdiv(A->width, 256), cdiv(A->nrows, VA_ROWS), v->nrows);
  timing_begin("mul_va_kernel", s, 2.0 * v->nrows * (double)A->nrows * A->ncols);
  mul_va_kernel<<<grid, 256, 0, s>>>(C->d, v->d, A->d, A->nrows, A->width, C->rowstride,
                                     v->rowstride, A->rowstride);
  M4B_LAUNCHED();
  timing_end(s);
}

// ---------------------------------------------------------------------------
// C (^)= A * Bt^T by AND + popcount parity: 64x64 output tile per CTA, 4x4
// outputs per thread, k staged through shared memory 16 words at a time.
// ---------------------------------------------------------------------------
constexpr int SK = 16;
__global__ void __launch_bounds__(256)
mul_abt_simple_kernel(word *__restrict__ C, const word *__restrict__ A, const word *__restrict__ Bt,
                      int m, int n, int kw, int cs, int as, int bs, int clear) {
  __shared__ word As[64][SK + 1];
  __shared__ word Bs[64][SK + 1];
  __shared__ unsigned long long Cs[64];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  word acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; a++)
#pragma unroll
    for (int b = 0; b < 4; b++) acc[a][b] = 0;
  if (threadIdx.x < 64) Cs[threadIdx.x] = 0;
  for (int k0 = 0; k0 < kw; k0 += SK) {
    for (int idx = threadIdx.x; idx < 64 * SK; idx += 256) {
      int r = idx / SK, c = idx % SK;
      As[r][c] = (i0 + r < m && k0 + c < kw) ? A[(size_t)(i0 + r) * as + k0 + c] : 0;
      Bs[r][c] = (j0 + r < n && k0 + c < kw) ? Bt[(size_t)(j0 + r) * bs + k0 + c] : 0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SK; kk++) {
      word av[4], bv[4];
#pragma unroll
      for (int a = 0; a < 4; a++) av[a] = As[ty + 16 * a][kk];
#pragma unroll
      for (int b = 0; b < 4; b++) bv[b] = Bs[tx + 16 * b][kk];
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) acc[a][b] ^= av[a] & bv[b];
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < 4; a++) {
    unsigned long long bits = 0;
#pragma unroll
    for (int b = 0; b < 4; b++)
      bits |= (unsigned long long)(__popcll(acc[a][b]) & 1) << (tx + 16 * b);
    if (bits) atomicOr(&Cs[ty + 16 * a], bits);
  }
  __syncthreads();
  if (threadIdx.x < 64 && i0 + threadIdx.x < m) {
    word *dst = &C[(size_t)(i0 + threadIdx.x) * cs + (j0 >> 6)];
    *dst = clear ? Cs[threadIdx.x] : (*dst ^ Cs[threadIdx.x]);
  }
}

void k_mul_abt_simple(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    return;
  }
  dim3 grid(cdiv(C->ncols, 64), cdiv(C->nrows, 64));
  mul_abt_simple_kernel<<<grid, 256, 0, s>>>(C->d, A->d, Bt->d, C->nrows, C->ncols, A->width,
                                             C->rowstride, A->rowstride, Bt->rowstride, clear);
  M4B_LAUNCHED();
}

}  // namespace m4b
