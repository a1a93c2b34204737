// noalloc.cu — C (^)= A*B on device operands without touching any allocator.
//
// Every other product entry point of this library gets its temporaries from the CUDA
// stream-ordered pool (dmat_new -> cudaMallocAsync): m4ri_b200_dmul needs B^T, n x k, because the
// tensor kernels contract two k-contiguous operands. Some hosts of this library cannot let it do
// that:
//   * a process whose device memory is already spoken for — another runtime's caching allocator
//     (a torch program handing us its tensors through m4ri_b200_dmat_wrap) or a service that sized
//     its matrices to fill the 180 GB of the part — where a pool allocation inside the library
//     either fails or takes memory the owner counted on;
//   * a latency-bound loop (many small LPN / decoding products per second) that must not meet the
//     implicit device-wide synchronisation of cudaMalloc/cudaFree or a pool that grows mid-loop.
// For them the B^T temporary lives in ONE fixed array in the library's device image. The price:
// a stated maximum size, 8 MiB held on every GPU that loads the library until the process exits,
// and no two calls in flight at once on one GPU (they would share the array).
#include <mutex>

#include "common.cuh"

namespace m4b {

constexpr size_t SCRATCH_WORDS = M4RI_B200_NOALLOC_SCRATCH_WORDS;

// The scratch itself. In host code `g_bt_scratch` is only a handle for the runtime's symbol
// table — its host-side address is NOT device memory; the kernels get the address that
// cudaGetSymbolAddress reports for the current device (scratch_on_current_device below).
static __device__ __align__(256) word g_bt_scratch[SCRATCH_WORDS];

static std::mutex g_scratch_mutex;
static word *g_scratch_ptr[64];  // per device ordinal, filled on the first call made on that device

static word *scratch_on_current_device() {
  int dev = -1;
  M4B_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) die("m4ri_b200: mul_noalloc: device ordinal %d is out of range\n", dev);
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_ptr[dev]) {
    // every device has its own instance of the array: resolve it with THIS device current
    void *p = nullptr;
    M4B_CUDA(cudaGetSymbolAddress(&p, g_bt_scratch));
    if (!p || ((uintptr_t)p & 255)) die("m4ri_b200: mul_noalloc: scratch symbol resolved to %p\n", p);
    g_scratch_ptr[dev] = (word *)p;
  }
  return g_scratch_ptr[dev];
}

static DMat view(const word *d, rci_t nrows, rci_t ncols, wi_t rowstride) {
  DMat M;
  M.d = const_cast<word *>(d);
  M.nrows = nrows;
  M.ncols = ncols;
  M.width = width_of(ncols);
  M.rowstride = rowstride;
  M.base = nullptr;
  return M;
}

}  // namespace m4b

using namespace m4b;

extern "C" int m4ri_b200_mul_noalloc(word *C, wi_t c_rowstride, const word *A, wi_t a_rowstride, const word *B,
                                     wi_t b_rowstride, rci_t m, rci_t k, rci_t n, int clear, void *stream) {
  if (m < 0 || k < 0 || n < 0) {
    fprintf(stderr, "m4ri_b200: mul_noalloc: negative dimension (%d x %d x %d)\n", m, k, n);
    return -3;
  }
  if (m == 0 || n == 0) return 0;
  const wi_t wa = width_of(k), wb = width_of(n);
  if (!C || (k > 0 && (!A || !B)) || c_rowstride < wb || (k > 0 && (a_rowstride < wa || b_rowstride < wb))) {
    fprintf(stderr, "m4ri_b200: mul_noalloc: null operand or a row stride below the row width\n");
    return -3;
  }
  if ((((uintptr_t)C | (uintptr_t)A | (uintptr_t)B) & 15) || ((c_rowstride | a_rowstride | b_rowstride) & 1)) {
    fprintf(stderr, "m4ri_b200: mul_noalloc: operands need 16-byte aligned rows (aligned base, even row stride)\n");
    return -2;
  }
  // B^T: n rows of ceil(k/64) words, stride rounded to 32 bytes so the 256-bit loads apply
  const wi_t bt_stride = (wa + 3) & ~3;
  const size_t need = (size_t)n * (size_t)(bt_stride ? bt_stride : 4);
  if (need > SCRATCH_WORDS) {
    fprintf(stderr,
            "m4ri_b200: mul_noalloc: B is %d x %d; its transpose needs %zu words of scratch, the built-in "
            "array holds %zu (8192 x 8192 bits). Use m4ri_b200_dmul, which allocates.\n",
            k, n, need, SCRATCH_WORDS);
    return -1;
  }
  cudaStream_t s = stream ? (cudaStream_t)stream : current_stream();
  DMat dC = view(C, m, n, c_rowstride);
  if (k == 0) {
    if (clear) k_zero(&dC, s);
    return 0;
  }
  DMat dA = view(A, m, k, a_rowstride), dB = view(B, k, n, b_rowstride);
  word *scratch = scratch_on_current_device();
  DMat dBt = view(scratch, n, k, bt_stride);
  // The array still holds the previous call's B^T. The transpose writes the ceil(k/64) valid
  // words of every row; the stride padding after them is read by the wide loads and has to be zero.
  M4B_CUDA(cudaMemsetAsync(scratch, 0, need * sizeof(word), s));
  k_transpose(&dBt, &dB, s);
  // the choice m4ri_b200_dmul makes, minus the Strassen levels (their temporaries are allocations)
  const bool small = (long long)m * n <= 64LL * 64 || k <= 64;
  if (!small && mma_available()) k_mul_abt_mma(&dC, &dA, &dBt, clear, s);
  else k_mul_abt_simple(&dC, &dA, &dBt, clear, s);
  return 0;
}
