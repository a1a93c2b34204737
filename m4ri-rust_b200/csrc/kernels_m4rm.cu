// kernels_m4rm.cu — path (a): Method of the Four Russians in shared memory.
//
// Replaces what M4RI does behind mzd_mul_m4rm / mzd_addmul_m4rm
// (m4ri-sys/src/brilliantrussian.rs:216,224; mechanism documented at
// brilliantrussian.rs:17-193: mzd_make_table builds the 2^k XOR-combinations of k
// rows of B, mzd_process_rows* look k bits of A up and XOR the table row into C).
//
// B200 design (not M4RI's): one CTA owns a 1024 x 1024-bit tile of C, held
// entirely in registers (1024 threads x 32 accumulators: thread (warp w, lane l)
// owns the 32-bit column word l of the tile rows 32w..32w+31). Per pass the CTA
// consumes 32 rows of B: it builds four 256-entry tables (8 rows of B each,
// 4 x 256 x 128 B = 128 KiB of shared memory), then every warp walks its 32
// rows: one A word per row (read once per lane, broadcast with a shuffle) gives
// four byte indices, each a conflict-free 128-byte table-row read (lane l reads
// bank l) that is XORed into the accumulators with LOP3.
//
// Roofline: shared-memory wavefronts. Per (row, pass) the warp
// issues 4 table reads + 1 shuffle = 5 crossbar wavefronts for 32 x 1024 bit-MACs;
// the table build adds 1024 store wavefronts per pass per CTA (25 % at 1024 rows).
#include "common.cuh"

namespace m4b {

constexpr int TM = 1024;          // tile rows
constexpr int TNW = 32;           // tile columns in 32-bit words (1024 bits)
constexpr int NTAB = 4;           // tables per pass (8 rows of B each)
constexpr int M4RM_SMEM = NTAB * 256 * TNW * 4;  // 131072

__global__ void __launch_bounds__(1024, 1)
m4rm_kernel(uint32_t *__restrict__ C, const uint32_t *__restrict__ A,
            const uint32_t *__restrict__ B, int m, int k, int aw32, int bw32, int cw32, int as32,
            int bs32, int cs32, int clear) {
  extern __shared__ __align__(16) uint32_t tab[];  // [NTAB][256][TNW]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int row0 = blockIdx.y * TM + warp * 32;  // first C row of this warp
  const int cw = blockIdx.x * TNW + lane;        // column word of this lane
  const bool col_ok = cw < bw32;

  uint32_t acc[32];
#pragma unroll
  for (int r = 0; r < 32; r++) acc[r] = 0;

  // table-build role of this warp: table bt, entries [be0, be0+32)
  const int bt = warp >> 3, be0 = (warp & 7) * 32;
  uint32_t *mytab = tab + (bt * 256 + be0) * TNW + lane;
  const uint32_t *tl = tab + lane;

  const int my_row = row0 + lane;
  const uint32_t *arow = A + (size_t)my_row * as32;
  const int npass = (k + 31) >> 5;
  uint32_t a_next = (my_row < m && 0 < aw32) ? arow[0] : 0;

  for (int p = 0; p < npass; p++) {
    // ---- build: 8 rows of B -> 32 of the 256 combinations (this warp's share) ----
    const int kb = p * 32 + bt * 8;
    uint32_t b[8];
#pragma unroll
    for (int j = 0; j < 8; j++)
      b[j] = (col_ok && kb + j < k) ? B[(size_t)(kb + j) * bs32 + cw] : 0;
    // high three index bits are fixed per warp, low five run in Gray-code order
    uint32_t cur = ((warp & 1) ? b[5] : 0) ^ ((warp & 2) ? b[6] : 0) ^ ((warp & 4) ? b[7] : 0);
    mytab[0] = cur;
#pragma unroll
    for (int g = 1; g < 32; g++) {
      const int bit = __ffs(g) - 1;           // Gray code: step g flips bit ctz(g)
      const int idx = g ^ (g >> 1);           // the index reached after the flip
      cur ^= b[bit];
      mytab[idx * TNW] = cur;
    }
    const uint32_t a_mine = a_next;
    if (p + 1 < npass) a_next = (my_row < m && p + 1 < aw32) ? arow[p + 1] : 0;
    __syncthreads();
    // ---- gather: 32 rows x 4 lookups per warp ----
#pragma unroll
    for (int r = 0; r < 32; r++) {
      const uint32_t a = __shfl_sync(0xffffffffu, a_mine, r);
      const uint32_t t0 = tl[((a & 0xff)) * TNW];
      const uint32_t t1 = tl[(256 + ((a >> 8) & 0xff)) * TNW];
      const uint32_t t2 = tl[(512 + ((a >> 16) & 0xff)) * TNW];
      const uint32_t t3 = tl[(768 + (a >> 24)) * TNW];
      acc[r] ^= t0 ^ t1 ^ t2 ^ t3;
    }
    __syncthreads();
  }
  if (cw < cw32) {
#pragma unroll
    for (int r = 0; r < 32; r++) {
      const int row = row0 + r;
      if (row < m) {
        uint32_t *dst = C + (size_t)row * cs32 + cw;
        *dst = clear ? acc[r] : (*dst ^ acc[r]);
      }
    }
  }
}

void k_mul_m4rm(DMat *C, const DMat *A, const DMat *B, int clear, cudaStream_t s) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    return;
  }
  static bool attr_set = false;
  if (!attr_set) {
    M4B_CUDA(cudaFuncSetAttribute(m4rm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  M4RM_SMEM));
    attr_set = true;
  }
  dim3 grid((C->width * 2 + TNW - 1) / TNW, (C->nrows + TM - 1) / TM);
  timing_begin("m4rm_kernel", s);
  m4rm_kernel<<<grid, 1024, M4RM_SMEM, s>>>((uint32_t *)C->d, (const uint32_t *)A->d,
                                            (const uint32_t *)B->d, C->nrows, A->ncols,
                                            A->width * 2, B->width * 2, C->width * 2,
                                            A->rowstride * 2, B->rowstride * 2, C->rowstride * 2,
                                            clear);
  M4B_LAUNCHED();
  timing_end(s);
}

}  // namespace m4b
