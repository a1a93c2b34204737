// kernels_m4rm.cu — path (a): Method of the Four Russians in shared memory.
//
// Replaces what M4RI does behind mzd_mul_m4rm / mzd_addmul_m4rm
// (m4ri-sys/src/brilliantrussian.rs:216,224; mechanism documented at
// brilliantrussian.rs:17-193: mzd_make_table builds the 2^k XOR-combinations of k
// rows of B in Gray-code order, mzd_process_rows* look k bits of A up and XOR
// the table row into C).
//
// B200 design (not M4RI's): one CTA owns a 1024 x 1024-bit tile of C, held
// entirely in registers (512 threads x 64 accumulator words; the other ~60
// registers keep eight LDS.128 in flight per warp). k is consumed in
// phases of 16 bits: two 256-entry tables (8 rows of B each) of 128-byte rows,
// double buffered (2 x 64 KiB of shared memory) so every phase both gathers from
// the current tables and builds the next ones — the LSU never idles on a
// build-only or gather-only stretch.
//
// Gather: a quarter-warp (8 lanes x 16 B) reads one 128-byte table row with
// LDS.128 — one conflict-free wavefront — so the four quarters of a warp follow
// FOUR different rows of A per instruction. Per (row, phase): 2 table reads,
// 1/4 shuffle, ~4 integer ops for the two byte indices, 4 LOP3 (3-input XOR).
// Build: lane = 32-bit column word, warp = (table, high nibble); the low nibble
// runs in Gray-code order: one LOP3 + one conflict-free STS.32 per entry.
//
// Roofline: shared-memory wavefronts (128 B/clk/SM measured). Per
// phase and CTA: 2048 gather + 256 shuffle + 512 build wavefronts for
// 1024 x 1024 x 16 bit-MACs.
#include "common.cuh"

namespace m4b {

constexpr int TM = 1024;   // tile rows
constexpr int TNW = 32;    // tile columns in 32-bit words (1024 bits)
constexpr int NTHR = 512;  // 16 warps x 64 rows; 64 accumulator registers per thread
constexpr int TAB_BYTES = 256 * TNW * 4;         // one table: 256 rows of 128 B = 32 KiB
constexpr int ASTAGE_BYTES = TM * 16;            // A staging: 16 B (8 phases of k) per row
constexpr int M4RM_SMEM = 2 * 2 * TAB_BYTES + 2 * ASTAGE_BYTES + 2 * 2048;  // 128 + 32 + 4 KiB

__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(saddr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t saddr) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(saddr));
  return v;
}
// 16-byte global -> shared copy that bypasses registers; bytes beyond src_bytes are zero-filled
__device__ __forceinline__ void cp_async16(uint32_t saddr, const void *gptr, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(saddr), "l"(gptr), "r"(src_bytes)
               : "memory");
}

__global__ void __launch_bounds__(NTHR, 1)
m4rm_kernel(uint32_t *__restrict__ C, const uint32_t *__restrict__ A,
            const uint32_t *__restrict__ B, int m, int k, int aw32, int bw32, int cw32, int as32,
            int bs32, int cs32, int clear) {
  extern __shared__ __align__(16) uint32_t tab[];  // [buf 2][table 2][256][TNW] | A stages
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q = lane >> 3, l8 = lane & 7;
  const int row0 = blockIdx.y * TM + warp * 64;  // first C row of this warp

  uint4 acc[16];
#pragma unroll
  for (int s = 0; s < 16; s++) acc[s] = make_uint4(0, 0, 0, 0);

  // ---- build role: table bt, entries [bhi*32, bhi*32+32), column word `lane` ----
  // The 16 rows of B a phase needs are staged in shared memory by cp.async two
  // phases ahead (no registers held across the gather, no exposed LDG latency).
  const int bt = warp >> 3, bhi = warp & 7;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(tab);
  const uint32_t bstage = sbase + 4 * TAB_BYTES + 2 * ASTAGE_BYTES;  // [2][16 rows][128 B]
  auto stage_b = [&](int ph) {
    if (threadIdx.x < 128) {
      const int j = threadIdx.x >> 3, c = threadIdx.x & 7;
      const int krow = ph * 16 + j, cw0 = blockIdx.x * TNW + c * 4;
      int bytes = (krow < k) ? (bw32 - cw0) * 4 : 0;
      bytes = bytes < 0 ? 0 : (bytes > 16 ? 16 : bytes);
      const uint32_t *src = B + (bytes ? (size_t)krow * bs32 + cw0 : 0);
      cp_async16(bstage + (ph & 1) * 2048 + j * 128 + c * 16, src, bytes);
    }
  };
  auto build = [&](int ph) {
    const uint32_t *bs = tab + (4 * TAB_BYTES + 2 * ASTAGE_BYTES) / 4 + (ph & 1) * 512 + bt * 256 + lane;
    const uint32_t b0 = bs[0 * TNW], b1 = bs[1 * TNW], b2 = bs[2 * TNW], b3 = bs[3 * TNW],
                   b4 = bs[4 * TNW], b5 = bs[5 * TNW], b6 = bs[6 * TNW], b7 = bs[7 * TNW];
    uint32_t *dst = tab + ((ph & 1) * 2 + bt) * (TAB_BYTES / 4) + (bhi * 32) * TNW + lane;
    uint32_t cur = ((bhi & 1) ? b5 : 0u) ^ ((bhi & 2) ? b6 : 0u) ^ ((bhi & 4) ? b7 : 0u);
    dst[0] = cur;
    // low five index bits in Gray-code order: step g flips bit ctz(g) and lands on
    // index g ^ (g >> 1); one XOR and one conflict-free store per entry
#define M4B_STEP(g, bb) cur ^= bb; dst[((g) ^ ((g) >> 1)) * TNW] = cur;
    M4B_STEP(1, b0) M4B_STEP(2, b1) M4B_STEP(3, b0) M4B_STEP(4, b2) M4B_STEP(5, b0) M4B_STEP(6, b1)
    M4B_STEP(7, b0) M4B_STEP(8, b3) M4B_STEP(9, b0) M4B_STEP(10, b1) M4B_STEP(11, b0) M4B_STEP(12, b2)
    M4B_STEP(13, b0) M4B_STEP(14, b1) M4B_STEP(15, b0) M4B_STEP(16, b4) M4B_STEP(17, b0)
    M4B_STEP(18, b1) M4B_STEP(19, b0) M4B_STEP(20, b2) M4B_STEP(21, b0) M4B_STEP(22, b1)
    M4B_STEP(23, b0) M4B_STEP(24, b3) M4B_STEP(25, b0) M4B_STEP(26, b1) M4B_STEP(27, b0)
    M4B_STEP(28, b2) M4B_STEP(29, b0) M4B_STEP(30, b1) M4B_STEP(31, b0)
#undef M4B_STEP
  };

  // ---- A staging: group g = phases [8g, 8g+8) = 16 bytes of every tile row ----
  const uint32_t abase = sbase + 4 * TAB_BYTES;
  const int nph = (k + 15) >> 4;
  const int ngrp = (nph + 7) >> 3;
  auto stage_a = [&](int g) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int r = threadIdx.x + h * NTHR;
      const int grow = blockIdx.y * TM + r;
      int bytes = (grow < m) ? (aw32 - g * 4) * 4 : 0;
      bytes = bytes < 0 ? 0 : (bytes > 16 ? 16 : bytes);
      const uint32_t *src = A + (size_t)(grow < m ? grow : 0) * as32 + (bytes ? g * 4 : 0);
      cp_async16(abase + (g & 1) * ASTAGE_BYTES + r * 16, src, bytes);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  // ---- gather role: rows row0 + 4s + q (s = 0..15), 16-byte column chunk l8 ----
  const uint32_t gbase = sbase + l8 * 16;
  const uint32_t arow_s = abase + (warp * 64 + q) * 16;  // + s*64 + stage + (ph&7)*2

  stage_b(0);
  if (nph > 1) stage_b(1);
  stage_a(0);  // commits the group
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  build(0);
  __syncthreads();
  if (nph > 2) stage_b(2);  // slot 0 is free again once build(0) has read it
  asm volatile("cp.async.commit_group;" ::: "memory");

  for (int ph = 0; ph < nph; ph++) {
    if (ph + 1 < nph) build(ph + 1);

    const uint32_t t0 = gbase + (ph & 1) * 2 * TAB_BYTES;
    const uint32_t ap = arow_s + ((ph >> 3) & 1) * ASTAGE_BYTES + (ph & 7) * 2;
    // four batches of four steps: 4 index loads, then 8 independent LDS.128 in flight
#pragma unroll
    for (int bch = 0; bch < 4; bch++) {
      uint32_t a[4];
#pragma unroll
      for (int i = 0; i < 4; i++) a[i] = lds_u16(ap + (bch * 4 + i) * 64);
      uint4 x[4], y[4];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        x[i] = lds128(t0 + ((a[i] & 0xffu) << 7));
        y[i] = lds128(t0 + TAB_BYTES + ((a[i] >> 8) << 7));
      }
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int s = bch * 4 + i;
        acc[s].x ^= x[i].x ^ y[i].x;
        acc[s].y ^= x[i].y ^ y[i].y;
        acc[s].z ^= x[i].z ^ y[i].z;
        acc[s].w ^= x[i].w ^ y[i].w;
      }
    }
    // stage B rows of phase ph+3 (their slot was last read by build(ph+1), a barrier
    // ago) and, every 8 phases, the next 16 bytes of every A row; all of it has the
    // whole next phase to land.
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (ph + 3 < nph) stage_b(ph + 3);
    if ((ph & 7) == 0 && (ph >> 3) + 1 < ngrp) stage_a((ph >> 3) + 1);
    else asm volatile("cp.async.commit_group;" ::: "memory");
  }

  // ---- epilogue: 16-byte stores, word-guarded at the right edge ----
  const int cw = blockIdx.x * TNW + l8 * 4;
#pragma unroll
  for (int s = 0; s < 16; s++) {
    const int row = row0 + 4 * s + q;
    if (row >= m || cw >= cw32) continue;
    uint32_t *dst = C + (size_t)row * cs32 + cw;
    if (cw + 4 <= cw32) {
      uint4 v = acc[s];
      if (!clear) {
        const uint4 o = *reinterpret_cast<const uint4 *>(dst);
        v.x ^= o.x; v.y ^= o.y; v.z ^= o.z; v.w ^= o.w;
      }
      *reinterpret_cast<uint4 *>(dst) = v;
    } else {
      const uint32_t v[4] = {acc[s].x, acc[s].y, acc[s].z, acc[s].w};
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (cw + j < cw32) dst[j] = clear ? v[j] : (dst[j] ^ v[j]);
    }
  }
}

void k_mul_m4rm(DMat *C, const DMat *A, const DMat *B, int clear, cudaStream_t s) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    return;
  }
  if (((uintptr_t)C->d & 15) || (C->rowstride & 1) || ((uintptr_t)A->d & 15) || (A->rowstride & 1))
    die("m4ri_b200: m4rm kernel needs 16-byte aligned rows of A and C\n");
  static std::atomic<unsigned long long> attr_seen{0};
  if (first_use_on_this_device(attr_seen)) {
    M4B_CUDA(cudaFuncSetAttribute(m4rm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  M4RM_SMEM));
  }
  dim3 grid((C->width * 2 + TNW - 1) / TNW, (C->nrows + TM - 1) / TM);
  timing_begin("m4rm_kernel", s, 2.0 * C->nrows * C->ncols * A->ncols);
  m4rm_kernel<<<grid, NTHR, M4RM_SMEM, s>>>((uint32_t *)C->d, (const uint32_t *)A->d,
                                            (const uint32_t *)B->d, C->nrows, A->ncols,
                                            A->width * 2, B->width * 2, C->width * 2,
                                            A->rowstride * 2, B->rowstride * 2, C->rowstride * 2,
                                            clear);
  M4B_LAUNCHED();
  timing_end(s);
}

}  // namespace m4b
