// kernels_mma.cu — path (b): the product as a dense contraction on the 5th-gen
// tensor cores. C (^)= A * Bt^T with A (m x k bits) and Bt (n x k bits) both
// k-contiguous, the layout _mzd_mul_naive takes (m4ri-sys/src/mzd.rs:163-168);
// mzd_mul gets here through one bit-transpose of B (api.cu).
//
// Bits are unpacked to 0/1 bytes inside the SM (never in HBM: 8x less L2/HBM
// traffic than an int8 GEMM on pre-expanded operands), multiplied by
// tcgen05.mma.kind::i8 with exact int32 accumulation in TMEM, and the parity
// (bit 0) of every accumulator is packed back into 64-bit words. Any k permutation
// applied identically to both operands leaves the product unchanged, so a 32-bit
// word w of k-bits expands to eight words (w >> s) & 0x01010101, s = 0..7 — two
// integer instructions per four elements, no table, no multiply.
//
// v1 (this file): one CTA per 128 x 128 tile, both operands expanded into
// 128B-swizzled K-major shared-memory tiles (SS form), 4-stage mbarrier ring:
//   warps 0-3  expand A rows (thread = row), then run the epilogue (TMEM lanes)
//   warps 4-7  expand Bt rows
//   warp  8    allocates TMEM, one elected lane issues tcgen05.mma / commit
#include "common.cuh"

namespace m4b {

namespace {

constexpr int BM = 128, BN = 128;      // C tile
constexpr int KB_BITS = 128;           // k bits per pipeline stage = one 128-byte swizzle row
constexpr int STAGES = 4;
constexpr int A_TILE = BM * 128;       // expanded bytes per stage
constexpr int B_TILE = BN * 128;
constexpr int MMA_THREADS = 288;       // 8 expander warps + 1 MMA warp
constexpr int SMEM_BYTES = STAGES * (A_TILE + B_TILE) + 1024 /*align*/ + 256 /*barriers*/;

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(bar), "r"(parity)
      : "memory");
}
// generic-proxy shared-memory writes -> visible to the async proxy (tensor core reads)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], int8 x int8 -> int32
__device__ __forceinline__ void umma_i8_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte swizzle shared-memory matrix descriptor (sm_100 format: start
// address >> 4 in bits [0,14), stride byte offset >> 4 in [32,46) = 1024 B between
// 8-row groups, version 1 in [46,48), layout SWIZZLE_128B = 2 in [61,64)).
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}
// instruction descriptor, kind::i8: D = S32 (2 << 4), A/B unsigned 8-bit K-major,
// N >> 3 in [17,23), M >> 4 in [24,29)
__host__ __device__ constexpr uint32_t make_idesc_i8(int M, int N) {
  return (2u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// One thread expands 128 k-bits of its row into the 128-byte swizzled row of the
// stage tile: eight 16-byte chunks, chunk c stored at c ^ (row & 7).
__device__ __forceinline__ void expand_row(uint32_t tile_saddr, int r, uint4 w) {
  const uint32_t rowaddr = tile_saddr + (r >> 3) * 1024 + (r & 7) * 128;
  const uint32_t sw = (uint32_t)(r & 7);
  const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const uint32_t x = ws[j];
    uint4 lo, hi;
    lo.x = x & 0x01010101u;
    lo.y = (x >> 1) & 0x01010101u;
    lo.z = (x >> 2) & 0x01010101u;
    lo.w = (x >> 3) & 0x01010101u;
    hi.x = (x >> 4) & 0x01010101u;
    hi.y = (x >> 5) & 0x01010101u;
    hi.z = (x >> 6) & 0x01010101u;
    hi.w = (x >> 7) & 0x01010101u;
    const uint32_t a0 = rowaddr + (((uint32_t)(2 * j) ^ sw) << 4);
    const uint32_t a1 = rowaddr + (((uint32_t)(2 * j + 1) ^ sw) << 4);
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a0), "r"(lo.x), "r"(lo.y), "r"(lo.z),
                 "r"(lo.w)
                 : "memory");
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a1), "r"(hi.x), "r"(hi.y), "r"(hi.z),
                 "r"(hi.w)
                 : "memory");
  }
}

__global__ void __launch_bounds__(MMA_THREADS, 1)
mma_i8_kernel(uint32_t *__restrict__ C, const uint4 *__restrict__ A, const uint4 *__restrict__ Bt,
              int m, int n, int k, int as16, int bs16, int cs32, int cw32, int tiles_m, int tiles_n,
              int clear) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B needs 1024-B alignment
  const uint32_t a_tiles = sbase, b_tiles = sbase + STAGES * A_TILE;
  const uint32_t bars = b_tiles + STAGES * B_TILE;
  const uint32_t full_bar = bars, empty_bar = bars + 8 * STAGES, accum_bar = bars + 16 * STAGES;
  const uint32_t tmem_slot = accum_bar + 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // tile of this CTA: groups of 8 m-tiles sweep the n-tiles together (L2 reuse)
  const int t = blockIdx.x;
  const int grp = t / (8 * tiles_n), in_grp = t % (8 * tiles_n);
  const int grp_rows = min(8, tiles_m - grp * 8);
  const int tm = grp * 8 + in_grp % grp_rows, tn = in_grp / grp_rows;
  const int m0 = tm * BM, n0 = tn * BN;
  const int nkb = (k + KB_BITS - 1) / KB_BITS;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(full_bar + 8 * s, 256);
      mbar_init(empty_bar + 8 * s, 1);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < 8) {
    // ---------------- expanders ----------------
    const bool isA = warp < 4;
    const int r = threadIdx.x & 127;
    const int grow = (isA ? m0 : n0) + r;
    const bool row_ok = grow < (isA ? m : n);
    const uint4 *src = (isA ? A : Bt) + (size_t)(row_ok ? grow : 0) * (isA ? as16 : bs16);
    const uint32_t tiles = isA ? a_tiles : b_tiles;
    const int tile_bytes = isA ? A_TILE : B_TILE;
    // bits of the last k-block beyond k are masked (a window's stride may hold foreign bits)
    const int tail = k - (nkb - 1) * KB_BITS;  // 1..128 valid bits in the last block
    uint4 w = make_uint4(0, 0, 0, 0);
    if (row_ok && nkb > 0) w = __ldg(src);
    for (int kb = 0; kb < nkb; kb++) {
      const int s = kb % STAGES;
      uint4 cur = w;
      if (kb + 1 < nkb && row_ok) w = __ldg(src + kb + 1);
      if (kb == nkb - 1 && tail < KB_BITS) {
        uint32_t *cw = reinterpret_cast<uint32_t *>(&cur);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int lo = j * 32;
          cw[j] = tail <= lo ? 0u : (tail >= lo + 32 ? cw[j] : (cw[j] & ((1u << (tail - lo)) - 1u)));
        }
      }
      mbar_wait(empty_bar + 8 * s, ((kb / STAGES) & 1) ^ 1);
      expand_row(tiles + s * tile_bytes, r, cur);
      fence_proxy_async();
      mbar_arrive(full_bar + 8 * s);
    }
    if (isA) {
      // ---------------- epilogue: TMEM lane = tile row ----------------
      mbar_wait(accum_bar, 0);
      tc_fence_after();
      const int row = m0 + warp * 32 + lane;
      uint32_t out[4];
#pragma unroll
      for (int c = 0; c < 4; c++) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + c * 32, v);
        uint32_t word = 0;
#pragma unroll
        for (int j = 0; j < 32; j++) word |= (v[j] & 1u) << j;
        out[c] = word;
      }
      const int cw = n0 / 32;
      if (row < m && cw < cw32) {
        uint32_t *dst = C + (size_t)row * cs32 + cw;
        if (cw + 4 <= cw32) {
          uint4 o = make_uint4(out[0], out[1], out[2], out[3]);
          if (!clear) {
            const uint4 old = *reinterpret_cast<const uint4 *>(dst);
            o.x ^= old.x; o.y ^= old.y; o.z ^= old.z; o.w ^= old.w;
          }
          *reinterpret_cast<uint4 *>(dst) = o;
        } else {
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (cw + j < cw32) dst[j] = clear ? out[j] : (dst[j] ^ out[j]);
        }
      }
    }
  } else if (lane == 0) {
    // ---------------- MMA issuer ----------------
    constexpr uint32_t idesc = make_idesc_i8(BM, BN);
    for (int kb = 0; kb < nkb; kb++) {
      const int s = kb % STAGES;
      mbar_wait(full_bar + 8 * s, (kb / STAGES) & 1);
      tc_fence_after();
      const uint64_t ad = make_desc_sw128(a_tiles + s * A_TILE);
      const uint64_t bd = make_desc_sw128(b_tiles + s * B_TILE);
#pragma unroll
      for (int j = 0; j < 4; j++)  // K = 32 bytes per instruction: +2 in the >>4 address field
        umma_i8_ss(tmem_base, ad + 2 * j, bd + 2 * j, idesc, (kb | j) ? 1u : 0u);
      umma_commit(empty_bar + 8 * s);
    }
    umma_commit(accum_bar);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) tmem_dealloc(tmem_base, 128);
}

}  // namespace

bool mma_available() { return true; }

bool mma_shape_ok(const DMat *C, const DMat *A, const DMat *Bt) {
  return !(((uintptr_t)A->d | (uintptr_t)Bt->d | (uintptr_t)C->d) & 15) && !(A->rowstride & 1) &&
         !(Bt->rowstride & 1) && !(C->rowstride & 1);
}

void k_mul_abt_mma1(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    return;
  }
  if (!has_tcgen05()) die("m4ri_b200: the tcgen05 MMA path needs an sm_100 device\n");
  if (!mma_shape_ok(C, A, Bt)) die("m4ri_b200: mma kernel needs 16-byte aligned rows\n");
  static std::atomic<unsigned long long> attr_seen{0};
  if (first_use_on_this_device(attr_seen)) {
    M4B_CUDA(cudaFuncSetAttribute(mma_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  }
  const int tiles_m = (C->nrows + BM - 1) / BM, tiles_n = (C->ncols + BN - 1) / BN;
  timing_begin("mma_i8_kernel", s, 2.0 * C->nrows * C->ncols * A->ncols);
  mma_i8_kernel<<<tiles_m * tiles_n, MMA_THREADS, SMEM_BYTES, s>>>(
      (uint32_t *)C->d, (const uint4 *)A->d, (const uint4 *)Bt->d, C->nrows, C->ncols, A->ncols,
      A->rowstride / 2, Bt->rowstride / 2, C->rowstride * 2, C->width * 2, tiles_m, tiles_n, clear);
  M4B_LAUNCHED();
  timing_end(s);
}

}  // namespace m4b
