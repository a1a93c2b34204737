// kernels_mma5.cu — path (b) v4: the FP4 contraction of kernels_mma4.cu with the packed
// bits read straight from global memory.
//
// C (^)= A * Bt^T, A (m x k bits) and Bt (n x k bits) both k-contiguous (the layout of
// _mzd_mul_naive, m4ri-sys/src/mzd.rs:163-168; mzd_mul, m4ri-sys/src/strassen.rs:18, gets
// here through one bit-transpose of B).
//
// Why a second FP4 kernel: with the tensor pipe at twice the int8 rate the SM's shared
// memory is the contended resource. Per 512-bit k-step and CTA the v3 kernel moves
// 96 KiB through it — TMA write of the packed bits (16), the expanders' LDS of them (16),
// the Bt expanders' STS (32), the MMA's reads of Bt (32) — 75 % of 128 B/clk, and removing
// the STS alone (perf experiment) buys 7 %. Here every expander thread owns one
// row and loads its 64 bytes per k-step with two 256-bit LDGs (L1 no-allocate: they never
// touch the shared-memory SRAM), one k-step ahead of use. Shared memory carries only the
// expanded Bt ring: 64 KiB per k-step. No TMA, no raw-bit barriers, one warp fewer.
//
//   warps 0-3    A rows: LDG 512 k-bits -> 512 E2M1 nibbles -> TENSOR MEMORY (tcgen05.st)
//   warps 4-7    Bt rows: the same into 128B-swizzled K-major shared memory
//   warp  8      one lane of the leader CTA issues 8 tcgen05.mma.kind::mxf4 (K = 64) per k-step
//   warps 9-12   epilogue: TMEM -> registers -> parity bits -> C
// Exactness: as kernels_mma4.cu (products 0 or 0.25, fp32 sums of at most 65536 of them).
#include <cuda.h>

#include "common.cuh"
#include "tcgen05_util.cuh"

namespace m4b {
namespace {

using namespace tc;

constexpr int KSTEP_BITS = 512;            // k bits per pipeline stage (8 MMAs of K = 64 nibbles)
constexpr int STAGES = 3;                  // expanded-operand ring (A: TMEM, Bt: smem)
constexpr int B_TILE = 128 * 128;          // 128 rows x 128 B = 256 nibbles per row, SWIZZLE_128B
constexpr int B_STAGE = 2 * B_TILE;        // two 256-nibble k-halves of the 512-bit k-step
constexpr int NEXP = 8;                    // expander warps: 4 A + 4 Bt
constexpr int NWARPS = NEXP + 5;           // + MMA issuer, 4 epilogue warps
constexpr int NTHREADS = NWARPS * 32;
constexpr int SMEM_V5 = STAGES * B_STAGE + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TMEM_A_COL = 256;            // A stages: 64 columns each in [256, 448)
constexpr int TMEM_SF_COL = 448;           // scale factors (all 2^0) in [448, 512)
constexpr int K_CHUNK_MAX = 65536;         // host-side chunking of the contraction
constexpr uint32_t NIB = 0x11111111u;

__host__ __device__ constexpr uint32_t make_idesc_mxf4(int M, int N) {  // see kernels_mma4.cu
  return (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (1u << 23) | ((uint32_t)(M >> 4) << 24);
}

template <int CG>
__device__ __forceinline__ void umma_mxf4_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc,
                                             uint32_t idesc, uint32_t sfa, uint32_t sfb,
                                             uint32_t accumulate) {
  if constexpr (CG == 1) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.scale_vec::2X [%0], [%1], %2, %3, [%5], [%6], p;\n\t}" ::
            "r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.scale_vec::2X [%0], [%1], %2, %3, [%5], [%6], p;\n\t}" ::
            "r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
        : "memory");
  }
}

__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int &tm, int &tn) {
  const int grp = t / (8 * tiles_n), in_grp = t - grp * 8 * tiles_n;
  const int grp_rows = min(8, tiles_m - grp * 8);
  tm = grp * 8 + in_grp % grp_rows;
  tn = in_grp / grp_rows;
}

// 32 bytes of packed bits, streaming (read once per tile pass; keep them out of L1/SRAM)
__device__ __forceinline__ void ldg256(uint32_t *x, const uint8_t *p) {
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(x[0]), "=r"(x[1]), "=r"(x[2]), "=r"(x[3]), "=r"(x[4]), "=r"(x[5]), "=r"(x[6]),
                 "=r"(x[7])
               : "l"(p));
}
__device__ __forceinline__ void ldg128(uint32_t *x, const uint8_t *p) {
  asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(x[0]), "=r"(x[1]), "=r"(x[2]), "=r"(x[3])
               : "l"(p));
}

// The 16 words (512 k-bits) of one row starting at byte `off`; `rem` = valid bytes of the row
// from `off` on (a multiple of 8; <= 0 for rows past the matrix). Words past the row — another
// quadrant's data when the operand is a window — read as zero.
template <bool WIDE>
__device__ __forceinline__ void load_kstep(uint32_t (&x)[16], const uint8_t *p, int rem) {
  if (rem >= 64) {
    if constexpr (WIDE) {
      ldg256(&x[0], p);
      ldg256(&x[8], p + 32);
    } else {
#pragma unroll
      for (int j = 0; j < 4; j++) ldg128(&x[4 * j], p + 16 * j);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (rem > 16 * j) {
        ldg128(&x[4 * j], p + 16 * j);  // rows are padded to 16 bytes: in bounds
        if (rem < 16 * j + 16) x[4 * j + 2] = x[4 * j + 3] = 0u;  // rem = 16 j + 8
      } else {
        x[4 * j] = x[4 * j + 1] = x[4 * j + 2] = x[4 * j + 3] = 0u;
      }
    }
  }
}

// bit j of the result = parity of 4 * v[j] (v[j] = popcount / 4 exactly, < 2^21): the mantissa
// LSB of v[j] + 2^21. Four independent shift-or chains of eight instead of one of 32.
__device__ __forceinline__ uint32_t pack_parity(const uint32_t (&v)[32]) {
  uint32_t w[4];
#pragma unroll
  for (int g = 0; g < 4; g++) {
    uint32_t acc = 0;
#pragma unroll
    for (int j = 7; j >= 0; j--)
      acc = (acc << 1) | (__float_as_uint(__uint_as_float(v[8 * g + j]) + 2097152.0f) & 1u);
    w[g] = acc;
  }
  return w[0] | (w[1] << 8) | (w[2] << 16) | (w[3] << 24);
}

template <int CG, bool WIDE>
__global__ void __launch_bounds__(NTHREADS, 1)
mma5_kernel(uint32_t *__restrict__ C, const uint8_t *__restrict__ Ab, const uint8_t *__restrict__ Bb,
            int m, int n, int k, long long a_rsb, long long b_rsb, int cs32, int cw32, int tiles_m,
            int tiles_n, int clear, int dbg) {
  constexpr int NCOL = 128 * CG;  // accumulator columns = tile columns
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B: 1024-B alignment
  const uint32_t bexp_base = sbase;
  const uint32_t bars = bexp_base + STAGES * B_STAGE;
  const uint32_t full_bar = bars, empty_bar = bars + 32, accum_bar = bars + 64, dempty_bar = bars + 72,
                 tmem_slot = bars + 80;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 1 ? 0u : cluster_ctarank();
  const int cluster_id = blockIdx.x / CG, nclusters = gridDim.x / CG;
  const int ntiles = tiles_m * tiles_n;
  const int nks = (k + KSTEP_BITS - 1) / KSTEP_BITS;  // k-steps per tile

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(full_bar + 8 * s, NEXP * CG);  // one arrival per expander warp of every CTA of the pair
      mbar_init(empty_bar + 8 * s, 1);         // tcgen05.commit
    }
    mbar_init(accum_bar, 1);                   // tcgen05.commit after the last k-step of a tile
    mbar_init(dempty_bar, 4 * CG);             // one arrival per epilogue warp of every CTA
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == NEXP) tmem_alloc_all<CG>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();  // the peer's barriers exist before anyone arrives on them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < NEXP) {
    // ================= expanders: one thread = one row of the CTA's 128-row slab =================
    const bool isA = warp < 4;
    const int r = (warp & 3) * 32 + lane;  // row of the slab = TMEM lane
    const uint32_t sw = (uint32_t)(r & 7);
    const uint32_t brow = bexp_base + (r >> 3) * 1024 + (r & 7) * 128;
    const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t a_lane = lane_base + TMEM_A_COL;
    const uint32_t full_leader = map_to_rank(full_bar, 0);
    const uint8_t *const base = isA ? Ab : Bb;
    const long long rsb = isA ? a_rsb : b_rsb;
    const int nrows = isA ? m : n;
    const int wbytes = ((k + 63) >> 6) << 3;  // valid bytes of a row (whole words; excess bits are zero)
    if (isA) {
      // every scale factor the MMAs can address is UE8M0 127 = 2^0 (any layout reads the same byte)
      uint32_t one[32];
#pragma unroll
      for (int j = 0; j < 32; j++) one[j] = 0x7F7F7F7Fu;
      tmem_st32(lane_base + TMEM_SF_COL, one);
      tmem_st32(lane_base + TMEM_SF_COL + 32, one);
      tmem_wait_st();
      tc_fence_before();
    }
    // row pointer of this thread in tile t (rows past the matrix: pointer to row 0, rem <= 0)
    auto row_of = [&](int t, const uint8_t *&p, int &w) {
      int tm, tn;
      tile_coords(t, tiles_m, tiles_n, tm, tn);
      const int row = ((isA ? tm : tn) * CG + (int)rank) * 128 + r;
      const bool valid = row < nrows;
      p = base + (size_t)(valid ? row : 0) * rsb;
      w = valid ? wbytes : 0;
    };
    uint32_t it = 0;
    // turns the 512 bits in x into nibbles and publishes stage it % STAGES
    auto expand = [&](uint32_t (&x)[16]) {
      const uint32_t s = it % STAGES;
      mbar_wait(empty_bar + 8 * s, ((it / STAGES) & 1) ^ 1);
      if (dbg & 1) {
      } else if (isA) {
        tc_fence_after();
#pragma unroll
        for (int h = 0; h < 2; h++) {
          uint32_t v[32];
#pragma unroll
          for (int c = 0; c < 8; c++)
#pragma unroll
            for (int j = 0; j < 4; j++) v[4 * c + j] = (x[8 * h + c] >> j) & NIB;
          tmem_st32(a_lane + s * 64 + h * 32, v);
        }
        tmem_wait_st();
        tc_fence_before();
      } else {
        const uint32_t dst = brow + s * B_STAGE;
#pragma unroll
        for (int h = 0; h < 2; h++)
#pragma unroll
          for (int c = 0; c < 8; c++) {
            const uint32_t w = x[8 * h + c];
            sts128(dst + h * B_TILE + (((uint32_t)c ^ sw) << 4), w & NIB, (w >> 1) & NIB,
                   (w >> 2) & NIB, (w >> 3) & NIB);
          }
        fence_proxy_async();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(full_leader + 8 * s);
      it++;
    };
    // flat walk over (tile, k-step) with the loads one k-step ahead of the expansion
    int t = cluster_id, ks = 0;
    const uint8_t *p = base;
    int w = 0;
    uint32_t xa[16], xb[16];
    if (t < ntiles) {
      row_of(t, p, w);
      load_kstep<WIDE>(xa, p, w);
    }
    while (t < ntiles) {
      // advance to the next k-step (possibly of the next tile) and start its loads
      if (++ks == nks) { ks = 0; t += nclusters; if (t < ntiles) row_of(t, p, w); }
      if (t < ntiles) load_kstep<WIDE>(xb, p + ks * 64, w - ks * 64);
      expand(xa);
      if (t >= ntiles) break;
      if (++ks == nks) { ks = 0; t += nclusters; if (t < ntiles) row_of(t, p, w); }
      if (t < ntiles) load_kstep<WIDE>(xa, p + ks * 64, w - ks * 64);
      expand(xb);
    }
  } else if (warp == NEXP) {
    // ================= MMA issuer: one lane of the leader CTA =================
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc_mxf4(128 * CG, NCOL);
      const uint32_t sfa = tmem_base + TMEM_SF_COL, sfb = tmem_base + TMEM_SF_COL + 32;
      uint32_t it = 0, tile_it = 0;
      for (int t = cluster_id; t < ntiles; t += nclusters, tile_it++) {
        if (tile_it > 0) mbar_wait(dempty_bar, (tile_it - 1) & 1);  // epilogue has drained D
        tc_fence_after();
        for (int ks = 0; ks < nks; ks++, it++) {
          const uint32_t s = it % STAGES;
          mbar_wait(full_bar + 8 * s, (it / STAGES) & 1);
          tc_fence_after();
          const int kleft = k - ks * KSTEP_BITS;
          const int nmma = kleft >= KSTEP_BITS ? 8 : (kleft + 63) >> 6;
          const uint32_t a_t = tmem_base + TMEM_A_COL + s * 64;
          const uint64_t bd = make_desc_sw128(bexp_base + s * B_STAGE);
#pragma unroll
          for (int i = 0; i < 8; i++) {
            if (i < nmma) {
              // K = 64 nibbles = 32 bytes per MMA: 8 TMEM columns of A; 32 bytes (+2 in the >>4
              // address field) within the 128-byte swizzle row of Bt, next 16 KiB tile after four
              const uint64_t bdi = bd + (uint64_t)((i >> 2) * (B_TILE >> 4) + (i & 3) * 2);
              umma_mxf4_ts<CG>(tmem_base, a_t + 8 * i, bdi, idesc, sfa, sfb, (ks | i) ? 1u : 0u);
            }
          }
          umma_commit<CG>(empty_bar + 8 * s);
        }
        umma_commit<CG>(accum_bar);
      }
    }
  } else {
    // ================= epilogue: the last four warps, TMEM lane quarter = warp % 4 =================
    const int qd = warp & 3;
    const uint32_t dempty_leader = map_to_rank(dempty_bar, 0);
    uint32_t tile_it = 0;
    for (int t = cluster_id; t < ntiles; t += nclusters, tile_it++) {
      int tm, tn;
      tile_coords(t, tiles_m, tiles_n, tm, tn);
      const int row = (tm * CG + (int)rank) * 128 + qd * 32 + lane;
      mbar_wait_long(accum_bar, tile_it & 1);
      tc_fence_after();
      uint32_t out[NCOL / 32];
      if (dbg & 4) {  // perf experiment: no drain
#pragma unroll
        for (int c = 0; c < NCOL / 32; c++) out[c] = 0;
      } else {
        // two chunks of 32 columns in flight: the load of chunk c+1 overlaps the packing of chunk c
        const uint32_t d_lane = tmem_base + ((uint32_t)(qd * 32) << 16);
        uint32_t va[32], vb[32];
        tmem_ld32_nowait(d_lane, va);
#pragma unroll
        for (int c = 0; c < NCOL / 32; c += 2) {
          tmem_wait_ld();
          tmem_ld32_nowait(d_lane + (c + 1) * 32, vb);
          out[c] = pack_parity(va);
          tmem_wait_ld();
          if (c + 2 < NCOL / 32) tmem_ld32_nowait(d_lane + (c + 2) * 32, va);
          out[c + 1] = pack_parity(vb);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(dempty_leader);  // D may be overwritten by the next tile
      const int cw0 = tn * (NCOL / 32);
      if (row < m) {
        uint32_t *dst = C + (size_t)row * cs32 + cw0;
#pragma unroll
        for (int g = 0; g < NCOL / 128; g++) {
          const int cw = cw0 + 4 * g;
          if (cw + 4 <= cw32) {
            uint4 o = make_uint4(out[4 * g], out[4 * g + 1], out[4 * g + 2], out[4 * g + 3]);
            if (!clear) {
              const uint4 old = *reinterpret_cast<const uint4 *>(dst + 4 * g);
              o.x ^= old.x; o.y ^= old.y; o.z ^= old.z; o.w ^= old.w;
            }
            *reinterpret_cast<uint4 *>(dst + 4 * g) = o;
          } else {
#pragma unroll
            for (int j = 0; j < 4; j++)
              if (cw + j < cw32) dst[4 * g + j] = clear ? out[4 * g + j] : (dst[4 * g + j] ^ out[4 * g + j]);
          }
        }
      }
    }
  }
  // ---- teardown: nobody leaves while the pair's MMAs may still read its memory ----
  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();
  if (warp == NEXP) tmem_dealloc_all<CG>(tmem_base);
}

}  // namespace

// ---- host side -------------------------------------------------------------------
template <int CG, bool WIDE>
static void launch_mma5(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s) {
  static std::atomic<unsigned long long> attr_seen{0};
  if (first_use_on_this_device(attr_seen)) {
    M4B_CUDA(cudaFuncSetAttribute(mma5_kernel<CG, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_V5));
  }
  const int tile = 128 * CG;
  const int tiles_m = (C->nrows + tile - 1) / tile, tiles_n = (C->ncols + tile - 1) / tile;
  const int ntiles = tiles_m * tiles_n;
  int nclusters = sm_count() / CG;
  if (nclusters > ntiles) nclusters = ntiles;
#ifdef M4RI_B200_PERF_HOOKS
  static const int dbg = getenv("M4RI_B200_DBG") ? atoi(getenv("M4RI_B200_DBG")) : 0;  // perf-hook builds only (build.py --perf)
#else
  const int dbg = 0;  // the shipped library never lets the environment alter results
#endif
  for (int k0 = 0; k0 < A->ncols; k0 += K_CHUNK_MAX) {
    const int kc = A->ncols - k0 < K_CHUNK_MAX ? A->ncols - k0 : K_CHUNK_MAX;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nclusters * CG);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = SMEM_V5;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    timing_begin(CG == 2 ? "mma5_kernel<2>" : "mma5_kernel<1>", s, 2.0 * C->nrows * C->ncols * kc);
    M4B_CUDA(cudaLaunchKernelEx(&cfg, mma5_kernel<CG, WIDE>, (uint32_t *)C->d,
                                (const uint8_t *)(A->d + k0 / 64), (const uint8_t *)(Bt->d + k0 / 64),
                                C->nrows, C->ncols, kc, (long long)A->rowstride * 8,
                                (long long)Bt->rowstride * 8, C->rowstride * 2, C->width * 2, tiles_m,
                                tiles_n, (clear && k0 == 0) ? 1 : 0, dbg));
    M4B_LAUNCHED();
    timing_end(s);
  }
}

// C (^)= A*Bt^T on the FP4 pipe, packed bits by direct global loads; cg = CTAs per MMA (1 or 2)
void k_mul_abt_mma5(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, cudaStream_t s) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    return;
  }
  if (!has_tcgen05()) die("m4ri_b200: the tcgen05 MMA path needs an sm_100 device\n");
  if (!mma_shape_ok(C, A, Bt)) die("m4ri_b200: mma kernel needs 16-byte aligned rows\n");
  // 256-bit loads need 32-byte aligned rows (true for every width that is a multiple of 256 bits)
  const bool wide = !(((uintptr_t)A->d | (uintptr_t)Bt->d) & 31) && !(A->rowstride & 3) && !(Bt->rowstride & 3);
  if (cg == 2) {
    if (wide) launch_mma5<2, true>(C, A, Bt, clear, s);
    else launch_mma5<2, false>(C, A, Bt, clear, s);
  } else {
    if (wide) launch_mma5<1, true>(C, A, Bt, clear, s);
    else launch_mma5<1, false>(C, A, Bt, clear, s);
  }
}

}  // namespace m4b
