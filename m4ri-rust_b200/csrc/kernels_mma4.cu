// kernels_mma4.cu — path (b) v3: the GF(2) product as an FP4 contraction on the
// 5th-gen tensor cores at twice the int8 rate.
//
// C (^)= A * Bt^T with A (m x k bits) and Bt (n x k bits) both k-contiguous, the
// layout _mzd_mul_naive takes (m4ri-sys/src/mzd.rs:163-168); mzd_mul
// (m4ri-sys/src/strassen.rs:18) gets here through one bit-transpose of B (api.cu).
//
// Every bit becomes one E2M1 element: nibble 0b0001 = 0.5 for a set bit, 0 for a
// clear one. tcgen05.mma.kind::mxf4 (block-scaled, all UE8M0 scale factors = 2^0)
// multiplies 64 such elements per row per instruction — twice the K of kind::i8 in
// the same issue slot — and adds the products (0 or 0.25, exactly representable)
// into fp32 accumulators in tensor memory. A sum of c products is c/4, exact in fp32
// for c < 2^24; the host wrapper chunks the contraction at 65536, so the result bit
// is bit 0 of 4*acc, read as the mantissa LSB of acc + 2^21.
//
// Pipeline (one CTA per SM, CTA pairs with cta_group::2 own a 256 x 256 tile of C,
// persistent over the tiles) — the same roles as kernels_mma2.cu:
//   warp 17        TMA: packed bits of 128 rows of A and of Bt, 1024 k-bits per stage
//   warps 0-3,8-11 A rows: 256 k-bits -> 256 nibbles -> TENSOR MEMORY (tcgen05.st)
//   warps 4-7,12-15 Bt rows: the same into 128B-swizzled K-major shared memory
//                  (warp / 8 = which 256-bit half of the 512-bit k-step: 16 warps fill one
//                  stage together, which halves the time a stage is being written)
//   warp 16        one lane of the leader CTA issues 8 MMAs (K = 64) per 512-bit k-step
//   warps 18-21    epilogue: TMEM -> registers -> parity bits -> C
#include <cuda.h>

#include "common.cuh"
#include "tcgen05_util.cuh"

namespace m4b {
namespace {

using namespace tc;

constexpr int KSTEP_BITS = 512;            // k bits per pipeline stage (8 MMAs of K = 64 nibbles)
constexpr int STAGES = 3;                  // expanded-operand ring (A: TMEM, Bt: smem)
constexpr int RAW_KSTEPS = 2;              // k-steps per TMA stage: 1024 bits = 128 B per row
constexpr int RAW_STAGES = 3;
constexpr int RAW_TILE = 128 * 128;        // 128 rows x 128 B of packed bits (A or Bt)
constexpr int RAW_STAGE = 2 * RAW_TILE;    // A rows then Bt rows
constexpr int B_TILE = 128 * 128;          // 128 rows x 128 B = 256 nibbles per row, SWIZZLE_128B
constexpr int B_STAGE = 2 * B_TILE;        // two 256-nibble k-halves of the 512-bit k-step
constexpr int HALVES = 1;                  // warps sharing one row's k-step (1 or 2)
constexpr int NEXP = 8 * HALVES;           // expander warps: HALVES x (4 A + 4 Bt)
constexpr int WBITS = KSTEP_BITS / HALVES; // k bits one warp expands per k-step
constexpr int WWORDS = WBITS / 32;
constexpr int NWARPS = NEXP + 6;           // + MMA issuer, TMA producer, 4 epilogue warps
constexpr int NTHREADS = NWARPS * 32;
constexpr int SMEM_V4 = RAW_STAGES * RAW_STAGE + STAGES * B_STAGE + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TMEM_A_COL = 256;            // A stages: 64 columns each in [256, 448)
constexpr int TMEM_SF_COL = 448;           // scale factors (all 2^0) in [448, 512)
constexpr int K_CHUNK_MAX = 65536;         // host-side chunking of the contraction
constexpr uint32_t NIB = 0x11111111u;

// kind::mxf4 block-scaled instruction descriptor: A and B E2M1 (1 << 7, 1 << 10), both
// K-major, N >> 3 in [17,23), UE8M0 scale factors (1 << 23), M >> 4 in [24,29), scale
// factor ids 0, K = 64 (bit 31 clear).
__host__ __device__ constexpr uint32_t make_idesc_mxf4(int M, int N) {
  return (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (1u << 23) | ((uint32_t)(M >> 4) << 24);
}

// D[tmem] (+)= A[tmem] * B[smem desc], E2M1 x E2M1 -> fp32, one UE8M0 scale per 32 elements
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

// tile index -> (tile row, tile column): groups of 8 tile-rows sweep the tile-columns
// together, so the CTAs resident at one time share a few panels of A and Bt in L2
__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int &tm, int &tn) {
  const int grp = t / (8 * tiles_n), in_grp = t - grp * 8 * tiles_n;
  const int grp_rows = min(8, tiles_m - grp * 8);
  tm = grp * 8 + in_grp % grp_rows;
  tn = in_grp / grp_rows;
}

template <int CG>
__global__ void __launch_bounds__(NTHREADS, 1)
mma4_kernel(uint32_t *__restrict__ C, const __grid_constant__ CUtensorMap tmA,
            const __grid_constant__ CUtensorMap tmB, int m, int n, int k, int cs32, int cw32,
            int tiles_m, int tiles_n, int clear, float *__restrict__ raw_acc, int dbg) {
  constexpr int NCOL = 128 * CG;  // accumulator columns = tile columns
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B: 1024-B alignment
  const uint32_t raw_base = sbase;
  const uint32_t bexp_base = sbase + RAW_STAGES * RAW_STAGE;
  const uint32_t bars = bexp_base + STAGES * B_STAGE;
  const uint32_t full_bar = bars, empty_bar = bars + 64, rawfull_bar = bars + 128,
                 rawempty_bar = bars + 160, accum_bar = bars + 192, dempty_bar = bars + 200,
                 tmem_slot = bars + 208;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 1 ? 0u : cluster_ctarank();
  const int cluster_id = blockIdx.x / CG, nclusters = gridDim.x / CG;
  const int ntiles = tiles_m * tiles_n;
  const int nks = (k + KSTEP_BITS - 1) / KSTEP_BITS;        // k-steps per tile
  const int nrs = (nks + RAW_KSTEPS - 1) / RAW_KSTEPS;      // TMA stages per tile

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(full_bar + 8 * s, NEXP * CG);  // one arrival per expander warp of every CTA of the pair
      mbar_init(empty_bar + 8 * s, 1);      // tcgen05.commit
    }
    for (int s = 0; s < RAW_STAGES; s++) {
      mbar_init(rawfull_bar + 8 * s, 1);    // expect_tx arrival + TMA bytes
      mbar_init(rawempty_bar + 8 * s, NEXP); // one arrival per expander warp
    }
    mbar_init(accum_bar, 1);                // tcgen05.commit after the last k-step of a tile
    mbar_init(dempty_bar, 4 * CG);          // one arrival per epilogue warp of every CTA
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
    // ===== expanders: warps with (warp & 4) == 0 -> A (TMEM), the others -> Bt (smem); warp / 8
    // selects which WBITS-bit part of every 512-bit k-step the warp turns into nibbles.
    // No tail masking: bits past k are zero in memory (excess bits of a row are kept zero by
    // every kernel of the library, columns past the row are zero-filled by TMA).
    const int half = warp >> 3;
    const bool isA = (warp & 4) == 0;
    const int r = (warp & 3) * 32 + lane;  // row of the 128-row slab = TMEM lane
    const uint32_t sw = (uint32_t)(r & 7);
    const uint32_t raw_row = raw_base + (isA ? 0 : RAW_TILE) + r * 128;
    const uint32_t brow = bexp_base + (half * WBITS / 256) * B_TILE + (r >> 3) * 1024 + (r & 7) * 128;
    const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t a_lane = lane_base + TMEM_A_COL + half * (WBITS / 8);
    const uint32_t full_leader = map_to_rank(full_bar, 0);
    if (isA && half == 0) {
      // every scale factor the MMAs can address is UE8M0 127 = 2^0 (any layout reads the same byte)
      uint32_t one[32];
#pragma unroll
      for (int j = 0; j < 32; j++) one[j] = 0x7F7F7F7Fu;
      tmem_st32(lane_base + TMEM_SF_COL, one);
      tmem_st32(lane_base + TMEM_SF_COL + 32, one);
      tmem_wait_st();
      tc_fence_before();
    }
    uint32_t it = 0, rit = 0;
    for (int t = cluster_id; t < ntiles; t += nclusters) {
      for (int ks = 0; ks < nks; ks++, it++) {
        const int q = ks & (RAW_KSTEPS - 1);
        const uint32_t rs = rit % RAW_STAGES;
        if (q == 0) mbar_wait(rawfull_bar + 8 * rs, (rit / RAW_STAGES) & 1);
        const uint32_t rrow = raw_row + rs * RAW_STAGE;
        uint32_t x[WWORDS];
#pragma unroll
        for (int c = 0; c < WWORDS / 4; c++) {
          const uint4 v4 = lds128(rrow + (((uint32_t)(4 * q + (WWORDS / 4) * half + c) ^ sw) << 4));
          x[4 * c] = v4.x; x[4 * c + 1] = v4.y; x[4 * c + 2] = v4.z; x[4 * c + 3] = v4.w;
        }
        const uint32_t s = it % STAGES;
        mbar_wait(empty_bar + 8 * s, ((it / STAGES) & 1) ^ 1);
        if (dbg & 1) {
        } else if (isA) {
          tc_fence_after();
#pragma unroll
          for (int h = 0; h < WWORDS / 8; h++) {
            uint32_t v[32];
#pragma unroll
            for (int c = 0; c < 8; c++)
#pragma unroll
              for (int j = 0; j < 4; j++) v[4 * c + j] = (x[8 * h + c] >> j) & NIB;
            if (dbg & 8) {  // perf experiment: expansion ALU work without the tensor-memory store
              uint32_t acc = 0;
#pragma unroll
              for (int j = 0; j < 32; j++) acc ^= v[j];
              if (acc == 0x12345678u) asm volatile("nanosleep.u32 1;");
            } else
            tmem_st32(a_lane + s * 64 + h * 32, v);
          }
          tmem_wait_st();
          tc_fence_before();
        } else if (!(dbg & 2)) {
          const uint32_t dst = brow + s * B_STAGE;
#pragma unroll
          for (int h = 0; h < WWORDS / 8; h++)
#pragma unroll
            for (int c = 0; c < 8; c++) {
              const uint32_t w = x[8 * h + c];
              if (dbg & 16) {
                if ((((w & NIB) ^ ((w >> 1) & NIB)) ^ (((w >> 2) & NIB) ^ ((w >> 3) & NIB))) == 0x12345678u)
                  asm volatile("nanosleep.u32 1;");
              } else
              sts128(dst + h * B_TILE + (((uint32_t)c ^ sw) << 4), w & NIB, (w >> 1) & NIB,
                     (w >> 2) & NIB, (w >> 3) & NIB);
            }
          fence_proxy_async();
        }
        __syncwarp();
        if (lane == 0) {
          mbar_arrive_cluster(full_leader + 8 * s);
          if (q == RAW_KSTEPS - 1 || ks == nks - 1) mbar_arrive_local(rawempty_bar + 8 * rs);
        }
        if (q == RAW_KSTEPS - 1 || ks == nks - 1) rit++;
      }
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
  } else if (warp == NEXP + 1) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t rit = 0;
      for (int t = cluster_id; t < ntiles; t += nclusters) {
        int tm, tn;
        tile_coords(t, tiles_m, tiles_n, tm, tn);
        const int arow = (tm * CG + (int)rank) * 128, brow = (tn * CG + (int)rank) * 128;
        for (int ri = 0; ri < nrs; ri++, rit++) {
          const uint32_t rs = rit % RAW_STAGES;
          mbar_wait_long(rawempty_bar + 8 * rs, ((rit / RAW_STAGES) & 1) ^ 1);
          mbar_expect_tx(rawfull_bar + 8 * rs, RAW_STAGE);
          tma_load_2d(raw_base + rs * RAW_STAGE, &tmA, ri * 128, arow, rawfull_bar + 8 * rs);
          tma_load_2d(raw_base + rs * RAW_STAGE + RAW_TILE, &tmB, ri * 128, brow, rawfull_bar + 8 * rs);
        }
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
      } else
#pragma unroll
      for (int c = 0; c < NCOL / 32; c++) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(qd * 32) << 16) + c * 32, v);
        if (raw_acc != nullptr && row < m) {  // diagnostic: the fp32 accumulators as they are
#pragma unroll
          for (int j = 0; j < 32; j++) {
            const int col = tn * NCOL + c * 32 + j;
            if (col < n) raw_acc[(size_t)row * n + col] = __uint_as_float(v[j]);
          }
        }
        uint32_t word = 0;
#pragma unroll
        for (int j = 0; j < 32; j++)
          word |= (__float_as_uint(__uint_as_float(v[j]) + 2097152.0f) & 1u) << j;
        out[c] = word;
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
template <int CG>
static void launch_mma4(DMat *C, const DMat *A, const DMat *Bt, int clear, float *raw_acc,
                        cudaStream_t s) {
  static std::atomic<unsigned long long> attr_seen{0};
  if (first_use_on_this_device(attr_seen)) {
    M4B_CUDA(cudaFuncSetAttribute(mma4_kernel<CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_V4));
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
    const CUtensorMap tmA = packed_rows_map(A->d + k0 / 64, A->nrows, width_of(kc), A->rowstride);
    const CUtensorMap tmB = packed_rows_map(Bt->d + k0 / 64, Bt->nrows, width_of(kc), Bt->rowstride);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nclusters * CG);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = SMEM_V4;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    timing_begin(CG == 2 ? "mma4_kernel<2>" : "mma4_kernel<1>", s, 2.0 * C->nrows * C->ncols * kc);
    M4B_CUDA(cudaLaunchKernelEx(&cfg, mma4_kernel<CG>, (uint32_t *)C->d, tmA, tmB, C->nrows, C->ncols,
                                kc, C->rowstride * 2, C->width * 2, tiles_m, tiles_n,
                                (clear && k0 == 0) ? 1 : 0, raw_acc, dbg));
    M4B_LAUNCHED();
    timing_end(s);
  }
}

// C (^)= A*Bt^T on the FP4 pipe; cg = CTAs per MMA (1 or 2). raw_acc (device, m x n floats,
// may be NULL) receives the fp32 accumulators of the last k-chunk: exactness diagnostics.
void k_mul_abt_mma4(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, float *raw_acc,
                    cudaStream_t s) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    return;
  }
  if (!has_tcgen05()) die("m4ri_b200: the tcgen05 MMA path needs an sm_100 device\n");
  if (!mma_shape_ok(C, A, Bt)) die("m4ri_b200: mma kernel needs 16-byte aligned rows\n");
  if (cg == 2) launch_mma4<2>(C, A, Bt, clear, raw_acc, s);
  else launch_mma4<1>(C, A, Bt, clear, raw_acc, s);
}

}  // namespace m4b
