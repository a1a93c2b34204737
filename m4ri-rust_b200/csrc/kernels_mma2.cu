// kernels_mma2.cu — path (b) v2: the GF(2) product as an int8 contraction on the
// 5th-gen tensor cores, CTA pairs (cta_group::2), A operand in tensor memory.
//
// C (^)= A * Bt^T with A (m x k bits) and Bt (n x k bits) both k-contiguous, the
// layout _mzd_mul_naive takes (m4ri-sys/src/mzd.rs:163-168); mzd_mul
// (m4ri-sys/src/strassen.rs:18) gets here through one bit-transpose of B (api.cu).
//
// Data flow per CTA (one CTA per SM, a cluster of CG CTAs owns a 128*CG x 128*CG
// tile of C and walks the tiles persistently):
//   TMA          packed bits of 128 rows of A and 128 rows of Bt, 1024 k-bits at a
//                time, into 128B-swizzled shared memory (zero fill past the edges)
//   warps 0-3    read their row's 256 k-bits, widen bit -> byte and store the bytes
//                straight into TENSOR MEMORY with tcgen05.st (the A operand never
//                touches shared memory again: "TS" form of the MMA)
//   warps 4-7    the same for Bt, into a 128B-swizzled K-major shared-memory tile
//   warp  8      one lane of the leader CTA issues tcgen05.mma.kind::i8 (M = N =
//                128*CG, K = 32 per instruction) with int32 accumulators in TMEM
//   warp  9      one lane issues the TMA loads
//   warps 10-13  epilogue: TMEM -> registers, keep bit 0 of every accumulator, pack
//                32 columns per word, store (or XOR into) C
//
// Widening trick: only bit 0 of a product matters, and (a*b) & 1 == (a&1) & (b&1)
// for two's-complement bytes, so byte j of (w >> s) — whose bit 0 is bit 8j+s of
// w, the other seven bits being whatever follows — is a valid operand byte: ONE
// shift per four operand bytes, no masking. The same permutation of k is applied
// to A and Bt, so the sum over k is unchanged. Signed bytes bound every product by
// 2^14, so the int32 accumulator is exact (no wrap) for k <= 131072; longer
// contractions are chunked by the host wrapper.
#include <cuda.h>

#include "common.cuh"
#include "tcgen05_util.cuh"

namespace m4b {
namespace {

constexpr int KSTEP_BITS = 256;            // k bits per pipeline stage (8 MMAs of K = 32 bytes)
constexpr int STAGES = 4;                  // expanded-operand ring (A: TMEM, Bt: smem)
constexpr int RAW_KSTEPS = 4;              // k-steps per TMA stage: 1024 bits = 128 B per row
constexpr int RAW_STAGES = 2;
constexpr int RAW_TILE = 128 * 128;        // 128 rows x 128 B of packed bits (A or Bt)
constexpr int RAW_STAGE = 2 * RAW_TILE;    // A rows then Bt rows
constexpr int B_TILE = 128 * 128;          // 128 rows x 128 expanded bytes, SWIZZLE_128B
constexpr int B_STAGE = 2 * B_TILE;        // two 128-byte k-halves of the 256-bit k-step
constexpr int NWARPS = 14;
constexpr int NTHREADS = NWARPS * 32;
constexpr int SMEM_V2 = RAW_STAGES * RAW_STAGE + STAGES * B_STAGE + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TMEM_A_COL = 256;            // A stages live in columns [256, 512)
constexpr int K_CHUNK_MAX = 65536;         // host-side chunking of the contraction

using namespace tc;

// D[tmem] (+)= A[tmem] * B[smem desc], signed int8 x signed int8 -> int32
template <int CG>
__device__ __forceinline__ void umma_i8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc,
                                           uint32_t idesc, uint32_t accumulate) {
  if constexpr (CG == 1) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
        "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// kind::i8 instruction descriptor: D = S32 (2 << 4), A and B signed 8-bit (1 << 7,
// 1 << 10), both K-major, N >> 3 in [17,23), M >> 4 in [24,29); saturate (bit 3) off.
__host__ __device__ constexpr uint32_t make_idesc_s8(int M, int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
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
mma2_kernel(uint32_t *__restrict__ C, const __grid_constant__ CUtensorMap tmA,
            const __grid_constant__ CUtensorMap tmB, int m, int n, int k, int cs32, int cw32,
            int tiles_m, int tiles_n, int clear) {
  constexpr int NCOL = 128 * CG;  // accumulator columns = tile columns
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B: 1024-B alignment
  const uint32_t raw_base = sbase;
  const uint32_t bexp_base = sbase + RAW_STAGES * RAW_STAGE;
  const uint32_t bars = bexp_base + STAGES * B_STAGE;
  const uint32_t full_bar = bars, empty_bar = bars + 32, rawfull_bar = bars + 64,
                 rawempty_bar = bars + 80, accum_bar = bars + 96, dempty_bar = bars + 104,
                 tmem_slot = bars + 112;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 1 ? 0u : cluster_ctarank();
  const int cluster_id = blockIdx.x / CG, nclusters = gridDim.x / CG;
  const int ntiles = tiles_m * tiles_n;
  const int nks = (k + KSTEP_BITS - 1) / KSTEP_BITS;        // k-steps per tile
  const int nrs = (nks + RAW_KSTEPS - 1) / RAW_KSTEPS;      // TMA stages per tile

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(full_bar + 8 * s, 8 * CG);  // one arrival per expander warp of every CTA of the pair
      mbar_init(empty_bar + 8 * s, 1);      // tcgen05.commit
    }
    for (int s = 0; s < RAW_STAGES; s++) {
      mbar_init(rawfull_bar + 8 * s, 1);    // expect_tx arrival + TMA bytes
      mbar_init(rawempty_bar + 8 * s, 8);   // one arrival per expander warp
    }
    mbar_init(accum_bar, 1);                // tcgen05.commit after the last k-step of a tile
    mbar_init(dempty_bar, 4 * CG);          // one arrival per epilogue warp of every CTA
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) tmem_alloc_all<CG>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();  // the peer's barriers exist before anyone arrives on them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < 8) {
    // ================= expanders: warps 0-3 -> A (TMEM), warps 4-7 -> Bt (smem) =================
    const bool isA = warp < 4;
    const int r = (warp & 3) * 32 + lane;  // row of the 128-row slab = TMEM lane
    const uint32_t sw = (uint32_t)(r & 7);
    const uint32_t raw_row = raw_base + (isA ? 0 : RAW_TILE) + r * 128;
    const uint32_t brow = bexp_base + (r >> 3) * 1024 + (r & 7) * 128;
    const uint32_t a_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + TMEM_A_COL;
    const uint32_t full_leader = map_to_rank(full_bar, 0);
    const int tail = k - (nks - 1) * KSTEP_BITS;  // 1..256 valid bits in the last k-step
    uint32_t it = 0, rit = 0;
    for (int t = cluster_id; t < ntiles; t += nclusters) {
      for (int ks = 0; ks < nks; ks++, it++) {
        const int q = ks & (RAW_KSTEPS - 1);
        const uint32_t rs = rit % RAW_STAGES;
        if (q == 0) mbar_wait(rawfull_bar + 8 * rs, (rit / RAW_STAGES) & 1);
        const uint32_t rrow = raw_row + rs * RAW_STAGE;
        const uint4 xa = lds128(rrow + (((uint32_t)(2 * q) ^ sw) << 4));
        const uint4 xb = lds128(rrow + (((uint32_t)(2 * q + 1) ^ sw) << 4));
        uint32_t x[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
        if (ks == nks - 1 && tail < KSTEP_BITS) {
#pragma unroll
          for (int j = 0; j < 8; j++) {
            const int lo = j * 32;
            x[j] = tail <= lo ? 0u : (tail >= lo + 32 ? x[j] : (x[j] & ((1u << (tail - lo)) - 1u)));
          }
        }
        const uint32_t s = it % STAGES;
        mbar_wait(empty_bar + 8 * s, ((it / STAGES) & 1) ^ 1);
        if (isA) {
          tc_fence_after();
          uint32_t v[32];
#pragma unroll
          for (int h = 0; h < 2; h++) {
#pragma unroll
            for (int j = 0; j < 4; j++)
#pragma unroll
              for (int sft = 0; sft < 8; sft++) v[8 * j + sft] = x[4 * h + j] >> sft;
            tmem_st32(a_lane + s * 64 + h * 32, v);
          }
          tmem_wait_st();
          tc_fence_before();
        } else {
          const uint32_t dst = brow + s * B_STAGE;
#pragma unroll
          for (int h = 0; h < 2; h++)
#pragma unroll
            for (int j = 0; j < 4; j++) {
              const uint32_t w = x[4 * h + j];
              sts128(dst + h * B_TILE + (((uint32_t)(2 * j) ^ sw) << 4), w, w >> 1, w >> 2, w >> 3);
              sts128(dst + h * B_TILE + (((uint32_t)(2 * j + 1) ^ sw) << 4), w >> 4, w >> 5, w >> 6, w >> 7);
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
  } else if (warp == 8) {
    // ================= MMA issuer: one lane of the leader CTA =================
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc_s8(128 * CG, NCOL);
      uint32_t it = 0, tile_it = 0;
      for (int t = cluster_id; t < ntiles; t += nclusters, tile_it++) {
        if (tile_it > 0) mbar_wait(dempty_bar, (tile_it - 1) & 1);  // epilogue has drained D
        tc_fence_after();
        for (int ks = 0; ks < nks; ks++, it++) {
          const uint32_t s = it % STAGES;
          mbar_wait(full_bar + 8 * s, (it / STAGES) & 1);
          tc_fence_after();
          const int kleft = k - ks * KSTEP_BITS;
          const int nmma = kleft >= KSTEP_BITS ? 8 : (kleft + 31) >> 5;
          const uint32_t a_t = tmem_base + TMEM_A_COL + s * 64;
          const uint64_t bd = make_desc_sw128(bexp_base + s * B_STAGE);
#pragma unroll
          for (int i = 0; i < 8; i++) {
            if (i < nmma) {
              // K = 32 bytes per MMA: 8 TMEM columns of A; 32 bytes (+2 in the >>4 address
              // field) within the 128-byte swizzle row of Bt, next 16 KiB tile after four
              const uint64_t bdi = bd + (uint64_t)((i >> 2) * (B_TILE >> 4) + (i & 3) * 2);
              umma_i8_ts<CG>(tmem_base, a_t + 8 * i, bdi, idesc, (ks | i) ? 1u : 0u);
            }
          }
          umma_commit<CG>(empty_bar + 8 * s);
        }
        umma_commit<CG>(accum_bar);
      }
    }
  } else if (warp == 9) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t rit = 0;
      for (int t = cluster_id; t < ntiles; t += nclusters) {
        int tm, tn;
        tile_coords(t, tiles_m, tiles_n, tm, tn);
        const int arow = (tm * CG + (int)rank) * 128, brow = (tn * CG + (int)rank) * 128;
        for (int ri = 0; ri < nrs; ri++, rit++) {
          const uint32_t rs = rit % RAW_STAGES;
          mbar_wait(rawempty_bar + 8 * rs, ((rit / RAW_STAGES) & 1) ^ 1);
          mbar_expect_tx(rawfull_bar + 8 * rs, RAW_STAGE);
          tma_load_2d(raw_base + rs * RAW_STAGE, &tmA, ri * 128, arow, rawfull_bar + 8 * rs);
          tma_load_2d(raw_base + rs * RAW_STAGE + RAW_TILE, &tmB, ri * 128, brow, rawfull_bar + 8 * rs);
        }
      }
    }
  } else {
    // ================= epilogue: warps 10-13, TMEM lane quarter = warp % 4 =================
    const int qd = warp & 3;
    const uint32_t dempty_leader = map_to_rank(dempty_bar, 0);
    uint32_t tile_it = 0;
    for (int t = cluster_id; t < ntiles; t += nclusters, tile_it++) {
      int tm, tn;
      tile_coords(t, tiles_m, tiles_n, tm, tn);
      mbar_wait(accum_bar, tile_it & 1);
      tc_fence_after();
      uint32_t out[NCOL / 32];
#pragma unroll
      for (int c = 0; c < NCOL / 32; c++) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(qd * 32) << 16) + c * 32, v);
        uint32_t word = 0;
#pragma unroll
        for (int j = 0; j < 32; j++) word |= (v[j] & 1u) << j;
        out[c] = word;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(dempty_leader);  // D may be overwritten by the next tile
      const int row = (tm * CG + (int)rank) * 128 + qd * 32 + lane;
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
  if (warp == 8) tmem_dealloc_all<CG>(tmem_base);
}

}  // namespace

// ---- host side -------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                    const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// libcuda is not linked: the encoder comes from the driver the runtime already loaded
static encode_tiled_fn tensor_map_encoder() {
  static encode_tiled_fn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    M4B_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    if (!p || q != cudaDriverEntryPointSuccess)
      die("m4ri_b200: cuTensorMapEncodeTiled is not available from this driver\n");
    fn = (encode_tiled_fn)p;
  }
  return fn;
}

// rows of packed bits as a 2-D byte tensor: box = 128 bytes (1024 k-bits) x 128 rows,
// 128B swizzle, out-of-range bytes and rows read as zero
CUtensorMap packed_rows_map(const word *d, int nrows, int width, int rowstride) {
  CUtensorMap map;
  cuuint64_t gdim[2] = {(cuuint64_t)width * 8, (cuuint64_t)nrows};
  cuuint64_t gstride[1] = {(cuuint64_t)rowstride * 8};
  cuuint32_t box[2] = {128, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = tensor_map_encoder()(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void *)d, gdim, gstride,
                                    box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) die("m4ri_b200: cuTensorMapEncodeTiled failed (%d)\n", (int)r);
  return map;
}

template <int CG>
static void launch_mma2(DMat *C, const DMat *A, const DMat *Bt, int clear, cudaStream_t s) {
  static std::atomic<unsigned long long> attr_seen{0};
  if (first_use_on_this_device(attr_seen)) {
    M4B_CUDA(cudaFuncSetAttribute(mma2_kernel<CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_V2));
  }
  const int tile = 128 * CG;
  const int tiles_m = (C->nrows + tile - 1) / tile, tiles_n = (C->ncols + tile - 1) / tile;
  const int ntiles = tiles_m * tiles_n;
  int nclusters = sm_count() / CG;
  if (nclusters > ntiles) nclusters = ntiles;
  for (int k0 = 0; k0 < A->ncols; k0 += K_CHUNK_MAX) {
    const int kc = A->ncols - k0 < K_CHUNK_MAX ? A->ncols - k0 : K_CHUNK_MAX;
    const CUtensorMap tmA = packed_rows_map(A->d + k0 / 64, A->nrows, width_of(kc), A->rowstride);
    const CUtensorMap tmB = packed_rows_map(Bt->d + k0 / 64, Bt->nrows, width_of(kc), Bt->rowstride);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nclusters * CG);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = SMEM_V2;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    timing_begin(CG == 2 ? "mma2_kernel<2>" : "mma2_kernel<1>", s, 2.0 * C->nrows * C->ncols * kc);
    M4B_CUDA(cudaLaunchKernelEx(&cfg, mma2_kernel<CG>, (uint32_t *)C->d, tmA, tmB, C->nrows, C->ncols,
                                kc, C->rowstride * 2, C->width * 2, tiles_m, tiles_n,
                                (clear && k0 == 0) ? 1 : 0));
    M4B_LAUNCHED();
    timing_end(s);
  }
}

// C (^)= A*Bt^T; cg = CTAs per MMA (1 or 2)
void k_mul_abt_mma2(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, cudaStream_t s) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    return;
  }
  if (!has_tcgen05()) die("m4ri_b200: the tcgen05 MMA path needs an sm_100 device\n");
  if (!mma_shape_ok(C, A, Bt)) die("m4ri_b200: mma kernel needs 16-byte aligned rows\n");
  if (cg == 2) launch_mma2<2>(C, A, Bt, clear, s);
  else launch_mma2<1>(C, A, Bt, clear, s);
}

}  // namespace m4b
