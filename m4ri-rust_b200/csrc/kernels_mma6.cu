// kernels_mma6.cu — path (b) v5, the default: the FP4 contraction of kernels_mma5.cu with a
// load path the compiler can pipeline, a cheaper drain, and a compile-time pipeline shape.
//
// C (^)= A * Bt^T, A (m x k bits) and Bt (n x k bits) both k-contiguous (the layout of
// _mzd_mul_naive, m4ri-sys/src/mzd.rs:163-168; mzd_mul, m4ri-sys/src/strassen.rs:18, gets
// here through one bit-transpose of B).
//
// What round 2's isolation experiments showed (profiles/r2a_dbg_isolation_mma5.log,
// profiles/r2c_mma6_explore*.log; leaf 16384^3, mma5 = 1.173 ms):
//   * mma5 without operand expansion 1.080, without the drain 1.139, with neither 1.033 ms — and
//     issue slots 35 %, ALU 46 %: latency, not throughput.
//   * The latency was the packed-bit loads: mma5 issues them CONDITIONALLY (ragged rows, k tails,
//     last tile), so ptxas cannot use counted scoreboard waits and the first use of a buffer
//     waits for ALL outstanding loads — the "one stage ahead" prefetch never overlapped
//     anything. With whole tiles and whole stages (template FULL) every load is one
//     unconditional 256-bit LDG: the expansion drops from 0.093 to 0.003 ms of exposed time
//     (1.173 -> 1.085 ms).
//   * More warps make it WORSE, not better: 16 expander warps 1.29 ms, 256-bit stages 1.48-1.60,
//     8 epilogue warps 1.10 (register cap 96 at 17 warps); deeper prefetch changes nothing.
//     The shape kept is mma5's: 512-bit stages x 3, 8 expander warps, 4 epilogue warps.
//   * Parity packing by fma(v, 2^j, 2^21) & (1 << j) (one FFMA + one LOP3 per accumulator instead
//     of FADD + AND + shift-or): 1.085 -> 1.068 ms. Needs accumulators <= 8192: k chunks of 32768.
//   * Tile order in groups of 16 tile-rows: 8.20 -> 8.17 ms on the 32768^3 single launch.
//   Floor of the MMA stream itself under these barriers: 1.033 ms (no expansion, no drain),
//   1.014 ms without the per-tile accumulator hand-over — 8.5-8.7 Pop/s, the measured FP4 pipe
//   rate of this MMA shape (kind::mxf4, cta_group::2, A from tensor memory).
//
//   warps 0-3    A rows: LDG 512 k-bits -> 512 E2M1 nibbles -> TENSOR MEMORY (tcgen05.st)
//   warps 4-7    Bt rows: the same into 128B-swizzled K-major shared memory
//   warp  8      one lane of the leader CTA issues 8 tcgen05.mma.kind::mxf4 (K = 64) per stage
//   warps 9-12   epilogue: TMEM -> registers -> parity bits -> C
// Exactness: products are 0 or 0.25, fp32 sums of at most 32768 of them (exact: tests
// test_fp4_accumulators_are_exact, test_fp4_accumulators_exact_on_dense_operands).
#include <cuda.h>

#include "common.cuh"
#include "tcgen05_util.cuh"

namespace m4b {
namespace {

using namespace tc;

// Pipeline shape (compile-time; the -D overrides exist for the A/B runs of tools/mma6_explore.sh):
//   MMA6_KS    k bits per stage (256 or 512); the ring always holds 1536 k-bits
//   MMA6_G     expander warp groups per operand (1: four warps, one per row quarter; 2: eight)
//   MMA6_NEPI  epilogue warps (4 or 8)
//   MMA6_PF    own-stages the packed-bit loads run ahead of their expansion (1 or 2)
//   MMA6_HINT  ns hint of the expanders' suspended mbarrier waits (0: plain try_wait polling)
#ifndef MMA6_KS
#define MMA6_KS 512
#endif
#ifndef MMA6_G
#define MMA6_G 1
#endif
#ifndef MMA6_NEPI
#define MMA6_NEPI 4
#endif
#ifndef MMA6_PF
#define MMA6_PF 1
#endif
#ifndef MMA6_HINT
#define MMA6_HINT 0
#endif
#ifndef MMA6_ST   // columns per tcgen05.st of the A expanders (16 or 32)
#define MMA6_ST 32
#endif
#ifndef MMA6_LD   // columns per tcgen05.ld of the epilogue (16 or 32)
#define MMA6_LD 32
#endif
#ifndef MMA6_PACK // 1: FFMA + LOP3 parity packing (needs accumulators <= 8192: k chunks of 32768)
#define MMA6_PACK 1
#endif
#ifndef MMA6_TG   // tile-rows per L2 group of the tile order
#define MMA6_TG 16
#endif
constexpr int KS_BITS = MMA6_KS;           // k bits per pipeline stage (KS_BITS / 64 MMAs of K = 64 nibbles)
constexpr int KS_BYTES = KS_BITS / 8;      // packed bytes of a row per stage
constexpr int STAGES = 1536 / KS_BITS;     // expanded-operand ring (A: TMEM, Bt: smem)
constexpr int UNITS = KS_BITS / 256;       // 256-bit units per stage
constexpr int G = MMA6_G;                  // expander warp groups per operand
// (KS 512, G 2): group g expands unit g of every stage. (KS 256, G 2): group g takes the stages
// with index = g (mod 2). G 1: the four warps expand every unit of every stage.
constexpr bool ALT_STAGES = (G == 2 && UNITS == 1);
constexpr int UPT = (G == 2 && UNITS == 2) ? 1 : UNITS;  // units a thread expands per own stage
constexpr int ARRIVALS = 2 * 4 * (ALT_STAGES ? 1 : G);   // expander warps arriving per stage and CTA
constexpr int A_COLS = KS_BITS / 8;        // TMEM columns of one A stage (8 nibbles per column)
constexpr int B_TILE = 128 * 128;          // 128 rows x 128 B = 256 nibbles per row, SWIZZLE_128B
constexpr int B_STAGE = UNITS * B_TILE;
constexpr int NEXP = 8 * G;                // expander warps: 4G A + 4G Bt
constexpr int NEPI = MMA6_NEPI;            // epilogue warps
constexpr int ISSUER = NEXP;               // warp index of the MMA issuer
constexpr int EPI0 = ISSUER + 1;           // first epilogue warp (TMEM lane quarter = warp % 4, any start works)
constexpr int NWARPS = EPI0 + NEPI;
constexpr int NTHREADS = NWARPS * 32;
constexpr int SMEM_V6 = STAGES * B_STAGE + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TMEM_A_COL = 256;            // A stages in [256, 448)
constexpr int TMEM_SF_COL = 448;           // scale factors (all 2^0) in [448, 512)
constexpr int K_CHUNK_MAX = MMA6_PACK == 1 ? 32768 : 65536;  // host-side chunking of the contraction (accumulator range)
constexpr int TILE_GROUP = MMA6_TG;        // tile-rows per L2 group
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
  const int grp = t / (TILE_GROUP * tiles_n), in_grp = t - grp * TILE_GROUP * tiles_n;
  const int grp_rows = min(TILE_GROUP, tiles_m - grp * TILE_GROUP);
  tm = grp * TILE_GROUP + in_grp % grp_rows;
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

// The 8 words (256 k-bits, one unit) of one row starting at p; `rem` = valid bytes of the row
// from p on (a multiple of 8; <= 0 for rows past the matrix). Words past the row — another
// quadrant's data when the operand is a window — read as zero.
template <bool WIDE>
__device__ __forceinline__ void load_unit(uint32_t *x, const uint8_t *p, int rem) {
  if (rem >= 32) {
    if constexpr (WIDE) {
      ldg256(&x[0], p);
    } else {
      ldg128(&x[0], p);
      ldg128(&x[4], p + 16);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 2; j++) {
      if (rem > 16 * j) {
        ldg128(&x[4 * j], p + 16 * j);  // rows are padded to 16 bytes: in bounds
        if (rem < 16 * j + 16) x[4 * j + 2] = x[4 * j + 3] = 0u;  // rem = 16 j + 8
      } else {
        x[4 * j] = x[4 * j + 1] = x[4 * j + 2] = x[4 * j + 3] = 0u;
      }
    }
  }
}

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
      "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}

// bit j of the result = parity of 4 * v[j] (v[j] = popcount / 4 exactly, < 2^21): the mantissa
// LSB of v[j] + 2^21. Two independent shift-or chains of eight.
__device__ __forceinline__ uint32_t pack_parity16(const uint32_t (&v)[16]) {
  uint32_t w[2];
#pragma unroll
  for (int g = 0; g < 2; g++) {
    uint32_t acc = 0;
#pragma unroll
    for (int j = 7; j >= 0; j--)
      acc = (acc << 1) | (__float_as_uint(__uint_as_float(v[8 * g + j]) + 2097152.0f) & 1u);
    w[g] = acc;
  }
  return w[0] | (w[1] << 8);
}

// the same for 32 columns.
// MMA6_PACK 1: fma(v, 2^j, 2^21) puts the parity bit (weight 0.25 of v) at mantissa bit j with
// nothing below it, so a value costs one FFMA and one LOP3 (acc | (t & (1 << j))) instead of
// FADD + AND + shift-or. Exact while v * 2^7 < 2^21, i.e. v <= 8192: the host chunks the
// contraction at 32768 bits for this variant. Four byte accumulators, assembled with two PRMTs.
#ifndef MMA6_PACK
#define MMA6_PACK 1
#endif
__device__ __forceinline__ uint32_t pack_parity(const uint32_t (&v)[32]) {
  uint32_t w[4];
#if MMA6_PACK == 1
#pragma unroll
  for (int g = 0; g < 4; g++) {
    uint32_t acc = 0;
#pragma unroll
    for (int j = 0; j < 8; j++)
      acc |= __float_as_uint(__fmaf_rn(__uint_as_float(v[8 * g + j]), (float)(1 << j), 2097152.0f)) & (1u << j);
    w[g] = acc;
  }
  return __byte_perm(__byte_perm(w[0], w[1], 0x0040), __byte_perm(w[2], w[3], 0x0040), 0x5410);
#else
#pragma unroll
  for (int g = 0; g < 4; g++) {
    uint32_t acc = 0;
#pragma unroll
    for (int j = 7; j >= 0; j--)
      acc = (acc << 1) | (__float_as_uint(__uint_as_float(v[8 * g + j]) + 2097152.0f) & 1u);
    w[g] = acc;
  }
  return w[0] | (w[1] << 8) | (w[2] << 16) | (w[3] << 24);
#endif
}

#ifdef M4RI_B200_PERF_HOOKS
#define M4B_DBG(bit) (dbg & (bit))
#else
#define M4B_DBG(bit) false
#endif

// FULL: m and n are whole tiles and k whole stages (the benchmark shapes): every packed-bit load
// is one unconditional 256-bit LDG, so that ptxas can keep several of them in flight behind
// counted scoreboard waits. (With conditional loads of unknown count the first use of a buffer
// waited for ALL outstanding loads — the prefetch was lost and the expansion cost 0.32 ms
// instead of 0.09 ms per 16384^3 leaf, profiles/r2c_mma6_explore*.log.)
// One launch = `count` products of ONE shape (m x k) * (n x k)^T, tiles numbered product-major:
// the seven products of a fused Strassen level run as one persistent grid (no launch gaps, one
// partial last wave instead of seven). Each product has one or two destinations of the same shape
// and stride, overwritten (clear) or accumulated into. Products of one launch may share
// destinations: then `atomic` is set, every destination is pre-zeroed by the caller and all of
// them accumulate with red.global.xor (XOR commutes: the result does not depend on the order).
constexpr int GROUP_MAX = 7;
struct Mma6Group {
  int count, atomic;
  const uint8_t *A[GROUP_MAX], *B[GROUP_MAX];
  long long a_rsb[GROUP_MAX], b_rsb[GROUP_MAX];
  uint32_t *C1[GROUP_MAX], *C2[GROUP_MAX];
  int clear1[GROUP_MAX], clear2[GROUP_MAX];
};
__device__ __forceinline__ void red_xor64(uint32_t *p, uint32_t lo, uint32_t hi) {
  asm volatile("red.global.xor.b64 [%0], %1;" ::"l"(p), "l"((unsigned long long)lo | ((unsigned long long)hi << 32))
               : "memory");
}
__device__ __forceinline__ void red_xor32(uint32_t *p, uint32_t v) {
  asm volatile("red.global.xor.b32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int CG, bool WIDE, bool FULL>
__global__ void __launch_bounds__(NTHREADS, 1)
mma6_kernel(const __grid_constant__ Mma6Group grp, int m, int n, int k, int cs32, int cw32, int tiles_m,
            int tiles_n, int dbg) {
  constexpr int NCOL = 128 * CG;  // accumulator columns = tile columns
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B: 1024-B alignment
  const uint32_t bexp_base = sbase;
  const uint32_t bars = bexp_base + STAGES * B_STAGE;
#ifdef MMA6_BARS_WIDE
  const uint32_t full_bar = bars, empty_bar = bars + 64, accum_bar = bars + 128, dempty_bar = bars + 136,
                 tmem_slot = bars + 144;
#else
  const uint32_t full_bar = bars, empty_bar = bars + 32, accum_bar = bars + 64, dempty_bar = bars + 72,
                 tmem_slot = bars + 80;
  static_assert(STAGES <= 4, "barrier block: 4 x 8 bytes per ring");
#endif
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 1 ? 0u : cluster_ctarank();
  const int cluster_id = blockIdx.x / CG, nclusters = gridDim.x / CG;
  const int tiles_per = tiles_m * tiles_n, ntiles = tiles_per * grp.count;
  const int nks = (k + KS_BITS - 1) / KS_BITS;  // stages per tile
  const int my_tiles = cluster_id < ntiles ? (ntiles - cluster_id + nclusters - 1) / nclusters : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(full_bar + 8 * s, ARRIVALS * CG);    // the expander warps that own (a share of) the stage, in every CTA
      mbar_init(empty_bar + 8 * s, 1);               // tcgen05.commit
    }
    mbar_init(accum_bar, 1);                         // tcgen05.commit after the last stage of a tile
    mbar_init(dempty_bar, NEPI * CG);                // one arrival per epilogue warp of every CTA
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == ISSUER) tmem_alloc_all<CG>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();  // the peer's barriers exist before anyone arrives on them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp < NEXP) {
    // ============ expanders: one thread = one row of the CTA's 128-row slab ============
    const bool isA = warp < 4 * G;
    const int g = G == 2 ? (warp >> 2) & 1 : 0;  // warp group within the operand
    const int r = (warp & 3) * 32 + lane;        // row of the slab = TMEM lane
    const uint32_t sw = (uint32_t)(r & 7);
    const uint32_t brow = bexp_base + (r >> 3) * 1024 + (r & 7) * 128;
    const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t a_lane = lane_base + TMEM_A_COL;
    const uint32_t full_leader = map_to_rank(full_bar, 0);
    const int nrows = isA ? m : n;
    const int wbytes = ((k + 63) >> 6) << 3;  // valid bytes of a row (whole words; excess bits are zero)
    if (isA && g == 0) {
      // every scale factor the MMAs can address is UE8M0 127 = 2^0 (any layout reads the same byte)
      uint32_t one[32];
#pragma unroll
      for (int j = 0; j < 32; j++) one[j] = 0x7F7F7F7Fu;
      tmem_st32(lane_base + TMEM_SF_COL, one);
      tmem_st32(lane_base + TMEM_SF_COL + 32, one);
      tmem_wait_st();
      tc_fence_before();
    }
    // own iteration j = global stage `it` (ALT_STAGES: 2j + g); within it the thread expands UPT
    // units, the first at unit offset u0 of the stage
    constexpr int STEP = ALT_STAGES ? 2 : 1;
    const int u0 = (G == 2 && UNITS == 2) ? g : 0;
    const int first = ALT_STAGES ? g : 0;
    const int total = my_tiles * nks;  // stages of this cluster (< 2^31: at most 2^16 tiles x 2^9 stages)
    const int own = total > first ? (total - first + STEP - 1) / STEP : 0;
    // ---- load cursor ----
    int l_ti = 0, l_ks = first, l_w = 0, l_cur_ti = -1;
    const uint8_t *l_p = isA ? grp.A[0] : grp.B[0];
    auto load_next = [&](uint32_t (&x)[8 * UPT]) {
      while (l_ks >= nks) { l_ks -= nks; l_ti++; }
      if constexpr (FULL) {
        // past the end the last tile's bits are loaded again (never expanded): no conditional load
        const int ti = l_ti < my_tiles ? l_ti : my_tiles - 1;
        if (ti != l_cur_ti) {
          int tm, tn;
          const int t = cluster_id + ti * nclusters, pr = t / tiles_per;
          tile_coords(t - pr * tiles_per, tiles_m, tiles_n, tm, tn);
          l_p = (isA ? grp.A[pr] : grp.B[pr]) +
                (size_t)(((isA ? tm : tn) * CG + (int)rank) * 128 + r) * (isA ? grp.a_rsb[pr] : grp.b_rsb[pr]);
          l_cur_ti = ti;
        }
        const uint8_t *p = l_p + (l_ti < my_tiles ? l_ks : 0) * KS_BYTES + u0 * 32;
#pragma unroll
        for (int u = 0; u < UPT; u++) ldg256(&x[8 * u], p + 32 * u);
      } else {
        if (l_ti >= my_tiles) return;  // past the end: the buffer is never expanded
        if (l_ti != l_cur_ti) {
          int tm, tn;
          const int t = cluster_id + l_ti * nclusters, pr = t / tiles_per;
          tile_coords(t - pr * tiles_per, tiles_m, tiles_n, tm, tn);
          const int row = ((isA ? tm : tn) * CG + (int)rank) * 128 + r;
          const bool valid = row < nrows;
          l_p = (isA ? grp.A[pr] : grp.B[pr]) + (size_t)(valid ? row : 0) * (isA ? grp.a_rsb[pr] : grp.b_rsb[pr]);
          l_w = valid ? wbytes : 0;
          l_cur_ti = l_ti;
        }
        const int off = l_ks * KS_BYTES + u0 * 32;
#pragma unroll
        for (int u = 0; u < UPT; u++) load_unit<WIDE>(&x[8 * u], l_p + off + 32 * u, l_w - off - 32 * u);
      }
      l_ks += STEP;
    };
    // turns the bits in x into nibbles and publishes this warp's share of the stage of own iteration j
    int j = 0;
    auto expand = [&](const uint32_t (&x)[8 * UPT]) {
      const uint32_t it = (uint32_t)(STEP * j + first);
      const uint32_t s = it % STAGES;
#if MMA6_HINT > 0
      mbar_wait_long(empty_bar + 8 * s, ((it / STAGES) & 1) ^ 1, MMA6_HINT);
#else
      mbar_wait(empty_bar + 8 * s, ((it / STAGES) & 1) ^ 1);
#endif
      if (M4B_DBG(1)) {
      } else if (isA) {
        tc_fence_after();
#if MMA6_ST == 16
#pragma unroll
        for (int h = 0; h < 2 * UPT; h++) {  // 16-column stores: few live registers
          uint32_t v[16];
#pragma unroll
          for (int c = 0; c < 4; c++)
#pragma unroll
            for (int b = 0; b < 4; b++) v[4 * c + b] = (x[4 * h + c] >> b) & NIB;
          tmem_st16(a_lane + s * A_COLS + u0 * 32 + h * 16, v);
        }
#else
#pragma unroll
        for (int h = 0; h < UPT; h++) {
          uint32_t v[32];
#pragma unroll
          for (int c = 0; c < 8; c++)
#pragma unroll
            for (int b = 0; b < 4; b++) v[4 * c + b] = (x[8 * h + c] >> b) & NIB;
          tmem_st32(a_lane + s * A_COLS + u0 * 32 + h * 32, v);
        }
#endif
        tmem_wait_st();
        tc_fence_before();
      } else {
#pragma unroll
        for (int u = 0; u < UPT; u++) {
          const uint32_t dst = brow + s * B_STAGE + (u0 + u) * B_TILE;
#pragma unroll
          for (int c = 0; c < 8; c++) {
            const uint32_t w = x[8 * u + c];
            sts128(dst + (((uint32_t)c ^ sw) << 4), w & NIB, (w >> 1) & NIB, (w >> 2) & NIB, (w >> 3) & NIB);
          }
        }
        fence_proxy_async();
      }
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(full_leader + 8 * s);
      j++;
    };
#if MMA6_PF == 2
    uint32_t x0[8 * UPT], x1[8 * UPT], x2[8 * UPT];
    load_next(x0);
    load_next(x1);
    while (j < own) {
      load_next(x2);
      expand(x0);
      if (j >= own) break;
      load_next(x0);
      expand(x1);
      if (j >= own) break;
      load_next(x1);
      expand(x2);
    }
#else
    uint32_t x0[8 * UPT], x1[8 * UPT];
    load_next(x0);
    while (j < own) {
      load_next(x1);
      expand(x0);
      if (j >= own) break;
      load_next(x0);
      expand(x1);
    }
#endif
  } else if (warp == ISSUER) {
    // ================= MMA issuer: one lane of the leader CTA =================
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc_mxf4(128 * CG, NCOL);
      constexpr int MPS = KS_BITS / 64;  // MMAs per stage
      const uint32_t sfa = tmem_base + TMEM_SF_COL, sfb = tmem_base + TMEM_SF_COL + 32;
      uint32_t it = 0;
      for (int ti = 0; ti < my_tiles; ti++) {
        if (ti > 0 && !M4B_DBG(8)) mbar_wait(dempty_bar, (ti - 1) & 1);  // epilogue has drained D
        tc_fence_after();
        for (int ks = 0; ks < nks; ks++, it++) {
          const uint32_t s = it % STAGES;
          mbar_wait(full_bar + 8 * s, (it / STAGES) & 1);
          tc_fence_after();
          const int kleft = k - ks * KS_BITS;
          const int nmma = kleft >= KS_BITS ? MPS : (kleft + 63) >> 6;
          const uint32_t a_t = tmem_base + TMEM_A_COL + s * A_COLS;
          const uint64_t bd = make_desc_sw128(bexp_base + s * B_STAGE);
#pragma unroll
          for (int i = 0; i < MPS; i++) {
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
  } else if (warp >= EPI0) {
    // ===== epilogue: NEPI warps, TMEM lane quarter = warp % 4, column part = (warp - EPI0) / 4 =====
    const int e = warp - EPI0;
    const int qd = warp & 3, part = e >> 2;
    constexpr int HCOL = NCOL / (NEPI / 4);  // columns per epilogue warp
    const uint32_t dempty_leader = map_to_rank(dempty_bar, 0);
    for (int ti = 0; ti < my_tiles; ti++) {
      int tm, tn;
      const int t = cluster_id + ti * nclusters, pr = t / tiles_per;
      tile_coords(t - pr * tiles_per, tiles_m, tiles_n, tm, tn);
      const int row = (tm * CG + (int)rank) * 128 + qd * 32 + lane;
#if defined(MMA6_SPIN_ACCUM)
      mbar_wait(accum_bar, ti & 1);
#else
      mbar_wait_long(accum_bar, ti & 1);
#endif
      tc_fence_after();
      uint32_t out[HCOL / 32];
#pragma unroll
      for (int c = 0; c < HCOL / 32; c++) out[c] = 0;
      if (!M4B_DBG(4)) {
        // two chunks in flight: the load of chunk c+1 overlaps the packing of chunk c
        const uint32_t d_lane = tmem_base + ((uint32_t)(qd * 32) << 16) + part * HCOL;
#if MMA6_LD == 16
        uint32_t va[16], vb[16];
        tmem_ld16_nowait(d_lane, va);
#pragma unroll
        for (int c = 0; c < HCOL / 16; c += 2) {
          tmem_wait_ld();
          tmem_ld16_nowait(d_lane + (c + 1) * 16, vb);
          out[c >> 1] = pack_parity16(va);
          tmem_wait_ld();
          if (c + 2 < HCOL / 16) {
            tmem_ld16_nowait(d_lane + (c + 2) * 16, va);
          } else {
            // the last load has landed in registers: D may be overwritten by the next tile
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(dempty_leader);
          }
          out[c >> 1] |= pack_parity16(vb) << 16;
        }
#else
        uint32_t va[32], vb[32];
        tmem_ld32_nowait(d_lane, va);
#pragma unroll
        for (int c = 0; c < HCOL / 32; c += 2) {
          tmem_wait_ld();
          tmem_ld32_nowait(d_lane + (c + 1) * 32, vb);
          out[c] = pack_parity(va);
          tmem_wait_ld();
          if (c + 2 < HCOL / 32) {
            tmem_ld32_nowait(d_lane + (c + 2) * 32, va);
          } else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(dempty_leader);
          }
          out[c + 1] = pack_parity(vb);
        }
#endif
      } else {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(dempty_leader);
      }
      const int cw0 = tn * (NCOL / 32) + part * (HCOL / 32);
      // one or two destinations with the same shape and stride (the fused Strassen level sends a
      // product into two quadrants of C), each overwritten or accumulated into
      for (int tgt = 0; tgt < 2; tgt++) {
        uint32_t *Ct = tgt ? grp.C2[pr] : grp.C1[pr];
        const int clr = tgt ? grp.clear2[pr] : grp.clear1[pr];
        if (Ct == nullptr || row >= m) continue;
        uint32_t *dst = Ct + (size_t)row * cs32 + cw0;
        if (grp.atomic) {  // destinations shared with other products of this launch (pre-zeroed)
#pragma unroll
          for (int jj = 0; jj < HCOL / 32; jj += 2) {
            if (cw0 + jj + 2 <= cw32) red_xor64(dst + jj, out[jj], out[jj + 1]);
            else if (cw0 + jj < cw32) red_xor32(dst + jj, out[jj]);
          }
          continue;
        }
        if constexpr (HCOL >= 128) {
#pragma unroll
          for (int q4 = 0; q4 < HCOL / 128; q4++) {
            const int cw = cw0 + 4 * q4;
            if (cw + 4 <= cw32) {
              uint4 o = make_uint4(out[4 * q4], out[4 * q4 + 1], out[4 * q4 + 2], out[4 * q4 + 3]);
              if (!clr) {
                const uint4 old = *reinterpret_cast<const uint4 *>(dst + 4 * q4);
                o.x ^= old.x; o.y ^= old.y; o.z ^= old.z; o.w ^= old.w;
              }
              *reinterpret_cast<uint4 *>(dst + 4 * q4) = o;
            } else {
#pragma unroll
              for (int jj = 0; jj < 4; jj++)
                if (cw + jj < cw32)
                  dst[4 * q4 + jj] = clr ? out[4 * q4 + jj] : (dst[4 * q4 + jj] ^ out[4 * q4 + jj]);
            }
          }
        } else {  // CG == 1 with eight epilogue warps: 64 columns = two words per warp
#pragma unroll
          for (int jj = 0; jj < HCOL / 32; jj++)
            if (cw0 + jj < cw32) dst[jj] = clr ? out[jj] : (dst[jj] ^ out[jj]);
        }
      }
    }
  }
  // ---- teardown: nobody leaves while the pair's MMAs may still read its memory ----
  tc_fence_before();
  __syncthreads();
  if constexpr (CG == 2) cluster_sync_all();
  if (warp == ISSUER) tmem_dealloc_all<CG>(tmem_base);
}

}  // namespace

// ---- host side -------------------------------------------------------------------
// One product per entry of the arrays (count <= GROUP_MAX), all of one shape.
struct Mma6Job {
  DMat *C1, *C2;
  const DMat *A, *Bt;
  int clear1, clear2;
};
template <int CG, bool WIDE, bool FULL>
static void launch_mma6(const Mma6Job *jobs, int count, int atomic, cudaStream_t s) {
  static std::atomic<unsigned long long> attr_seen{0};
  if (first_use_on_this_device(attr_seen)) {
    M4B_CUDA(cudaFuncSetAttribute(mma6_kernel<CG, WIDE, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_V6));
  }
  const DMat *C = jobs[0].C1, *A = jobs[0].A;
  const int tile = 128 * CG;
  const int tiles_m = (C->nrows + tile - 1) / tile, tiles_n = (C->ncols + tile - 1) / tile;
  const int ntiles = tiles_m * tiles_n * count;
  int nclusters = sm_count() / CG;
  if (nclusters > ntiles) nclusters = ntiles;
#ifdef M4RI_B200_PERF_HOOKS
  static const int dbg = getenv("M4RI_B200_DBG") ? atoi(getenv("M4RI_B200_DBG")) : 0;  // perf-hook builds only
#else
  const int dbg = 0;
#endif
  for (int k0 = 0; k0 < A->ncols; k0 += K_CHUNK_MAX) {
    const int kc = A->ncols - k0 < K_CHUNK_MAX ? A->ncols - k0 : K_CHUNK_MAX;
    Mma6Group grp = {};
    grp.count = count;
    grp.atomic = atomic;
    for (int i = 0; i < count; i++) {
      grp.A[i] = (const uint8_t *)(jobs[i].A->d + k0 / 64);
      grp.B[i] = (const uint8_t *)(jobs[i].Bt->d + k0 / 64);
      grp.a_rsb[i] = (long long)jobs[i].A->rowstride * 8;
      grp.b_rsb[i] = (long long)jobs[i].Bt->rowstride * 8;
      grp.C1[i] = (uint32_t *)jobs[i].C1->d;
      grp.C2[i] = jobs[i].C2 ? (uint32_t *)jobs[i].C2->d : nullptr;
      grp.clear1[i] = (jobs[i].clear1 && k0 == 0) ? 1 : 0;
      grp.clear2[i] = (jobs[i].clear2 && k0 == 0) ? 1 : 0;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nclusters * CG);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = SMEM_V6;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    timing_begin(CG == 2 ? "mma6_kernel<2>" : "mma6_kernel<1>", s, 2.0 * C->nrows * C->ncols * kc * count);
    M4B_CUDA(cudaLaunchKernelEx(&cfg, mma6_kernel<CG, WIDE, FULL>, grp, C->nrows, C->ncols, kc, C->rowstride * 2,
                                C->width * 2, tiles_m, tiles_n, dbg));
    M4B_LAUNCHED();
    timing_end(s);
  }
}

static bool my_tiles_ok(int m, int n, int k, int tile) {
  return m % tile == 0 && n % tile == 0 && k % KS_BITS == 0 && (k <= K_CHUNK_MAX || K_CHUNK_MAX % KS_BITS == 0);
}

static void dispatch_mma6(const Mma6Job *jobs, int count, int atomic, int cg, cudaStream_t s) {
  const DMat *C = jobs[0].C1, *A = jobs[0].A;
  bool wide = true;
  for (int i = 0; i < count; i++) {
    const Mma6Job &j = jobs[i];
    if (!mma_shape_ok(j.C1, j.A, j.Bt)) die("m4ri_b200: mma kernel needs 16-byte aligned rows\n");
    if (j.C1->nrows != C->nrows || j.C1->ncols != C->ncols || j.A->ncols != A->ncols || j.C1->rowstride != C->rowstride ||
        ((uintptr_t)j.C1->d & 15) ||
        (j.C2 && (j.C2->nrows != C->nrows || j.C2->ncols != C->ncols || j.C2->rowstride != C->rowstride ||
                  ((uintptr_t)j.C2->d & 15))))
      die("m4ri_b200: the destinations of one launch must share shape, stride and alignment\n");
    // 256-bit loads need 32-byte aligned rows (true for every width that is a multiple of 256 bits)
    wide = wide && !(((uintptr_t)j.A->d | (uintptr_t)j.Bt->d) & 31) && !(j.A->rowstride & 3) && !(j.Bt->rowstride & 3);
  }
  const int tile = 128 * cg;
  // whole tiles and whole stages in every k-chunk of the launch loop: the unconditional-load kernel
  const bool full = wide && my_tiles_ok(C->nrows, C->ncols, A->ncols, tile);
  if (cg == 2) {
    if (full) launch_mma6<2, true, true>(jobs, count, atomic, s);
    else if (wide) launch_mma6<2, true, false>(jobs, count, atomic, s);
    else launch_mma6<2, false, false>(jobs, count, atomic, s);
  } else {
    if (full) launch_mma6<1, true, true>(jobs, count, atomic, s);
    else if (wide) launch_mma6<1, true, false>(jobs, count, atomic, s);
    else launch_mma6<1, false, false>(jobs, count, atomic, s);
  }
}

// C (^)= A*Bt^T on the FP4 pipe; cg = CTAs per MMA. With C2: the same product also goes to
// (clear2) or into (!clear2) C2, a matrix of C's shape and row stride.
void k_mul_abt_mma6(DMat *C, const DMat *A, const DMat *Bt, int clear, int cg, cudaStream_t s, DMat *C2, int clear2) {
  if (C->nrows == 0 || C->ncols == 0) return;
  if (A->ncols == 0) {
    if (clear) k_zero(C, s);
    if (C2 && clear2) k_zero(C2, s);
    return;
  }
  if (!has_tcgen05()) die("m4ri_b200: the tcgen05 MMA path needs an sm_100 device\n");
  Mma6Job job = {C, C2, A, Bt, clear, clear2};
  dispatch_mma6(&job, 1, 0, cg, s);
}

// `count` products of one shape in ONE launch; every C1[i]/C2[i] (C2[i] may be NULL) is
// accumulated into with atomic XORs, so destinations may repeat across the products (the caller
// zeroes them first when it wants C = instead of C ^=).
void k_mul_abt_mma6_group(int count, DMat *const *C1, DMat *const *C2, const DMat *const *A, const DMat *const *Bt,
                          int cg, cudaStream_t s) {
  if (count < 1 || count > GROUP_MAX) die("m4ri_b200: a product group holds 1 to %d products\n", GROUP_MAX);
  if (C1[0]->nrows == 0 || C1[0]->ncols == 0 || A[0]->ncols == 0) return;
  if (!has_tcgen05()) die("m4ri_b200: the tcgen05 MMA path needs an sm_100 device\n");
  Mma6Job jobs[GROUP_MAX];
  for (int i = 0; i < count; i++) jobs[i] = {C1[i], C2[i], A[i], Bt[i], 0, 0};
  dispatch_mma6(jobs, count, 1, cg, s);
}

}  // namespace m4b
