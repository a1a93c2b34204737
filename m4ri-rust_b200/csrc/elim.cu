// elim.cu — the consumers of multiplication: reduced row echelon form, inverse and
// linear solve on the GPU, behind the reference's names
//   mzd_echelonize / mzd_echelonize_pluq / mzd_echelonize_m4ri   m4ri-sys/src/echelonform.rs:16,27,36
//   mzd_inv_m4ri                                                  m4ri-sys/src/brilliantrussian.rs:208
//   mzd_solve_left                                                m4ri-sys/src/solve.rs:23
// (call sites in m4ri-rust: BinMatrix::echelonize / rank / inverted and solve_left,
// m4ri-rust/src/friendly/binary_matrix.rs:247-268,575-586).
//
// Algorithm (B200 design, not M4RI's PLE recursion): blocked Gauss-Jordan whose bulk work is ONE
// accumulating GF(2) product per column block on the tensor path — the same "trailing update is
// a matrix product" structure that makes M4RI's PLE fast, with the product being
// m4ri_b200's dmul(clear = 0).
//
// For a block of B columns [c0, c0 + B):
//   1. The block's columns of ALL rows are copied into a panel PT = [P | T], T = 0 ("tags").
//   2. 64 columns at a time (one word column j of P): one CTA finds up to 64 pivots among the
//      candidate rows (rows that are no pivot yet) by eliminating on that single word column,
//      and reduces the pivot words among themselves; the augmented pivot rows are combined into
//      R' (identity on the pivot columns) and every other row of the panel gets
//      row ^= sum_k bit(row, pivot column k) * R'_k — a rank-64 update confined to the panel.
//      A pivot row's tag gets its own selection index, so after the block T holds, for every
//      row, the combination of the block's ORIGINAL pivot rows that the elimination added to it.
//   3. Everything right of the block is updated with one product:
//         W[:, c0+B:] ^= T (m x B) * OldPivotRows[:, c0+B:] (B x rest)        (tensor path)
//      (columns left of the block need nothing: candidate rows are zero there).
//   4. The panel is written back and the pivot rows are moved up to rows rank .. rank+kb-1 in
//      pivot-column order.
// The result is THE reduced row echelon form (unique), so parity with any other implementation
// is bit-exact by construction; rank and pivot columns come with it.
#include <algorithm>
#include <vector>

#include "common.cuh"

namespace m4b {
namespace {

constexpr int EB_DEFAULT = 1024;  // block width in columns (multiple of 128)
constexpr int EB_MAX = 4096;

struct ElimState {
  int kp;                        // pivots of the current 64-column step
  int kb;                        // pivots of the block before the current step
  int kbn;                       // ... including the current step
  int pr[64];                    // their rows
  int pc[64];                    // their bit positions in the word column
  unsigned long long ctag[64];   // row k of the inverse of the pivot words on the pivot columns
  int blk_rows[EB_MAX];          // pivots of the block in selection order: row, column
  int blk_cols[EB_MAX];
  // PLE only (lu_pivot_search_kernel): the step's pivots by COLUMN of the word
  unsigned long long pivmask;    // columns that hold a pivot
  word pwk[64];                  // pivot k's word j as the pivot row keeps it: [L | 1 | E]
  word pwcol[64];                // the same words by pivot column
  int kcol[64];                  // selection index by pivot column
};

#ifdef M4RI_B200_PERF_HOOKS
__device__ long long g_clk[16];
__device__ long long g_it[64];
#define CLK(i) do { if (threadIdx.x == 0 && c0 == 0 && j == 3) g_clk[i] = clock64(); } while (0)
#else
#define CLK(i) do {} while (0)
#endif

// State that lives across the blocks of one elimination (the per-block ElimState is cleared at
// every block): nothing of it is needed on the host before the very end, so the whole
// elimination is enqueued without a host round trip.
struct ElimGlobal {
  int rank;    // pivots found in the blocks done so far = first candidate row of the next block
  int moves;   // rows the current block's permutation moves
  int max_id;  // inversion: largest ORIGINAL index of a row that has served as a pivot row
};

static inline unsigned cdiv(long long a, long long b) { return (unsigned)((a + b - 1) / b); }

// The 64-column steps are chains of small dependent kernels (search -> prepare -> apply, ~1 500
// launches per elimination): they are launched with programmatic stream serialization, so that a
// kernel's launch overlaps its predecessor's tail. Every kernel of the chain starts with
// PDL_WAIT() — it touches nothing before the kernels ahead of it have completed and flushed — and
// lets its successor be scheduled early with PDL_GO().
#define PDL_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
#define PDL_GO() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")
template <typename... KArgs, typename... Args>
static void launch_chained(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t s, Args... args) {
  static const bool off = getenv("M4RI_B200_ELIM_NO_PDL") != nullptr;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = off ? 0 : 1;
  M4B_CUDA(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
  M4B_LAUNCHED();
}

// PT[i] = [ W[i][cw0 .. cw0+bw) | 0 ... 0 ]  (words past the matrix width read as zero)
__global__ void extract_panel_kernel(word *PT, int pstride, const word *W, int ws, int wwidth, int m, int cw0,
                                     int bw) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)m * 2 * bw) return;
  const int i = (int)(idx / (2 * bw)), w = (int)(idx % (2 * bw));
  word v = 0;
  if (w < bw && cw0 + w < wwidth) v = W[(size_t)i * ws + cw0 + w];
  PT[(size_t)i * pstride + w] = v;
}

__global__ void writeback_panel_kernel(word *W, int ws, int wwidth, const word *PT, int pstride, int m, int cw0,
                                       int bw) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)m * bw) return;
  const int i = (int)(idx / bw), w = (int)(idx % bw);
  if (cw0 + w < wwidth) W[(size_t)i * ws + cw0 + w] = PT[(size_t)i * pstride + w];
}

// One CTA: pivots of word column j of the panel among the candidate rows (row >= r0, not yet a
// pivot of this block), 1024 rows at a time.
//   all threads  forward-eliminate their row's word with the pivots found so far (tracking which
//                pivots they absorbed) and compact the rows that are still non-zero;
//   warp 0       takes the compacted rows 32 at a time and extracts pivots with ballot/shuffle
//                (first non-zero lane, its lowest set bit). The pivots live in warp 0's
//                registers (lane l holds pivots l and l + 32) and every new pivot is cleared
//                from the earlier ones at once, so the set is always fully reduced — no
//                shared-memory round trip per pivot and no reduction pass afterwards.
//                Stops when all 64 columns have a pivot: for a random matrix after ~70 rows.
// ctag[k] = the combination of the selected rows (selection order) that gives the k-th reduced
// pivot word.
__global__ void __launch_bounds__(1024)
pivot_search_kernel(const word *PT, int pstride, int j, int m, const ElimGlobal *gl, unsigned char *ispiv, int *pidx,
                    ElimState *st, int c0, word colmask /* valid columns of this word (pivot limit) */) {
  PDL_WAIT();
  PDL_GO();
  const int r0 = gl->rank;
  __shared__ word pw[64], sw[1024];
  __shared__ unsigned long long ctag[64], slt[1024];
  __shared__ int pc[64], pr[64], srow[1024];
  __shared__ int kp_s, cnt;
  const int tid = threadIdx.x, lane = tid & 31;
  CLK(0);
  if (tid < 64 && tid < st->kp) pidx[st->pr[tid]] = 0;  // the previous step's pivot marks
  __syncthreads();
  if (tid == 0) {
    kp_s = 0;
    cnt = 0;
    st->kb = st->kbn;  // the previous step's pivots are counted from now on
  }
  // warp 0: pivot `lane` in the A registers, pivot `lane + 32` in the B registers
  word pwA = 0, pwB = 0;
  unsigned long long ctA = 0, ctB = 0;
  int pcA = 0, pcB = 0, prA = -1, prB = -1, kp = 0;
  __syncthreads();
  CLK(1);
  for (int chunk0 = r0; chunk0 < m; chunk0 += 1024) {
    const int kp0 = kp_s;
    if (kp0 == 64) break;
    const int row = chunk0 + tid;
    word w = 0;
    if (row < m) {
      const unsigned char taken = ispiv[row];
      w = PT[(size_t)row * pstride + j] & colmask;
      if (taken) w = 0;
    }
    unsigned long long lt = 0;
    for (int k = 0; k < kp0; k++) {  // branch-free: the shared-memory reads pipeline
      const word sel = (word)0 - ((w >> pc[k]) & 1);
      w ^= pw[k] & sel;
      lt ^= ctag[k] & sel;
    }
    if (chunk0 == r0) CLK(2);
    const unsigned nz = __ballot_sync(0xffffffffu, w != 0);
    int base = 0;
    if (lane == 0 && nz) base = atomicAdd(&cnt, __popc(nz));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (w) {
      const int e = base + __popc(nz & ((1u << lane) - 1));
      sw[e] = w;
      slt[e] = lt;
      srow[e] = row;
    }
    __syncthreads();
    if (chunk0 == r0) CLK(3);
    if (tid < 32) {
      const int n = cnt;
#ifdef M4RI_B200_PERF_HOOKS
      long long t_fwd = 0, t_ext = 0; int n_groups = 0;
#endif
      for (int g = 0; g < n && kp < 64; g += 32) {
#ifdef M4RI_B200_PERF_HOOKS
        long long t0 = clock64(); n_groups++;
#endif
        // (loads unconditional and the warp explicitly converged: the full-mask ballots and
        // shuffles below cost ~4x more when the compiler leaves the lanes past n diverged)
        const int e = g + lane, ec = e < n ? e : 0;
        word v = sw[ec];
        unsigned long long vt = slt[ec];
        const int vrow = srow[ec];
        if (e >= n) { v = 0; vt = 0; }
        __syncwarp();
        const int kp_g = kp;
        for (int k = kp0; k < kp_g; k++) {  // pivots found earlier in this chunk (published below)
          const word sel = (word)0 - ((v >> pc[k]) & 1);
          v ^= pw[k] & sel;
          vt ^= ctag[k] & sel;
        }
#ifdef M4RI_B200_PERF_HOOKS
        long long t1 = clock64(); t_fwd += t1 - t0;
#endif
        for (;;) {
#ifdef M4RI_B200_PERF_HOOKS
          if (kp < 8 && tid == 0 && c0 == 0 && j == 3) g_it[kp * 4] = clock64();
#endif
          const unsigned ball = __ballot_sync(0xffffffffu, v != 0);
          if (!ball || kp == 64) break;
#ifdef M4RI_B200_PERF_HOOKS
          if (kp < 8 && tid == 0 && c0 == 0 && j == 3) g_it[kp * 4 + 1] = clock64();
#endif
          const int src = __ffs(ball) - 1;
          const word pv = __shfl_sync(0xffffffffu, v, src);
          const unsigned long long pt = __shfl_sync(0xffffffffu, vt, src) ^ (1ull << kp);
          const int prow = __shfl_sync(0xffffffffu, vrow, src);
          const int tp = __ffsll((long long)pv) - 1;
#ifdef M4RI_B200_PERF_HOOKS
          if (kp < 8 && tid == 0 && c0 == 0 && j == 3) g_it[kp * 4 + 2] = clock64();
#endif
          if ((pwA >> tp) & 1) { pwA ^= pv; ctA ^= pt; }  // keep the earlier pivots reduced
          if ((pwB >> tp) & 1) { pwB ^= pv; ctB ^= pt; }
          {
            const word sel = lane == src ? 0 : (word)0 - ((v >> tp) & 1);
            v = lane == src ? 0 : v ^ (pv & sel);
            vt ^= pt & sel;
          }
          if (lane == (kp & 31)) {
            if (kp < 32) { pwA = pv; ctA = pt; pcA = tp; prA = prow; }
            else         { pwB = pv; ctB = pt; pcB = tp; prB = prow; }
          }
          kp++;
        }
#ifdef M4RI_B200_PERF_HOOKS
        t_ext += clock64() - t1;
#endif
        // mark this group's pivots; publish the (reduced) set for the next group, the next
        // chunk and the output
        if (lane >= kp_g && lane < kp) { ispiv[prA] = 1; pidx[prA] = lane + 1; }
        if (lane + 32 >= kp_g && lane + 32 < kp) { ispiv[prB] = 1; pidx[prB] = lane + 33; }
        pw[lane] = pwA; pw[lane + 32] = pwB;
        ctag[lane] = ctA; ctag[lane + 32] = ctB;
        pc[lane] = pcA; pc[lane + 32] = pcB;
        pr[lane] = prA; pr[lane + 32] = prB;
        __syncwarp();
      }
      if (lane == 0) {
        kp_s = kp;
        cnt = 0;
      }
      if (chunk0 == r0) CLK(4);
#ifdef M4RI_B200_PERF_HOOKS
      if (chunk0 == r0 && tid == 0 && c0 == 0 && j == 3) { g_clk[8] = n; g_clk[9] = n_groups; g_clk[10] = t_fwd; g_clk[11] = t_ext; g_clk[12] = kp; }
#endif
    }
    __syncthreads();
  }
  CLK(5);
  const int n = kp_s;
  if (tid < 64) {
    const int kb = st->kbn;  // == the new kb (set above by thread 0; kbn itself is rewritten below)
    st->pr[tid] = tid < n ? pr[tid] : -1;
    st->pc[tid] = tid < n ? pc[tid] : 0;
    st->ctag[tid] = tid < n ? ctag[tid] : 0;
    if (tid < n) {
      st->blk_rows[kb + tid] = pr[tid];
      st->blk_cols[kb + tid] = c0 + 64 * j + pc[tid];
    }
  }
  __syncthreads();
  if (tid == 0) {
    st->kp = n;
    st->kbn = st->kbn + n;
  }
  CLK(6);
}

// Step preparation, one launch:
//   CTAs [0, nrc)   4 words w of [j, L) per CTA: R'[k][w] = XOR over l in ctag[k] of
//                   PT[pr[l]][w], plus the tag bits kb + l (the pivot rows' words are staged in
//                   shared memory), then the 16 four-bit tables of the step
//                       Tab[g][v][w] = XOR over bits b of v of R'[4g + b][w]
//   CTAs [nrc, ..)  mult[i] bit k = bit pc[k] of the OLD word j of row i, for every row
constexpr int RP_XW = 4;
// The table part of a step, one CTA of 256 threads per RP_XW words (shared by both flavours).
__device__ __forceinline__ void step_tables(word *Tab, int L, const word *PT, int pstride, int j, int bw,
                                            const ElimState *st, word *srows, word *srp) {
  const int kp = st->kp, t = threadIdx.x;
  const int k = t / RP_XW, xx = t % RP_XW, w = j + blockIdx.x * RP_XW + xx;
  const bool live = w < L;
  if (live && k < kp) srows[t] = PT[(size_t)st->pr[k] * pstride + w];
  __syncthreads();
  word acc = 0;
  if (live && k < kp) {
    const unsigned long long c_all = st->ctag[k];
    unsigned long long c = c_all;
    while (c) {
      const int l = __ffsll((long long)c) - 1;
      c &= c - 1;
      acc ^= srows[l * RP_XW + xx];
    }
    if (w >= bw) {  // tag part: selected row l carries tag bit kb + l
      const int kb = st->kb, sh = kb & 63, tw = w - bw;
      if (tw == (kb >> 6)) acc ^= c_all << sh;
      else if (sh && tw == (kb >> 6) + 1) acc ^= c_all >> (64 - sh);
    }
  }
  srp[t] = acc;
  __syncthreads();
  if (!live) return;
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const int e = k + 64 * q, g = e >> 4, v = e & 15;  // table g, entry v
    const word *r4 = srp + (4 * g) * RP_XW + xx;
    word o = 0;
    if (v & 1) o ^= r4[0];
    if (v & 2) o ^= r4[RP_XW];
    if (v & 4) o ^= r4[2 * RP_XW];
    if (v & 8) o ^= r4[3 * RP_XW];
    Tab[(size_t)e * L + w] = o;
  }
}

__global__ void __launch_bounds__(256)
prepare_step_kernel(word *Tab, unsigned long long *mult, int L, const word *PT, int pstride, int j, int bw, int m,
                    int nrc, const ElimState *st) {
  PDL_WAIT();
  PDL_GO();
  __shared__ word srows[64 * RP_XW], srp[64 * RP_XW];
  __shared__ int spc[64];
  const int kp = st->kp, t = threadIdx.x;
  if ((int)blockIdx.x >= nrc) {
    if (t < 64) spc[t] = st->pc[t];
    __syncthreads();
    const int i = (blockIdx.x - nrc) * 256 + t;
    if (i >= m) return;
    const word w = PT[(size_t)i * pstride + j];
    unsigned long long mm = 0;
    for (int k = 0; k < kp; k++) mm |= ((w >> spc[k]) & 1ull) << k;
    mult[i] = mm;
    return;
  }
  step_tables(Tab, L, PT, pstride, j, bw, st, srows, srp);
}

// Every (row, word): PT[i][w] ^= sum over k in mult[i] of R'[k][w], w in [j, L), as 16 table
// look-ups (4 multiplier bits each); the pivot rows of the step become R'. One thread per word;
// the tables (256 x L words, 64 KB at the default block width) are read through L1 and all 16
// loads are in flight together.
__global__ void __launch_bounds__(256)
apply_rprime_kernel(word *PT, int pstride, int j, int L, int m, const int *pidx, const unsigned long long *mult,
                    const word *__restrict__ Tab, const ElimState *st) {
  PDL_WAIT();
  PDL_GO();
  if (st->kp == 0) return;
  const int LW = L - j;
  const long long idx = blockIdx.x * 256LL + threadIdx.x;
  if (idx >= (long long)m * LW) return;
  const int i = (int)(idx / LW), w = j + (int)(idx % LW);
  const int pk = pidx[i];
  const unsigned long long mm = mult[i];
  word *dst = PT + (size_t)i * pstride + w;
  if (pk) {  // pivot pk - 1 of this step (the mark is cleared by the next step's search)
    *dst = Tab[(size_t)(((pk - 1) >> 2) * 16 + (1 << ((pk - 1) & 3))) * L + w];
    return;
  }
  if (!mm) return;
  const word *tp = Tab + w;
  word acc = 0;
#pragma unroll
  for (int g = 0; g < 16; g++)
    acc ^= __ldg(tp + (size_t)(g * 16 + (int)((mm >> (4 * g)) & 15)) * L);
  *dst ^= acc;
}

// The same with two words per thread (16-byte table look-ups and a 16-byte read-modify-write of
// the row): pairs start at the even word at or below j; a pair's first word is skipped when it
// lies left of j.
__global__ void __launch_bounds__(256)
apply_rprime2_kernel(word *PT, int pstride, int j, int L, int m, const int *pidx, const unsigned long long *mult,
                     const word *__restrict__ Tab, const ElimState *st) {
  PDL_WAIT();
  PDL_GO();
  if (st->kp == 0) return;
  const int j0 = j & ~1, LP = (L - j0) / 2;
  const long long idx = blockIdx.x * 256LL + threadIdx.x;
  if (idx >= (long long)m * LP) return;
  const int i = (int)(idx / LP), w = j0 + 2 * (int)(idx % LP);
  const int pk = pidx[i];
  const unsigned long long mm = mult[i];
  ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(PT + (size_t)i * pstride + w);
  const bool first = w >= j;  // (w + 1 >= j always)
  if (pk) {
    const ulonglong2 t = *reinterpret_cast<const ulonglong2 *>(Tab + (size_t)(((pk - 1) >> 2) * 16 + (1 << ((pk - 1) & 3))) * L + w);
    if (first) *dst = t;
    else reinterpret_cast<word *>(dst)[1] = t.y;
    return;
  }
  if (!mm) return;
  const word *tp = Tab + w;
  ulonglong2 acc = make_ulonglong2(0, 0);
#pragma unroll
  for (int g = 0; g < 16; g++) {
    const ulonglong2 t = __ldg(reinterpret_cast<const ulonglong2 *>(tp + (size_t)(g * 16 + (int)((mm >> (4 * g)) & 15)) * L));
    acc.x ^= t.x;
    acc.y ^= t.y;
  }
  if (first) {
    ulonglong2 o = *dst;
    o.x ^= acc.x;
    o.y ^= acc.y;
    *dst = o;
  } else {
    reinterpret_cast<word *>(dst)[1] ^= acc.y;
  }
}

// OP[t][:] = W[blk_rows[t]][tw0 .. tw0+tw) for t < kb (other rows of OP stay zero);
// and T[blk_rows[t]] ^= e_t: a pivot row is REPLACED by its combination, i.e. it also loses itself
__global__ void gather_pivot_rows_kernel(word *OP, int ostride, const word *W, int ws, int tw0, int tw,
                                         word *PT, int pstride, int bw, const ElimState *st) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int kb = st->kbn;  // (the grid is sized for a full block of pivots)
  if (idx >= (long long)kb * (tw + 1)) return;
  const int t = (int)(idx / (tw + 1)), w = (int)(idx % (tw + 1));
  const int row = st->blk_rows[t];
  if (w < tw) OP[(size_t)t * ostride + w] = W[(size_t)row * ws + tw0 + w];
  else PT[(size_t)row * pstride + bw + (t >> 6)] ^= (word)1 << (t & 63);
}

// End of a block, one CTA: where every row goes. The block's pivot rows move up to positions
// rank .. rank+kb-1 in pivot-column order; the rows they displace there (which are no pivots of
// this block) take the positions the pivot rows vacate outside that range — which displaced row
// goes to which free position does not matter (non-pivot rows are only candidates of later
// blocks or end up zero). Writes the move list (src -> dst, through a temporary), the pivot
// columns in order, and advances rank.
__global__ void __launch_bounds__(1024)
plan_permutation_kernel(const ElimState *st, ElimGlobal *gl, int *src, int *dst, int *piv_all) {
  __shared__ int s_cols[EB_MAX];
  __shared__ unsigned char s_flag[EB_MAX];
  __shared__ int s_warp[32], s_total;
  const int kb = st->kbn, rank = gl->rank, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int t = tid; t < kb; t += 1024) {
    s_cols[t] = st->blk_cols[t];
    s_flag[t] = 0;
  }
  __syncthreads();
  for (int t = tid; t < kb; t += 1024) {
    const int c = s_cols[t], r = st->blk_rows[t];
    int pos = 0;
    for (int u = 0; u < kb; u++) pos += s_cols[u] < c;  // pivot columns are distinct
    src[t] = r;
    dst[t] = rank + pos;
    piv_all[rank + pos] = c;
    if (r >= rank && r < rank + kb) s_flag[r - rank] = 1;  // this target position holds a pivot row
  }
  __syncthreads();
  // two compactions over [0, kb) with one block-wide exclusive scan each: pass 0 lists the
  // displaced rows (positions of the target range without a pivot row), pass 1 the vacated
  // positions (pivot rows outside the target range)
  const int per = (kb + 1023) / 1024, i0 = tid * per, i1 = min(kb, i0 + per);
  for (int pass = 0; pass < 2; pass++) {
    int mine = 0;
    for (int i = i0; i < i1; i++) {
      const int r = pass ? st->blk_rows[i] : 0;
      mine += pass ? !(r >= rank && r < rank + kb) : !s_flag[i];
    }
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = s_warp[lane], wi = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += v;
      }
      s_warp[lane] = wi - w;
      if (lane == 31) s_total = wi;
    }
    __syncthreads();
    int at = kb + s_warp[warp] + incl - mine;
    for (int i = i0; i < i1; i++) {
      if (pass == 0) {
        if (!s_flag[i]) src[at++] = rank + i;
      } else {
        const int r = st->blk_rows[i];
        if (!(r >= rank && r < rank + kb)) dst[at++] = r;
      }
    }
    __syncthreads();
  }
  if (tid == 0) {
    gl->moves = kb + s_total;  // (both passes count the same number of rows)
    gl->rank = rank + kb;
  }
}

// rows dst[q] of W := rows src[q] (through a compact temporary)
// (rowid, PLE only: which original row sits at every position — moved with the rows)
__global__ void gather_rows_kernel(word *tmp, const word *W, int ws, int width, const int *src, const ElimGlobal *gl,
                                   int *tmpid, const int *rowid) {
  const int cnt = gl->moves;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)cnt * width) return;
  const int q = (int)(idx / width), w = (int)(idx % width);
  tmp[idx] = W[(size_t)src[q] * ws + w];
  if (rowid && w == 0) tmpid[q] = rowid[src[q]];
}
__global__ void scatter_rows_kernel(word *W, int ws, int width, const word *tmp, const int *dst, const ElimGlobal *gl,
                                    const int *tmpid, int *rowid) {
  const int cnt = gl->moves;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)cnt * width) return;
  const int q = (int)(idx / width), w = (int)(idx % width);
  W[(size_t)dst[q] * ws + w] = tmp[idx];
  if (rowid && w == 0) rowid[dst[q]] = tmpid[q];
}

// ---- PLE (mzd_ple / mzd_pluq, m4ri-sys/src/ple.rs:35,65): the LU flavour of a 64-column step ----
// Columns are taken in increasing order; a row is eliminated only on the bits ABOVE a pivot
// column and keeps the bit at the pivot column itself: that bit is its multiplier (the entry of
// L). A row "stops" at the first non-pivot column where it still has a bit — there it can become
// the pivot of that column.
__device__ __forceinline__ word bits_above(int b) { return b >= 63 ? (word)0 : ~(word)0 << (b + 1); }

// Continue the column scan of one row's word from bit `from` against the pivots published in
// shared memory: returns the stop column (64: the row depends on the pivots).
__device__ __forceinline__ int lu_scan(word &w, unsigned long long &lt, int from, word pivmask, word colmask,
                                       const word *pwcol, const unsigned long long *ctcol) {
  word consider = from >= 64 ? (word)0 : ~(word)0 << from;
  for (;;) {
    const word t = w & consider & colmask;
    if (!t) return 64;
    const int b = __ffsll((long long)t) - 1;
    if (!((pivmask >> b) & 1)) return b;
    w ^= pwcol[b] & bits_above(b);
    lt ^= ctcol[b];
    consider = bits_above(b);
  }
}

// One CTA: the pivots of word column j for the PLE. 1024 candidate rows at a time are scanned
// against the pivots found so far and the rows that stop somewhere are compacted; warp 0 takes
// them 32 at a time: the smallest stop column of the group gets its pivot (first lane that stops
// there), the other lanes stopping there absorb it and scan on. A pivot found later at a smaller
// column never touches the earlier pivots: every row scanned before had a zero there.
// ctag[k] = the combination of the ORIGINAL selected rows that makes pivot row k (row k of the
// inverse of the step's unit lower triangular L, in selection indices).
__global__ void __launch_bounds__(1024)
lu_pivot_search_kernel(const word *PT, int pstride, int j, int m, const ElimGlobal *gl, unsigned char *ispiv, int *pidx,
                       ElimState *st, int c0, word colmask) {
  PDL_WAIT();
  PDL_GO();
  const int r0 = gl->rank;
  __shared__ word pwcol[64], pwk[64], sw[1024];
  __shared__ unsigned long long ctcol[64], ctk[64], slt[1024];
  __shared__ int pc[64], pr[64], kcol[64], srow[1024], sstop[1024];
  __shared__ word pmask_s;
  __shared__ int kp_s, cnt, wcnt[32];
  const int tid = threadIdx.x, lane = tid & 31;
  if (tid < 64 && tid < st->kp) pidx[st->pr[tid]] = 0;  // the previous step's pivot marks
  __syncthreads();
  if (tid == 0) {
    kp_s = 0;
    cnt = 0;
    pmask_s = 0;
    st->kb = st->kbn;
  }
  if (tid < 64) {
    pwcol[tid] = 0;
    ctcol[tid] = 0;
    kcol[tid] = 0;
  }
  int kp = 0;  // warp 0's count
  __syncthreads();
  for (int chunk0 = r0; chunk0 < m; chunk0 += 1024) {
    const word pm = pmask_s;
    if ((pm & colmask) == colmask) break;  // every column of the word has its pivot
    const int row = chunk0 + tid;
    word w = 0;
    if (row < m && !ispiv[row]) w = PT[(size_t)row * pstride + j];
    unsigned long long lt = 0;
    const int stop = lu_scan(w, lt, 0, pm, colmask, pwcol, ctcol);
    // compaction in row order (deterministic pivots): per-warp counts, then a prefix over them
    const unsigned nz = __ballot_sync(0xffffffffu, stop < 64);
    if (lane == 0) wcnt[tid >> 5] = __popc(nz);
    __syncthreads();
    const int wc = wcnt[lane];
    const int base = __reduce_add_sync(0xffffffffu, lane < (tid >> 5) ? wc : 0);
    if (tid < 32) {
      const int total = __reduce_add_sync(0xffffffffu, wc);
      if (lane == 0) cnt = total;
    }
    if (stop < 64) {
      const int e = base + __popc(nz & ((1u << lane) - 1));
      sw[e] = w;
      slt[e] = lt;
      srow[e] = row;
      sstop[e] = stop;
    }
    __syncthreads();
    if (tid < 32) {
      const int n = cnt;
      word pmask = pm;
      for (int g = 0; g < n; g += 32) {
        if ((pmask & colmask) == colmask) break;
        const int e = g + lane, ec = e < n ? e : 0;
        word v = sw[ec];
        unsigned long long vt = slt[ec];
        const int vrow = srow[ec];
        int vstop = sstop[ec], mine = -1;
        if (e >= n) { v = 0; vt = 0; vstop = 64; }
        __syncwarp();
        if (vstop < 64) vstop = lu_scan(v, vt, vstop, pmask, colmask, pwcol, ctcol);  // pivots of earlier groups
        __syncwarp();
        // The group's pivots stay in registers (lane l holds the l-th pivot of the group) and are
        // published once, at the end: inside the group the smallest stop column comes first, so a
        // lane that absorbs a pivot only scans columns ABOVE every pivot of its own group — the
        // look-ups of lu_scan only ever hit pivots of earlier groups, which are in shared memory.
        const int kp_g = kp;
        word myp = 0;
        unsigned long long myt = 0;
        int myr = -1, myc = 0;
        for (;;) {
          const int cmin = __reduce_min_sync(0xffffffffu, vstop);
          if (cmin >= 64) break;
          const int src = __ffs(__ballot_sync(0xffffffffu, vstop == cmin)) - 1;
          const word pv = __shfl_sync(0xffffffffu, v, src);
          const unsigned long long pt = __shfl_sync(0xffffffffu, vt, src) ^ (1ull << kp);
          const int prow = __shfl_sync(0xffffffffu, vrow, src);
          if (lane == kp - kp_g) { myp = pv; myt = pt; myr = prow; myc = cmin; }
          pmask |= (word)1 << cmin;
          if (lane == src) {
            mine = kp;
            vstop = 64;
          } else if (vstop == cmin) {
            v ^= pv & bits_above(cmin);
            vt ^= pt;
            vstop = lu_scan(v, vt, cmin + 1, pmask, colmask, pwcol, ctcol);
          }
          kp++;
        }
        if (lane < kp - kp_g) {
          const int k = kp_g + lane;
          pwcol[myc] = myp; ctcol[myc] = myt; kcol[myc] = k;
          pwk[k] = myp; ctk[k] = myt; pr[k] = myr; pc[k] = myc;
        }
        __syncwarp();
        if (mine >= 0) { ispiv[vrow] = 1; pidx[vrow] = mine + 1; }
      }
      if (lane == 0) {
        pmask_s = pmask;
        kp_s = kp;
        cnt = 0;
      }
    }
    __syncthreads();
  }
  const int n = kp_s;
  if (tid < 64) {
    const int kb = st->kbn;
    st->pr[tid] = tid < n ? pr[tid] : -1;
    st->pc[tid] = tid < n ? pc[tid] : 0;
    st->ctag[tid] = tid < n ? ctk[tid] : 0;
    st->pwk[tid] = tid < n ? pwk[tid] : 0;
    st->pwcol[tid] = pwcol[tid];
    st->kcol[tid] = kcol[tid];
    if (tid < n) {
      st->blk_rows[kb + tid] = pr[tid];
      st->blk_cols[kb + tid] = c0 + 64 * j + pc[tid];
    }
  }
  __syncthreads();
  if (tid == 0) {
    st->pivmask = pmask_s;
    st->kp = n;
    st->kbn = st->kbn + n;
  }
}

// Word j of every row after the step, and the multipliers of the rank-64 update of the words
// right of it: candidate rows are scanned against the complete pivot set (the bits left at the
// pivot columns are the entries of L, mult = the same bits in selection indices), the step's
// pivot rows get their [L | 1 | E] word, rows above the block's first candidate and the pivots of
// earlier steps are left alone (mult = 0).
// (one launch with the step's tables: CTAs [0, nrc) build the tables of the words right of j —
// they read words > j of the pivot rows, the row part writes word j only)
__global__ void __launch_bounds__(256)
lu_prepare_kernel(word *Tab, int L, int bw, int nrc, word *PT, int pstride, int j, int m, const ElimGlobal *gl,
                  const unsigned char *ispiv, const int *pidx, unsigned long long *mult, const ElimState *st) {
  PDL_WAIT();
  PDL_GO();
  __shared__ word srows[64 * RP_XW], srp[64 * RP_XW];
  if ((int)blockIdx.x < nrc) {
    step_tables(Tab, L, PT, pstride, j + 1, bw, st, srows, srp);
    return;
  }
  word *pwcol = srows;                // (the row part reuses the table part's shared memory)
  int *kcol = (int *)srp;
  const int t = threadIdx.x;
  if (t < 64) {
    pwcol[t] = st->pwcol[t];
    kcol[t] = st->kcol[t];
  }
  __syncthreads();
  const int i = (blockIdx.x - nrc) * 256 + t;
  if (i >= m) return;
  const int pk = pidx[i];
  if (pk) {
    PT[(size_t)i * pstride + j] = st->pwk[pk - 1];
    mult[i] = 0;
    return;
  }
  if (i < gl->rank || ispiv[i] || st->kp == 0) {
    mult[i] = 0;
    return;
  }
  const word pivmask = st->pivmask;
  word w = PT[(size_t)i * pstride + j], consider = ~(word)0;
  unsigned long long mm = 0;
  for (;;) {
    const word tt = w & consider & pivmask;
    if (!tt) break;
    const int b = __ffsll((long long)tt) - 1;
    w ^= pwcol[b] & bits_above(b);
    mm |= 1ull << kcol[b];
    consider = bits_above(b);
  }
  PT[(size_t)i * pstride + j] = w;
  mult[i] = mm;
}

// ---- PLE, the end: L compressed into the first `rank` columns ---------------------------------
// Lc[i] bit k = W[i][Q[k]] for k < min(i, rank): row i's multiplier for pivot k sits at pivot
// column Q[k] during the elimination; the documented storage (ple.rs:44-51, "L and S stored in
// place over A") has column k of L in column k. pm[w] = the pivot columns of word w as a mask,
// pcnt[w] = pivots left of word w: a destination word walks the set bits of the masks from the
// word that holds pivot 64*dw on — shifts and masks on a few source words, no load per bit.
__global__ void pivot_masks_kernel(word *pm, const int *Q, int rank) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < rank) atomicOr((unsigned long long *)&pm[Q[k] >> 6], 1ull << (Q[k] & 63));
}
__global__ void pivot_prefix_kernel(int *pcnt, const word *pm, int width) {  // width is a few hundred words
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= width) return;
  int acc = 0;
  for (int u = 0; u < w; u++) acc += __popcll(pm[u]);
  pcnt[w] = acc;
}
__global__ void gather_l_kernel(word *Lc, int lcs, int lcw, const word *W, int ws, int width, int m, int rank,
                                const int *Q, const word *pm, const int *pcnt) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)m * lcw) return;
  const int i = (int)(idx / lcw), dw = (int)(idx % lcw);
  const int kmax = min(i, rank), k0 = 64 * dw;
  word o = 0;
  if (k0 < kmax) {
    const word *row = W + (size_t)i * ws;
    const int want = min(64, kmax - k0);
    int w = Q[k0] >> 6, filled = 0;
    int skip = k0 - pcnt[w];  // pivots of word w that belong to earlier destination words
    while (filled < want && w < width) {
      word mask = pm[w];
      const word v = row[w];
      for (; skip > 0; skip--) mask &= mask - 1;
      while (mask && filled < want) {
        const int b = __ffsll((long long)mask) - 1;
        o |= ((v >> b) & 1) << filled;
        filled++;
        mask &= mask - 1;
      }
      w++;
    }
  }
  Lc[(size_t)i * lcs + dw] = o;
}
// Rows below the rank hold nothing but multipliers; a pivot row keeps its columns from Q[i] on.
__global__ void strip_l_kernel(word *W, int ws, int width, int m, int rank, const int *Q) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)m * width) return;
  const int i = (int)(idx / width), w = (int)(idx % width);
  word *p = W + (size_t)i * ws + w;
  if (i >= rank) { *p = 0; return; }
  const int c = Q[i];
  if (w < (c >> 6)) *p = 0;
  else if (w == (c >> 6)) *p &= ~(word)0 << (c & 63);
}
// out[i][c] = in[i][colsrc[c]] for the first `rows` rows (PLUQ: U = E with its columns permuted)
__global__ void permute_cols_kernel(word *out, int os, const word *in, int is, int rows, int width, int ncols,
                                    const int *colsrc) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)rows * width) return;
  const int i = (int)(idx / width), w = (int)(idx % width);
  const word *row = in + (size_t)i * is;
  word o = 0;
  for (int b = 0; b < 64; b++) {
    const int c = 64 * w + b;
    if (c >= ncols) break;
    const int sc = colsrc[c];
    o |= ((row[sc >> 6] >> (sc & 63)) & 1) << b;
  }
  out[(size_t)i * os + w] = o;
}
__global__ void merge_l_kernel(word *W, int ws, const word *Lc, int lcs, int lcw, int m) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)m * lcw) return;
  const int i = (int)(idx / lcw), w = (int)(idx % lcw);
  W[(size_t)i * ws + w] |= Lc[(size_t)i * lcs + w];
}
// out[i] = in[src[i]] (mzd_apply_p_left*: the composed row permutation in one pass)
__global__ void permute_rows_kernel(word *out, int os, const word *in, int is, int rows, int width, const int *src) {
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)rows * width) return;
  const int i = (int)(idx / width), w = (int)(idx % width);
  out[(size_t)i * os + w] = in[(size_t)src[i] * is + w];
}
// dst (r x r, zeroed) = the unit lower (lower != 0) or the upper triangle of the packed factors:
// bits c < i plus the diagonal, or bits c >= i, of the first r columns of row i
__global__ void tri_extract_kernel(word *dst, int ds, const word *src, int ss, int r, int lower) {
  const int rw = (r + 63) / 64;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= (long long)r * rw) return;
  const int i = (int)(idx / rw), w = (int)(idx % rw);
  word v = src[(size_t)i * ss + w];
  const int lo = 64 * w;                       // first column of this word
  word below = 0;                              // mask of the columns c < i in this word
  if (i >= lo + 64) below = ~(word)0;
  else if (i > lo) below = ((word)1 << (i - lo)) - 1;
  if (lower) {
    v &= below;
    if (i >= lo && i < lo + 64) v |= (word)1 << (i - lo);
  } else {
    v &= ~below;
  }
  if (w == rw - 1 && (r & 63)) v &= ((word)1 << (r & 63)) - 1;
  dst[(size_t)i * ds + w] = v;
}
// Inversion ([A | I]): the right half of a row is its own identity bit plus the identity bits of
// the pivot rows it has absorbed, so the only columns of the right half that a trailing update
// can touch are those of rows that have been pivot rows — this records how far they reach.
__global__ void max_pivot_id_kernel(const ElimState *st, const int *rowid, ElimGlobal *gl) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < st->kbn) atomicMax(&gl->max_id, rowid[st->blk_rows[t]]);
}
__global__ void iota_kernel(int *a, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = i;
}

__global__ void set_identity_kernel(word *W, int ws, int n, int col0) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) W[(size_t)i * ws + (col0 + i) / 64] |= (word)1 << ((col0 + i) % 64);
}

static int block_width() {
  static const int b = [] {
    const char *e = getenv("M4RI_B200_ELIM_BLOCK");
    int v = e ? atoi(e) : EB_DEFAULT;
    v = v / 128 * 128;
    return v < 128 ? 128 : (v > EB_MAX ? EB_MAX : v);
  }();
  return b;
}

}  // namespace

// Non-blocking "has this event happened yet". The host runs ahead of the device here, so
// cudaErrorNotReady is the usual answer; it is not a failure, and it must not stay behind as the
// thread's last error for the next M4B_LAUNCHED() to find.
static bool event_done(cudaEvent_t e) {
  const cudaError_t r = cudaEventQuery(e);
  if (r == cudaSuccess) return true;
  if (r == cudaErrorNotReady) {
    (void)cudaGetLastError();
    return false;
  }
  M4B_CUDA(r);
  return false;
}

// Reduced row echelon form of W in place, pivots searched in columns [0, pivot_cols) only (the
// rest is carried along: the right-hand side of a solve, the identity of an inversion).
// Returns the rank; pivots (if given) receives the pivot columns in increasing order.
//
// lu = true is the PLE flavour of the same blocked schedule (k_ple below): pivots column by
// column, rows eliminated only below and right of a pivot, multipliers left at the pivot columns;
// d_rowid (m ints on the device, identity on entry) follows the rows through the permutations
// and d_pivots_out (may be NULL, >= min(m, pivot_cols) + block ints) receives the pivot columns.
// ident_col0 >= 0: columns [ident_col0, ident_col0 + m) hold the identity on entry (inversion) —
// the trailing products then stop at the last column a pivot row can have touched.
static int eliminate(DMat *W, int pivot_cols, bool lu, std::vector<int> *pivots, int *d_rowid, int *d_pivots_out,
                     cudaStream_t s, int ident_col0 = -1) {
  const int m = W->nrows, n = W->ncols;
  if (pivots) pivots->clear();
  if (m == 0 || n == 0 || pivot_cols <= 0) return 0;
  if (pivot_cols > n) pivot_cols = n;
  if (((uintptr_t)W->d & 15) || (W->rowstride & 1)) die("m4ri_b200: rref needs 16-byte aligned rows\n");
  const int B = block_width(), bw = B / 64, L = 2 * bw;
  static const bool apply2 = getenv("M4RI_B200_ELIM_APPLY1") == nullptr;  // two words per thread in the apply kernel
  DMat *PT = dmat_new(m, 2 * B, s);       // [panel | tags], stride L
  DMat *OP = dmat_new(B, n, s);           // original pivot rows of the block (trailing columns)
  const int pstride = PT->rowstride, ws = W->rowstride;
  const int nblocks = (pivot_cols + B - 1) / B, maxrank = std::min(m, pivot_cols);
  ElimState *st;
  ElimGlobal *gl;
  unsigned char *ispiv;
  int *pidx, *perm, *piv_all, *tmpid = nullptr;
  unsigned long long *mult;
  word *Rp, *tmp;
  M4B_CUDA(cudaMallocAsync(&st, sizeof(ElimState), s));
  M4B_CUDA(cudaMallocAsync(&gl, sizeof(ElimGlobal), s));
  M4B_CUDA(cudaMemsetAsync(gl, 0, sizeof(ElimGlobal), s));
  M4B_CUDA(cudaMallocAsync(&ispiv, (size_t)m, s));
  M4B_CUDA(cudaMallocAsync(&mult, (size_t)m * 8, s));
  M4B_CUDA(cudaMallocAsync(&pidx, (size_t)m * sizeof(int), s));
  M4B_CUDA(cudaMallocAsync(&Rp, (size_t)256 * L * 8, s));  // the step's 16 four-bit tables
  M4B_CUDA(cudaMallocAsync(&tmp, (size_t)2 * B * W->width * 8, s));
  M4B_CUDA(cudaMallocAsync(&perm, (size_t)4 * B * sizeof(int), s));
  if (d_pivots_out) piv_all = d_pivots_out;
  else M4B_CUDA(cudaMallocAsync(&piv_all, (size_t)(maxrank + B) * sizeof(int), s));
  int *own_rowid = nullptr;
  if (ident_col0 >= 0 && !d_rowid) {
    M4B_CUDA(cudaMallocAsync(&own_rowid, (size_t)m * sizeof(int), s));
    iota_kernel<<<cdiv(m, 256), 256, 0, s>>>(own_rowid, m);
    M4B_LAUNCHED();
    d_rowid = own_rowid;
  }
  if (d_rowid) M4B_CUDA(cudaMallocAsync(&tmpid, (size_t)2 * B * sizeof(int), s));
  // rank after every block, copied home asynchronously: the host never waits for it, it only
  // looks (two blocks late) whether the rank has reached the row count and stops enqueueing
  static thread_local int *h_rank = nullptr;  // pinned, kept for the thread's lifetime
  static thread_local int h_rank_cap = 0;
  if (h_rank_cap < nblocks + 1) {
    if (h_rank) M4B_CUDA(cudaFreeHost(h_rank));
    h_rank_cap = std::max(nblocks + 1, 1024);
    M4B_CUDA(cudaMallocHost(&h_rank, (size_t)2 * h_rank_cap * sizeof(int)));  // second half: max_id per block
  }
  int *h_maxid = h_rank + h_rank_cap;
  std::vector<cudaEvent_t> ev(nblocks);
  for (auto &e : ev) M4B_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  int blocks_done = 0;
  for (int b = 0, c0 = 0; c0 < pivot_cols; b++, c0 += B) {
    if (b >= 2 && event_done(ev[b - 2]) && h_rank[b - 2] >= m) break;  // every row is a pivot row
    const int cw0 = c0 / 64;
    const int bcols = std::min(B, pivot_cols - c0);            // columns of this block that may hold pivots
    const int steps = (bcols + 63) / 64;
    M4B_CUDA(cudaMemsetAsync(st, 0, sizeof(ElimState), s));
    M4B_CUDA(cudaMemsetAsync(ispiv, 0, (size_t)m, s));
    M4B_CUDA(cudaMemsetAsync(pidx, 0, (size_t)m * sizeof(int), s));
    extract_panel_kernel<<<cdiv((long long)m * L, 256), 256, 0, s>>>(PT->d, pstride, W->d, ws, W->width, m, cw0, bw);
    M4B_LAUNCHED();
    for (int j = 0; j < steps; j++) {
      const int valid = std::min(64, bcols - 64 * j);
      const word colmask = valid == 64 ? ~(word)0 : (((word)1 << valid) - 1);
      if (lu) {
        // the word column itself is finished by lu_prepare_kernel; the rank-64 update starts right of it
        launch_chained(lu_pivot_search_kernel, 1, 1024, s, PT->d, pstride, j, m, gl, ispiv, pidx, st, c0, colmask);
        const int nrc = (int)cdiv(L - j - 1, RP_XW);
        launch_chained(lu_prepare_kernel, nrc + cdiv(m, 256), 256, s, Rp, L, bw, nrc, PT->d, pstride, j, m, gl, ispiv, pidx,
                       mult, st);
        if (apply2)
          launch_chained(apply_rprime2_kernel, cdiv((long long)m * ((L - ((j + 1) & ~1)) / 2), 256), 256, s, PT->d, pstride,
                         j + 1, L, m, pidx, mult, Rp, st);
        else
          launch_chained(apply_rprime_kernel, cdiv((long long)m * (L - j - 1), 256), 256, s, PT->d, pstride, j + 1, L, m,
                         pidx, mult, Rp, st);
        continue;
      }
      launch_chained(pivot_search_kernel, 1, 1024, s, PT->d, pstride, j, m, gl, ispiv, pidx, st, c0, colmask);
      const int nrc = (int)cdiv(L - j, RP_XW);
      launch_chained(prepare_step_kernel, nrc + cdiv(m, 256), 256, s, Rp, mult, L, PT->d, pstride, j, bw, m, nrc, st);
      if (apply2)
        launch_chained(apply_rprime2_kernel, cdiv((long long)m * ((L - (j & ~1)) / 2), 256), 256, s, PT->d, pstride, j, L, m,
                       pidx, mult, Rp, st);
      else
        launch_chained(apply_rprime_kernel, cdiv((long long)m * (L - j), 256), 256, s, PT->d, pstride, j, L, m, pidx, mult, Rp,
                       st);
    }
    // trailing update: everything right of the block, as ONE product of fixed shape (the pivot
    // count stays on the device: rows of OP and columns of T beyond it are zero)
    const int t0 = c0 + B;  // first trailing column (multiple of 128)
    int n_end = n;          // one past the last trailing column that can change
    if (ident_col0 >= 0 && t0 < n) {
      // (one host round trip per block: the steps of the block must have finished anyway before
      // its product can start, and the product it shortens is worth far more)
      max_pivot_id_kernel<<<cdiv(B, 256), 256, 0, s>>>(st, d_rowid, gl);
      M4B_LAUNCHED();
      M4B_CUDA(cudaMemcpyAsync(&h_maxid[b], &gl->max_id, sizeof(int), cudaMemcpyDeviceToHost, s));
      M4B_CUDA(cudaStreamSynchronize(s));
      n_end = std::min(n, std::max(t0, ident_col0 + (h_maxid[b] + 1 + 127) / 128 * 128));
    }
    if (t0 < n_end) {
      const int tw0 = t0 / 64, tw = width_of(n_end) - tw0;
      M4B_CUDA(cudaMemset2DAsync(OP->d, (size_t)OP->rowstride * 8, 0, (size_t)tw * 8, B, s));
      gather_pivot_rows_kernel<<<cdiv((long long)B * (tw + 1), 256), 256, 0, s>>>(OP->d, OP->rowstride, W->d, ws, tw0, tw,
                                                                                 PT->d, pstride, bw, st);
      M4B_LAUNCHED();
      // PLE: rows above the rank are finished (their tags are zero) — skip them, using the newest
      // rank that has already come home (a lower bound of the current one)
      // (the host stays at most three blocks ahead of the device: enough queued work to keep it
      // busy, and a rank that is at most three blocks old)
      int row_lb = 0;
      if (lu && b >= 3) {
        M4B_CUDA(cudaEventSynchronize(ev[b - 3]));
        int q = b - 3;
        while (q + 1 < b && event_done(ev[q + 1])) q++;
        row_lb = h_rank[q] / 256 * 256;
      }
      if (row_lb < m) {
        DMat Wt = dwin(W, row_lb, t0, m - row_lb, n_end - t0), Tg = dwin(PT, row_lb, B, m - row_lb, B),
             Ot = dwin(OP, 0, 0, B, n_end - t0);
        dmul(&Wt, &Tg, &Ot, 0, M4RI_B200_ALGO_AUTO, s);
      }
    }
    writeback_panel_kernel<<<cdiv((long long)m * bw, 256), 256, 0, s>>>(W->d, ws, W->width, PT->d, pstride, m, cw0, bw);
    M4B_LAUNCHED();
    // pivot rows up to rows rank .. rank+kb-1 in pivot-column order, planned and done on the device
    plan_permutation_kernel<<<1, 1024, 0, s>>>(st, gl, perm, perm + 2 * B, piv_all);
    M4B_LAUNCHED();
    gather_rows_kernel<<<cdiv((long long)2 * B * W->width, 256), 256, 0, s>>>(tmp, W->d, ws, W->width, perm, gl,
                                                                                 tmpid, d_rowid);
    M4B_LAUNCHED();
    scatter_rows_kernel<<<cdiv((long long)2 * B * W->width, 256), 256, 0, s>>>(W->d, ws, W->width, tmp, perm + 2 * B,
                                                                                  gl, tmpid, d_rowid);
    M4B_LAUNCHED();
    M4B_CUDA(cudaMemcpyAsync(&h_rank[b], &gl->rank, sizeof(int), cudaMemcpyDeviceToHost, s));
    M4B_CUDA(cudaEventRecord(ev[b], s));
    blocks_done = b + 1;
  }
  M4B_CUDA(cudaStreamSynchronize(s));
  const int rank = blocks_done ? h_rank[blocks_done - 1] : 0;
  if (pivots && rank > 0) {
    pivots->resize(rank);
    M4B_CUDA(cudaMemcpyAsync(pivots->data(), piv_all, (size_t)rank * sizeof(int), cudaMemcpyDeviceToHost, s));
    M4B_CUDA(cudaStreamSynchronize(s));
  }
  for (auto &e : ev) M4B_CUDA(cudaEventDestroy(e));
  M4B_CUDA(cudaFreeAsync(st, s));
  M4B_CUDA(cudaFreeAsync(gl, s));
  M4B_CUDA(cudaFreeAsync(ispiv, s));
  M4B_CUDA(cudaFreeAsync(pidx, s));
  M4B_CUDA(cudaFreeAsync(mult, s));
  M4B_CUDA(cudaFreeAsync(Rp, s));
  M4B_CUDA(cudaFreeAsync(tmp, s));
  M4B_CUDA(cudaFreeAsync(perm, s));
  if (!d_pivots_out) M4B_CUDA(cudaFreeAsync(piv_all, s));
  if (tmpid) M4B_CUDA(cudaFreeAsync(tmpid, s));
  if (own_rowid) M4B_CUDA(cudaFreeAsync(own_rowid, s));
  dmat_delete(PT, s);
  dmat_delete(OP, s);
  return rank;
}

int k_rref(DMat *W, int pivot_cols, std::vector<int> *pivots, cudaStream_t s) {
  return eliminate(W, pivot_cols, false, pivots, nullptr, nullptr, s);
}
// The same for [A | I] with the identity starting at column ident_col0 (inversion).
static int k_rref_identity(DMat *W, int pivot_cols, int ident_col0, cudaStream_t s) {
  static const bool off = getenv("M4RI_B200_INV_FULL_WIDTH") != nullptr;
  return eliminate(W, pivot_cols, false, nullptr, nullptr, nullptr, s, off ? -1 : ident_col0);
}

// PLE / PLUQ decomposition of W in place (m4ri-sys/src/ple.rs:35,65). On return, with r the rank:
//   rows 0 .. r-1    row i holds L[i][0 .. i) in columns [0, i) and, from column Q[i] on, row i of
//                    the echelon form E (E[i][Q[i]] = 1, zero left of it) — for pluq, from column
//                    i on, row i of U = E with columns i and Q[i] swapped for i = 0 .. r-1 in turn
//   rows r .. m-1    L[i][0 .. r) in columns [0, r), zero elsewhere
//   P (m entries)    row transpositions: swapping rows i and P[i] of the INPUT for i = 0 .. m-1 in
//                    turn gives L * E (so A = P L E with P the product of the swaps; P[i] >= i)
//   Q (n entries)    Q[i] = pivot column of row i for i < r (the column rank profile, increasing),
//                    Q[i] = i beyond
// Pivot rows are chosen deterministically (lowest row index among the candidates that stop at
// the column, in chunks of 1024 rows); M4RI's choice differs in general — the decomposition is
// not unique, its defining identity is what the tests check.
int k_ple(DMat *W, int *P, int *Q, bool pluq, cudaStream_t s) {
  const int m = W->nrows, n = W->ncols;
  for (int i = 0; i < m; i++) P[i] = i;
  for (int c = 0; c < n; c++) Q[c] = c;
  if (m == 0 || n == 0) return 0;
  const int B = block_width();
  int *rowid, *dq;
  M4B_CUDA(cudaMallocAsync(&rowid, (size_t)m * sizeof(int), s));
  M4B_CUDA(cudaMallocAsync(&dq, (size_t)(std::min(m, n) + B) * sizeof(int), s));
  iota_kernel<<<cdiv(m, 256), 256, 0, s>>>(rowid, m);
  M4B_LAUNCHED();
  std::vector<int> piv;
  const int r = eliminate(W, n, true, &piv, rowid, dq, s);
  std::vector<int> ids(m);
  M4B_CUDA(cudaMemcpyAsync(ids.data(), rowid, (size_t)m * sizeof(int), cudaMemcpyDeviceToHost, s));
  const int ws = W->rowstride, width = W->width;
  if (r > 0) {
    const int lcw = width_of(r);
    DMat *Lc = dmat_new(m, r, s);
    word *pm;
    int *pcnt;
    M4B_CUDA(cudaMallocAsync(&pm, (size_t)width * 8, s));
    M4B_CUDA(cudaMallocAsync(&pcnt, (size_t)width * sizeof(int), s));
    M4B_CUDA(cudaMemsetAsync(pm, 0, (size_t)width * 8, s));
    pivot_masks_kernel<<<cdiv(r, 256), 256, 0, s>>>(pm, dq, r);
    M4B_LAUNCHED();
    pivot_prefix_kernel<<<cdiv(width, 128), 128, 0, s>>>(pcnt, pm, width);
    M4B_LAUNCHED();
    gather_l_kernel<<<cdiv((long long)m * lcw, 256), 256, 0, s>>>(Lc->d, Lc->rowstride, lcw, W->d, ws, width, m, r, dq,
                                                                 pm, pcnt);
    M4B_LAUNCHED();
    M4B_CUDA(cudaFreeAsync(pm, s));
    M4B_CUDA(cudaFreeAsync(pcnt, s));
    strip_l_kernel<<<cdiv((long long)m * width, 256), 256, 0, s>>>(W->d, ws, width, m, r, dq);
    M4B_LAUNCHED();
    if (pluq) {
      // columns: swap i <-> Q[i] for i = 0 .. r-1 in turn; colsrc[c] = the column of E that lands at c
      std::vector<int> colsrc(n);
      for (int c = 0; c < n; c++) colsrc[c] = c;
      bool moved = false;
      for (int i = 0; i < r; i++)
        if (piv[i] != i) { std::swap(colsrc[i], colsrc[piv[i]]); moved = true; }
      if (moved) {
        int *dsrc;
        M4B_CUDA(cudaMallocAsync(&dsrc, (size_t)n * sizeof(int), s));
        M4B_CUDA(cudaMemcpyAsync(dsrc, colsrc.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
        DMat *U = dmat_new(r, n, s);
        permute_cols_kernel<<<cdiv((long long)r * width, 256), 256, 0, s>>>(U->d, U->rowstride, W->d, ws, r, width, n,
                                                                           dsrc);
        M4B_LAUNCHED();
        DMat top = dwin(W, 0, 0, r, n);
        k_copy(&top, U, s);
        M4B_CUDA(cudaStreamSynchronize(s));  // colsrc is read by the copy above
        dmat_delete(U, s);
        M4B_CUDA(cudaFreeAsync(dsrc, s));
      }
    }
    merge_l_kernel<<<cdiv((long long)m * lcw, 256), 256, 0, s>>>(W->d, ws, Lc->d, Lc->rowstride, lcw, m);
    M4B_LAUNCHED();
    dmat_delete(Lc, s);
  } else {
    k_zero(W, s);
  }
  M4B_CUDA(cudaStreamSynchronize(s));
  M4B_CUDA(cudaFreeAsync(rowid, s));
  M4B_CUDA(cudaFreeAsync(dq, s));
  for (int i = 0; i < r; i++) Q[i] = piv[i];
  // the final arrangement (position p holds input row ids[p]) as transpositions P[i] >= i
  std::vector<int> cur(m), pos(m);
  for (int i = 0; i < m; i++) cur[i] = pos[i] = i;
  for (int i = 0; i < m; i++) {
    const int want = ids[i], at = pos[want];
    P[i] = at;
    const int other = cur[i];
    cur[at] = other;
    pos[other] = at;
    cur[i] = want;
    pos[want] = i;
  }
  return r;
}

void k_set_identity(DMat *W, int n, int col0, cudaStream_t s) {
  if (n <= 0) return;
  set_identity_kernel<<<cdiv(n, 256), 256, 0, s>>>(W->d, W->rowstride, n, col0);
  M4B_LAUNCHED();
}

}  // namespace m4b

using namespace m4b;

extern "C" {

#ifdef M4RI_B200_PERF_HOOKS
void m4ri_b200_elim_clocks(long long *out) { M4B_CUDA(cudaMemcpyFromSymbol(out, g_clk, sizeof(long long) * 16)); }
void m4ri_b200_elim_iters(long long *out) { M4B_CUDA(cudaMemcpyFromSymbol(out, g_it, sizeof(long long) * 64)); }
#endif

// ---- device-resident forms (additive API) --------------------------------------------
int m4ri_b200_drref(m4ri_b200_dmat_t *W, int pivot_cols, int *pivot_columns, int max_pivots) {
  std::vector<int> pv;
  const int r = k_rref(W, pivot_cols <= 0 ? W->ncols : pivot_cols, &pv, current_stream());
  if (pivot_columns)
    for (int i = 0; i < r && i < max_pivots; i++) pivot_columns[i] = pv[i];
  return r;
}

int m4ri_b200_dple(m4ri_b200_dmat_t *W, rci_t *P, rci_t *Q, int pluq) {
  return k_ple(W, P, Q, pluq != 0, current_stream());
}

// ---- host ABI: the reference's names ---------------------------------------------------
static rci_t host_ple(mzd_t *A, mzp_t *P, mzp_t *Q, bool pluq, const char *who) {
  if (P->length != A->nrows || Q->length != A->ncols)
    die("m4ri_b200: %s: Permutation P length (%d) must match A nrows (%d) and Q length (%d) A ncols (%d)\n", who,
        P->length, A->nrows, Q->length, A->ncols);
  if (A->nrows == 0 || A->ncols == 0) {
    for (rci_t i = 0; i < P->length; i++) P->values[i] = i;
    for (rci_t i = 0; i < Q->length; i++) Q->values[i] = i;
    return 0;
  }
  cudaStream_t s = current_stream();
  DMat *dA = dmat_new(A->nrows, A->ncols, s);
  upload(dA, A, s);
  const int r = k_ple(dA, P->values, Q->values, pluq, s);
  download(A, dA, s);
  dmat_delete(dA, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return r;
}
rci_t mzd_ple(mzd_t *A, mzp_t *P, mzp_t *Q, int cutoff) {
  (void)cutoff;
  return host_ple(A, P, Q, false, "mzd_ple");
}
rci_t mzd_pluq(mzd_t *A, mzp_t *P, mzp_t *Q, int cutoff) {
  (void)cutoff;
  return host_ple(A, P, Q, true, "mzd_pluq");
}

// forward declarations (defined below)
static std::vector<int> compose_swaps(const mzp_t *P, int n, bool reverse);

// rows [0, n) of D permuted in place by a gather list
static void permute_rows_inplace(DMat *D, int n, const std::vector<int> &src, cudaStream_t s) {
  bool moved = false;
  for (int i = 0; i < n && !moved; i++) moved = src[i] != i;
  if (!moved) return;
  int *dsrc;
  M4B_CUDA(cudaMallocAsync(&dsrc, (size_t)n * sizeof(int), s));
  M4B_CUDA(cudaMemcpyAsync(dsrc, src.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
  DMat *T = dmat_new(n, D->ncols, s);
  permute_rows_kernel<<<cdiv((long long)n * D->width, 256), 256, 0, s>>>(T->d, T->rowstride, D->d, D->rowstride, n,
                                                                      D->width, dsrc);
  M4B_LAUNCHED();
  DMat top = dwin(D, 0, 0, n, D->ncols);
  k_copy(&top, T, s);
  M4B_CUDA(cudaStreamSynchronize(s));  // src (host) and dsrc are done with
  M4B_CUDA(cudaFreeAsync(dsrc, s));
  dmat_delete(T, s);
}

// Solves (P L U Q) X = B with the packed factors mzd_pluq left in A (solve.rs:53): rows of B
// permuted by P, Y = L11^-1 B[:r] and U11^-1 Y as reduced row echelon forms of [L11 | .] and
// [U11 | .] on the GPU, the consistency check B[r:] == L21 Y as one product, X = Z permuted by Q
// (free variables zero). X overwrites B (its other rows are zeroed); -1 and B untouched when the
// system is inconsistent.
int mzd_pluq_solve_left(const mzd_t *A, rci_t rank, const mzp_t *P, const mzp_t *Q, mzd_t *B, int cutoff,
                        int inconsistency_check) {
  (void)cutoff;
  (void)inconsistency_check;  // always checked
  const rci_t m = A->nrows, n = A->ncols, k = B->ncols, r = rank;
  if (P->length != m || Q->length != n) die("m4ri_b200: mzd_pluq_solve_left: permutation lengths must match A\n");
  if (B->nrows < m || B->nrows < n) die("m4ri_b200: mzd_pluq_solve_left: B needs at least max(nrows(A), ncols(A)) rows\n");
  if (r < 0 || r > m || r > n) die("m4ri_b200: mzd_pluq_solve_left: rank %d out of range\n", r);
  if (k == 0 || B->nrows == 0) return 0;
  cudaStream_t s = current_stream();
  DMat *dB = dmat_new(B->nrows, k, s);
  upload(dB, B, s);
  permute_rows_inplace(dB, m, compose_swaps(P, m, false), s);
  DMat *dX = dmat_new(B->nrows, k, s);  // zero: free variables and the rows beyond ncols(A)
  int ok = 1;
  if (r > 0) {
    DMat *dA = dmat_new(m, n, s);
    upload(dA, A, s);
    const rci_t off = (r + 127) / 128 * 128;
    DMat *W = dmat_new(r, off + k, s);
    DMat left = dwin(W, 0, 0, r, r), right = dwin(W, 0, off, r, k), btop = dwin(dB, 0, 0, r, k);
    tri_extract_kernel<<<cdiv((long long)r * width_of(r), 256), 256, 0, s>>>(W->d, W->rowstride, dA->d, dA->rowstride, r, 1);
    M4B_LAUNCHED();
    k_copy(&right, &btop, s);
    if (k_rref(W, r, nullptr, s) != r) die("m4ri_b200: mzd_pluq_solve_left: L is not unit lower triangular\n");
    if (m > r) {  // B[r:] ^= L21 * Y must vanish
      DMat rest = dwin(dB, r, 0, m - r, k), L21 = dwin(dA, r, 0, m - r, r);
      dmul(&rest, &L21, &right, 0, M4RI_B200_ALGO_AUTO, s);
      ok = k_is_zero(&rest, s) ? 1 : 0;
    }
    if (ok) {
      // [U11 | Y] -> [I | Z]; Y sits in the right part already
      k_zero(&left, s);
      tri_extract_kernel<<<cdiv((long long)r * width_of(r), 256), 256, 0, s>>>(W->d, W->rowstride, dA->d, dA->rowstride, r, 0);
      M4B_LAUNCHED();
      if (k_rref(W, r, nullptr, s) != r) die("m4ri_b200: mzd_pluq_solve_left: U has a zero on its diagonal\n");
      DMat xtop = dwin(dX, 0, 0, r, k);
      k_copy(&xtop, &right, s);
      permute_rows_inplace(dX, n, compose_swaps(Q, n, true), s);
    }
    dmat_delete(W, s);
    dmat_delete(dA, s);
  } else if (m > 0) {
    DMat rest = dwin(dB, 0, 0, m, k);
    ok = k_is_zero(&rest, s) ? 1 : 0;
  }
  if (ok) download(B, dX, s);
  dmat_delete(dB, s);
  dmat_delete(dX, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return ok ? 0 : -1;
}

// The swaps i <-> P[i] (i ascending, or descending for `reverse`) composed into one gather list:
// position p of the result holds element src[p] of the input.
static std::vector<int> compose_swaps(const mzp_t *P, int n, bool reverse) {
  std::vector<int> src(n);
  for (int i = 0; i < n; i++) src[i] = i;
  const int len = std::min((int)P->length, n);
  for (int t = 0; t < len; t++) {
    const int i = reverse ? len - 1 - t : t, j = P->values[i];
    if (j < 0 || j >= n) die("m4ri_b200: permutation entry %d out of range at %d\n", j, i);
    std::swap(src[i], src[j]);
  }
  return src;
}
static void host_apply_p(mzd_t *A, const mzp_t *P, bool columns, bool reverse) {
  if (A->nrows == 0 || A->ncols == 0) return;
  const int n = columns ? A->ncols : A->nrows;
  const std::vector<int> src = compose_swaps(P, n, reverse);
  bool moved = false;
  for (int i = 0; i < n && !moved; i++) moved = src[i] != i;
  if (!moved) return;
  cudaStream_t s = current_stream();
  DMat *dA = dmat_new(A->nrows, A->ncols, s), *dO = dmat_new(A->nrows, A->ncols, s);
  upload(dA, A, s);
  int *dsrc;
  M4B_CUDA(cudaMallocAsync(&dsrc, (size_t)n * sizeof(int), s));
  M4B_CUDA(cudaMemcpyAsync(dsrc, src.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice, s));
  const long long work = (long long)A->nrows * dA->width;
  if (columns)
    permute_cols_kernel<<<cdiv(work, 256), 256, 0, s>>>(dO->d, dO->rowstride, dA->d, dA->rowstride, A->nrows, dA->width,
                                                       A->ncols, dsrc);
  else
    permute_rows_kernel<<<cdiv(work, 256), 256, 0, s>>>(dO->d, dO->rowstride, dA->d, dA->rowstride, A->nrows, dA->width,
                                                       dsrc);
  M4B_LAUNCHED();
  download(A, dO, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  M4B_CUDA(cudaFreeAsync(dsrc, s));
  dmat_delete(dA, s);
  dmat_delete(dO, s);
}
// rows i <-> P[i], i = 0 .. len-1 in turn (mzp.rs:39); _trans: the same swaps in reverse order
void mzd_apply_p_left(mzd_t *A, const mzp_t *P) { host_apply_p(A, P, false, false); }
void mzd_apply_p_left_trans(mzd_t *A, const mzp_t *P) { host_apply_p(A, P, false, true); }
// columns i <-> P[i]: _trans in turn (mzp.rs:48), plain in reverse order (mzp.rs:45)
void mzd_apply_p_right_trans(mzd_t *A, const mzp_t *P) { host_apply_p(A, P, true, false); }
void mzd_apply_p_right(mzd_t *A, const mzp_t *P) { host_apply_p(A, P, true, true); }

static rci_t host_echelonize(mzd_t *A, const char *who) {
  (void)who;
  if (A->nrows == 0 || A->ncols == 0) return 0;
  cudaStream_t s = current_stream();
  DMat *dA = dmat_new(A->nrows, A->ncols, s);
  upload(dA, A, s);
  const int r = k_rref(dA, A->ncols, nullptr, s);
  download(A, dA, s);
  dmat_delete(dA, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return r;
}

// (Reduced) row echelon form in place; returns the rank. `full == 0` asks only for an upper
// triangular form: the reduced form returned here is one (echelonform.rs:9-16).
rci_t mzd_echelonize(mzd_t *A, int full) {
  (void)full;
  return host_echelonize(A, "mzd_echelonize");
}
rci_t mzd_echelonize_pluq(mzd_t *A, int full) {
  (void)full;
  return host_echelonize(A, "mzd_echelonize_pluq");
}
rci_t mzd_echelonize_m4ri(mzd_t *A, int full, int k) {
  (void)full;
  (void)k;
  return host_echelonize(A, "mzd_echelonize_m4ri");
}

// dst = src^-1 (NULL: allocate), NULL when src is singular. [src | I] -> [I | src^-1].
mzd_t *mzd_inv_m4ri(mzd_t *dst, const mzd_t *src, int k) {
  (void)k;
  const rci_t n = src->nrows;
  if (src->ncols != n) die("m4ri_b200: mzd_inv_m4ri: matrix must be square (%d x %d)\n", src->nrows, src->ncols);
  if (dst && (dst->nrows != n || dst->ncols != n)) die("m4ri_b200: mzd_inv_m4ri: dst has wrong dimensions\n");
  if (dst == src) die("m4ri_b200: mzd_inv_m4ri: dst must not alias src\n");
  if (n == 0) return dst ? dst : mzd_init(0, 0);
  cudaStream_t s = current_stream();
  // the identity starts on a 128-bit boundary so that the inverse is an aligned window
  const rci_t off = (n + 127) / 128 * 128;
  DMat *W = dmat_new(n, off + n, s);
  DMat left = dwin(W, 0, 0, n, n);
  upload(&left, src, s);
  k_set_identity(W, n, off, s);
  const int r = k_rref_identity(W, n, off, s);
  mzd_t *out = nullptr;
  if (r == n) {
    out = dst ? dst : mzd_init_impl(n, n, 0);
    DMat right = dwin(W, 0, off, n, n);
    download(out, &right, s);
  }
  dmat_delete(W, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return out;
}

// mzd.rs:218 — inverse by elimination; the identity argument only saved M4RI an allocation.
// NULL when A is singular (M4RI's naive routine does not notice and returns garbage).
mzd_t *mzd_invert_naive(mzd_t *inv, const mzd_t *A, const mzd_t *identity) {
  (void)identity;
  return mzd_inv_m4ri(inv, A, 0);
}
// brilliantrussian.rs:199 — completes an upper triangular form to the reduced one: the reduced
// row echelon form of whatever is passed in (the same unique object); returns the rank.
rci_t mzd_top_echelonize_m4ri(mzd_t *A, int k) {
  (void)k;
  return host_echelonize(A, "mzd_top_echelonize_m4ri");
}

// Solves A X = B; X overwrites the first ncols(A) rows of B (the rest of B is zeroed). Returns 0,
// or -1 when the system is inconsistent (B is then left untouched). Free variables are zero.
// A is "overwritten" in the reference (solve.rs:18); here it is left as it was.
int mzd_solve_left(mzd_t *A, mzd_t *B, int cutoff, int inconsistency_check) {
  (void)cutoff;
  (void)inconsistency_check;  // the check is free here: always done
  const rci_t m = A->nrows, n = A->ncols, k = B->ncols;
  if (B->nrows < n) die("m4ri_b200: mzd_solve_left: A ncols (%d) need to be lower than B nrows (%d)\n", n, B->nrows);
  if (B->nrows < m) die("m4ri_b200: mzd_solve_left: A nrows (%d) need to be lower than B nrows (%d)\n", m, B->nrows);
  if (m == 0 || n == 0 || k == 0) return 0;
  cudaStream_t s = current_stream();
  const rci_t off = (n + 127) / 128 * 128;
  DMat *W = dmat_new(m, off + k, s);
  DMat left = dwin(W, 0, 0, m, n), right = dwin(W, 0, off, m, k);
  upload(&left, A, s);
  {  // rows [0, m) of B
    mzd_t Bv = *B;
    Bv.nrows = m;
    DMat Rt = right;
    upload(&Rt, &Bv, s);
  }
  std::vector<int> piv;
  const int r = k_rref(W, n, &piv, s);
  // consistent iff the rows below the rank are zero in the right part
  int ok = 1;
  if (r < m) {
    DMat rest = dwin(W, r, off, m - r, k);
    ok = k_is_zero(&rest, s) ? 1 : 0;
  }
  if (ok) {
    // X[pivot column t] = right part of row t; other rows of X (free variables) are zero
    mzd_t *R = mzd_init_impl(r > 0 ? r : 1, k, 1);
    if (r > 0) {
      DMat top = dwin(W, 0, off, r, k);
      mzd_t Rv = *R;
      Rv.nrows = r;
      download(&Rv, &top, s);
    }
    M4B_CUDA(cudaStreamSynchronize(s));
    for (rci_t i = 0; i < B->nrows; i++) {
      word *row = B->rows[i];
      for (wi_t w = 0; w + 1 < B->width; w++) row[w] = 0;
      if (B->width) row[B->width - 1] &= ~B->high_bitmask;
    }
    for (int t = 0; t < r; t++) {
      word *row = B->rows[piv[t]];
      const word *srcw = R->rows[t];
      for (wi_t w = 0; w + 1 < B->width; w++) row[w] = srcw[w];
      if (B->width) row[B->width - 1] |= srcw[B->width - 1] & B->high_bitmask;
    }
    mzd_free(R);
  }
  dmat_delete(W, s);
  M4B_CUDA(cudaStreamSynchronize(s));
  return ok ? 0 : -1;
}

}  // extern "C"
