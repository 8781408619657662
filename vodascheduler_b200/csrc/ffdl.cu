// ffdl.cu — FfDLOptimizer.Schedule (pkg/algorithm/ffdl_optimizer.go:32-130) for a batch of
// rounds, hand-written for sm_100a.
//
// One WARP (a 32-thread CTA) owns one round at a time (persistent grid striding over rounds).
// Per round:
//   1. stable sort of the jobs by SubmitTime (:45-47)            -> bitonic argsort
//   2. trim to the first J = min(n, K) jobs (:54-65)
//   3. DP  P[j][k] = max_g S_j[g] + P[j-1][k-g], g in [1, max_j], k-g >= 0, rows initialised
//      to -10000 (:78-105) -> ONE ROW PER STEP, IN PLACE: the row lives in shared memory
//      (float4 groups); a cell only reads cells to its left, so the warp takes the row's tiles
//      from the right end, 32 at a time, every lane reading its window of row j-1 before row j
//      lands on it (a __syncwarp, no CTA barrier anywhere in the row loop).  A lane owns 8
//      adjacent cells and slides a three-group register window of row j-1 over g, so a block of
//      4 g-steps costs one LDS.128 of P for 32 add+max pairs.  The forward pass keeps ONLY the
//      values: per candidate it is one FADD and half a three-input FMNMX3.
//   4. The reference also records SOL[j][k] = the g that won (strict '>', so the smallest g
//      among equals).  Tracking it costs two more instructions per candidate on the pipe that
//      already bounds the loop, and only J of the J*K entries are ever read.  Instead the rows
//      of P are streamed to the round's slab of global scratch, and the backtrack (:116-125),
//      a SECOND kernel, re-evaluates the candidates of the J cells on its path from the stored
//      row j-1 — the same float additions in the same order with the same strict compare, so
//      it finds the same g bit for bit.
//   5. P[J][K] <= 0 -> "infeasible" (:109-114); deferred validateResult (:42).
//
// Two exact trims keep the table small (properties of the recurrence, not heuristics):
//   saturation: for k >= klim_j = sum of the first j jobs' max sizes the GPU budget no longer
//     binds and P[j][k] is one constant c_j — bit for bit, because every candidate then adds
//     the same two numbers; tiles beyond klim_j are filled with c_j, not computed, not stored;
//   cone: the backtrack starts at k = K and gives up at most max_i per row, so row j is only
//     ever read at k >= klo_j = K - (sum of the later jobs' max sizes), and such a cell reads
//     row j-1 at k-g >= klo_{j-1}; tiles wholly below klo_j are not computed (what the row
//     buffer holds there is never read by a cell that matters).
//
// float32 evaluation order: p = S_j[g] + P[j-1][k-g] is a single rounded add (-fmad=false;
// there is nothing to contract anyway).  Cells with k-g < 0 are fed -inf, which can never win
// (and -inf + +inf = NaN never does either), which reproduces the reference's `if k-g >= 0`
// guard without a branch.  As far as the VALUE goes fmaxf(best, p) equals the reference's
// strict-'>' select for every input (equal values are interchangeable; a NaN p never replaces
// best in either form).
#include "common.cuh"
#include "kernels.h"

namespace voda {

namespace {

#ifndef VODA_FFDL_MINBLOCKS
#define VODA_FFDL_MINBLOCKS 20
#endif
// One warp per round: the row loop needs no CTA barrier, and the warp updates ONE row buffer in
// place (tiles are taken from the right end of the band: a cell only reads cells to its left).
constexpr int kThreads = 32;
constexpr int kCells = 8;  // adjacent cells per thread: two float4 groups, a three-group window
// Checkpointing: when no job of the round can take more than kCkptGmax GPUs (the usual case:
// the speedup tables stop at 32), only every kCkpt-th row of P is stored, and the backtrack
// recomputes the rows in between on the narrow band its walk can reach (kCkpt * kCkptGmax
// cells at most).  That cuts the scratch traffic, the DP kernel's main memory cost, by kCkpt.
#ifndef VODA_FFDL_CKPT
#define VODA_FFDL_CKPT 4
#endif
constexpr int kCkpt = VODA_FFDL_CKPT;
constexpr int kCkptGmax = 32;
static_assert(kCkpt == 4, "ffdl_walk_sparse dispatches the topmost segment on 1..4 rows");
constexpr int kBand = kCkpt * kCkptGmax;  // cells per band row
constexpr int kBandPad = kCkptGmax;       // -inf cells in front of a band row (columns below 0)
constexpr int kBandRow = kBandPad + kBand + 32;

// Per-row constants, worked out once per round after the sort and read back with one LDS.64
// and one LDS.32 per row: four 16-bit fields packed in a word, and the row maximum.
//   bits  0..15  nblk   blocks of 4 g-steps served from the table; bit 15: the row also has candidates g >= W
//   bits 16..31  tc0    tiles tc0 .. tcE-1 are computed
//   bits 32..47  tcE
//   bits 48..63  fillE  tiles tcE .. fillE-1 are filled with the saturated constant
__device__ __forceinline__ unsigned long long pack_row(int nblk, bool slow, int tc0, int tcE, int fillE) {
  return (unsigned long long)(unsigned)(nblk | (slow ? 0x8000 : 0)) | ((unsigned long long)(unsigned)tc0 << 16) |
         ((unsigned long long)(unsigned)tcE << 32) | ((unsigned long long)(unsigned)fillE << 48);
}

// Shared memory per CTA: one row buffer and 20 bytes per job; the sort keys' space is reused
// for the row constants once the sort is done.
struct FfdlSmem {
  float4* row;     // the row buffer (see below); the pointer is E[0]
  int64_t* khi;    // [Jp]  sort key (seconds); after the sort: the packed row constants
  int32_t* klo;    // [Jp]  sort key (nanoseconds); after the sort: srmax, max over the row's usable g of S_j[g]
  int* perm;       // [Jp]  sorted position -> job; then smaxc = clamp(max_proc, 0, K) in sorted order
  int32_t* pre;    // [Jp]  inclusive prefix sums of smaxc
  float4* sring;   // [4][SW/4]  S rows of the current and the next rows
};

// What the DP kernel leaves per round for the backtrack kernel ...
struct FfdlMeta {
  int32_t J;   // rows of the DP (jobs kept after the trim)
  int32_t st;  // VODA_ST_OK, or the panic raised before / by the DP
  int32_t K;
  int32_t sparse;  // 1: only every kCkpt-th row of P was stored (see the backtrack kernel)
};
// ... and per DP row j (1-based; record j-1), 16 bytes so that one load fetches it
struct __align__(16) FfdlRow {
  int32_t job;   // index of the row's job within the round (input order)
  int32_t gmax;  // min(max_proc, K): candidates are g = 1 .. min(gmax, k)
  int32_t klim;  // saturation point of THIS row: P[j][k] == c for k >= klim
  float c;       // the saturated value c_j
};

struct FfdlScratch {
  FfdlMeta* meta;     // [count]
  FfdlRow* rows;      // [count][Jcap]
  float* srows;       // [count][Jcap][SW]  S rows in sorted order, -inf beyond the table limit
  float* P;           // [count][slabRows][rowFloats].  A round that stores every row keeps P[j][1..K]
                      // in slab row j-1 (float k-1); a sparse round keeps P[c], c a multiple of kCkpt,
                      // in slab row c/kCkpt - 1: its stored rows lie together
  size_t rowFloats;   // 8 * KTcap
  size_t slabFloats;  // slabRows * rowFloats
  int Jcap;
  int SW;
  int sparseOnly;     // the slab holds checkpoint rows only (the declared max_max_proc makes every round sparse)
};

// A DP row in shared memory.  Group q holds P[4q-3 .. 4q] (group 0 = three cells left of the
// row and P[j][0]).  A thread's tile is two groups (2tq+1, 2tq+2) and its window three
// (2tq .. 2tq+2), so with the groups stored in order the lanes of a warp would touch every
// other 16 bytes: a two-way bank conflict on every LDS.128 / STS.128.  The groups are therefore
// stored DE-INTERLEAVED: the even groups E[i] = group 2i in one array, the odd groups
// O[i] = group 2i+1 in another (O = E + ost), and consecutive lanes touch consecutive 16 bytes.
// Both arrays carry pad_half(SW) groups of -inf in front, so the window can slide left of the
// row (k-g < 0) without a branch.
__host__ __device__ constexpr int pad_half(int SW) { return (SW / 4 + 1) / 2 + 1; }
__host__ __device__ constexpr int row_groups(int SW, int KTcap) { return 2 * pad_half(SW) + 2 * KTcap + 2; }
__device__ __forceinline__ const float4* grp(const float4* E, int ost, int q) {
  return E + (q >> 1) + ((q & 1) ? ost : 0);
}
__device__ __forceinline__ float cell(const float4* E, int ost, int x) {  // P[x], x >= 0
  return reinterpret_cast<const float*>(grp(E, ost, (x + 3) >> 2))[(x + 3) & 3];
}


// 4 g-steps (g = 4m+1 .. 4m+4) for the 8 cells of a tile: cell c at step s reads element
// 3+c-s of the window (lo, mid, hi) = 12 consecutive floats of row j-1; two candidates per
// three-input maximum.
__device__ __forceinline__ void ffdl_block8(float (&best)[8], const float4 sq, const float4 lo, const float4 mid,
                                            const float4 hi) {
  const float v[12] = {lo.x, lo.y, lo.z, lo.w, mid.x, mid.y, mid.z, mid.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
  for (int c = 0; c < 8; c++) {
    best[c] = fmaxf(fmaxf(best[c], sq.x + v[3 + c]), sq.y + v[2 + c]);
    best[c] = fmaxf(fmaxf(best[c], sq.z + v[1 + c]), sq.w + v[c]);
  }
}

__device__ __forceinline__ void ffdl_store_tile(float4* row, int ost, int tq, const float (&best)[8],
                                                float4* __restrict__ grow, bool act) {
  const float4 lo = make_float4(best[0], best[1], best[2], best[3]);
  const float4 hi = make_float4(best[4], best[5], best[6], best[7]);
  __syncwarp();  // every lane of the pass has read its window of row j-1 before row j lands on it
  if (!act) return;
  row[tq + ost] = lo;  // group 2tq+1
  row[tq + 1] = hi;    // group 2tq+2
  if (grow) {  // the slab row is in cell order; streamed: nothing reads it before the backtrack kernel
    __stcs(grow + 2 * tq, lo);
    __stcs(grow + 2 * tq + 1, hi);
  }
}

// One DP row for column tile `tq` (cells k = 8*tq+1 .. 8*tq+8) when every candidate comes from
// the first NB blocks of the table (g <= 4*NB): the common shapes, fully unrolled, S in registers.
// For the block of g = 4m+1 .. 4m+4 the cells need P[base-3 .. base+8], base = 8*tq - 4m, i.e.
// groups q0 .. q0+2 with q0 = 2*tq - m.
template <int NB>
__device__ __forceinline__ void ffdl_tile_nb(float4* row, int ost, int tq, const float4 sA,
                                             const float4 sB, float4* __restrict__ grow, bool act) {
  float best[8];
#pragma unroll
  for (int c = 0; c < 8; c++) best[c] = -10000.0f;
  const float4 g0 = row[tq], g1 = row[tq + ost], g2 = row[tq + 1];
  ffdl_block8(best, sA, g0, g1, g2);
  if (NB >= 2) {
    const float4 gm = row[tq - 1 + ost];
    ffdl_block8(best, sB, gm, g0, g1);
  }
  ffdl_store_tile(row, ost, tq, best, grow, act);
}

// The general tile: any number of blocks (S read from the round's sorted S rows in global
// memory), then the candidates g in [gslow0, gslow1] the dense table does not hold (the Go map
// misses, S = 0; rare: max_j >= W).
__device__ __forceinline__ void ffdl_tile_any(float4* row, int ost, int tq,
                                              const float4* __restrict__ s4, int nblk, int gslow0, int gslow1,
                                              float4* __restrict__ grow, bool act) {
  float best[8];
#pragma unroll
  for (int c = 0; c < 8; c++) best[c] = -10000.0f;
  const int qtop = 2 * tq;
  // the window is three groups; instead of shifting registers after every block the three
  // roles rotate through three unrolled copies of the block
  float4 g0 = *grp(row, ost, qtop), g1 = *grp(row, ost, qtop + 1), g2 = *grp(row, ost, qtop + 2);
  int m = 0;
  for (; m + 3 <= nblk; m += 3) {
    ffdl_block8(best, s4[m], g0, g1, g2);
    g2 = *grp(row, ost, qtop - m - 1);
    ffdl_block8(best, s4[m + 1], g2, g0, g1);
    g1 = *grp(row, ost, qtop - m - 2);
    ffdl_block8(best, s4[m + 2], g1, g2, g0);
    g0 = *grp(row, ost, qtop - m - 3);
  }
  if (m < nblk) {
    ffdl_block8(best, s4[m], g0, g1, g2);
    if (m + 1 < nblk) {
      g2 = *grp(row, ost, qtop - m - 1);
      ffdl_block8(best, s4[m + 1], g2, g0, g1);
    }
  }
  if (gslow1 >= gslow0) {
#pragma unroll
    for (int c = 0; c < 8; c++) {
      const int k = 8 * tq + 1 + c;
      for (int g = gslow0; g <= gslow1 && g <= k; g++) best[c] = fmaxf(best[c], 0.0f + cell(row, ost, k - g));
    }
  }
  ffdl_store_tile(row, ost, tq, best, grow, act);
}

// 16-byte asynchronous copy global -> shared (LDGSTS)
__device__ __forceinline__ void cp_async16(float4* dst, const float4* src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Everything the row loop and the backtrack need to know about the rows of one round, worked
// out once after the sort: the sorted S rows (to scratch), the row maxima, the two running
// sums behind the trims, the tile ranges.
__device__ int ffdl_setup_rows(const AllocArgs& a, const FfdlSmem& sm, int o, int J, int K, int KT, int SW,
                                float* gs, FfdlRow* grows) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int W = a.stride;
  unsigned long long* rowq = reinterpret_cast<unsigned long long*>(sm.khi);  // the sort keys are dead
  float* srmax = reinterpret_cast<float*>(sm.klo);
  int32_t* smaxc = sm.perm;  // slot j is read (perm) and then written (smaxc) by the same thread
  int32_t* pre = sm.pre;
  // a. a thread per row.  First the row's size limit (one dependent load per row, all rows of
  //    the thread in flight together), then S_j[1..] with -inf beyond the row's table limit:
  //    only the entries the row can use are read.
  int gtop = 0;
#pragma unroll 4
  for (int j = tid; j < J; j += kThreads) {
    const int pj = sm.perm[j];
    const int gmax = min(a.max_proc[o + pj], K);
    smaxc[j] = max(0, gmax);
    gtop = max(gtop, gmax);
    grows[j].job = pj;
    grows[j].gmax = gmax;
  }
  for (int j = tid; j < J; j += kThreads) {
    const int pj = grows[j].job;  // this thread's own stores
    const int gmax = grows[j].gmax;
    const int gt = min(gmax, W - 1);      // last g served from the table
    const bool slow = gmax >= max(W, 1);  // candidates the table does not hold read S = 0 (a Go map miss)
    const float* sj = a.speedup + (a.sp_off ? (size_t)a.sp_off[o + pj] : (size_t)(o + pj) * (size_t)W);
    float4* gj = reinterpret_cast<float4*>(gs + (size_t)j * SW);
    float rm = slow ? 0.0f : VODA_NEG_INF;
#pragma unroll 3
    for (int m = 0; m < SW / 4; m++) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int g = 4 * m + u + 1;
        v[u] = g <= gt ? sj[g] : VODA_NEG_INF;
        rm = fmaxf(rm, v[u]);
      }
      gj[m] = make_float4(v[0], v[1], v[2], v[3]);
    }
    srmax[j] = rm;
  }
  __syncthreads();
  // b. inclusive prefix sums of the usable sizes (warp 0, a contiguous chunk per lane)
  if (warp == 0) {
    const int per = (J + 31) / 32;
    const int lo = min(lane * per, J), hi = min(lo + per, J);
    int s = 0;
    for (int i = lo; i < hi; i++) s += smaxc[i];
    int incl = s;
    for (int d = 1; d < 32; d <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    int run = incl - s;
    for (int i = lo; i < hi; i++) {
      run += smaxc[i];
      pre[i] = run;
    }
  }
  __syncthreads();
  // c. the trims as tile ranges
  const int total = J > 0 ? pre[J - 1] : 0;
  for (int j = tid; j < J; j += kThreads) {
    const int gmax = grows[j].gmax;
    const int gt = min(gmax, W - 1);
    const int p = pre[j];
    const int klim = min(K, p);                 // saturation point of the row
    const int klo = max(0, K - (total - p));    // cone: the lowest k the backtrack can reach in this row
    const int klimNext = j + 1 < J ? min(K, pre[j + 1]) : K;  // how far the next row (or P[J][K]) reads
    const int tc0 = max(0, (klo + kCells - 1) / kCells - 1);  // first tile with a cell >= klo
    int tcE = klim >= 1 ? min(KT, (klim - 1) / kCells + 1) : 0;  // one past the last tile starting at or below klim
    if (tcE < tc0) tcE = tc0;
    int fillE = klimNext >= 1 ? min(KT, (klimNext - 1) / kCells + 1) : 0;
    if (fillE < tcE) fillE = tcE;
    rowq[j] = pack_row(gt > 0 ? (gt + 3) >> 2 : 0, gmax >= max(W, 1), tc0, tcE, fillE);
    grows[j].klim = klim;
  }
  __syncthreads();
  return __reduce_max_sync(0xffffffffu, gtop);  // kThreads == 32; the most GPUs any job of the round can take
}

// All DP rows of one round.  Returns with the last row in `*last` (shared memory).
__device__ void ffdl_rows(const AllocArgs& a, const FfdlSmem& sm, int J, int KT, int SW, int ost, bool sparse,
                          float* __restrict__ slab, size_t rowFloats, const float* gs, FfdlRow* grows,
                          const float4** last) {
  const int tid = threadIdx.x;
  float4* row = sm.row;
  // row 0: P[0][k] = 0 for every k (:80-82)
  for (int i = tid; i <= KT; i += kThreads) {
    row[i] = i == 0 ? make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, 0.0f) : make_float4(0.f, 0.f, 0.f, 0.f);
    row[i + ost] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const unsigned long long* rowq = reinterpret_cast<const unsigned long long*>(sm.khi);
  const float* srmax = reinterpret_cast<const float*>(sm.klo);
  // S rows reach shared memory through a four-slot ring filled by 16-byte asynchronous copies
  // (cp.async: global -> shared without a register in between) issued TWO rows ahead: a row of
  // the loop is shorter than an L2 round trip.  Everything that advances by a constant per row
  // is a running pointer (the row loop is short enough for its address arithmetic to show).
  const int sGroups = SW / 4;
  const bool stager = tid < sGroups;
  const float4* sgp = reinterpret_cast<const float4*>(gs) + tid;  // S row to stage next, this thread's 16 bytes
  float4* ringt = sm.sring + tid;
  int stageRow = 0;
  auto stage_s = [&]() {  // the S row of DP row stageRow (0-based) goes to slot stageRow & 3
    if (stager && stageRow < J) cp_async16(ringt + (stageRow & 3) * sGroups, sgp);
    cp_async_commit();
    sgp += sGroups;
    stageRow++;
  };
  stage_s();
  stage_s();
  cp_async_wait<0>();
  __syncthreads();
  const int sB = SW >= 8 ? 1 : 0;  // SW >= 4: the second block is clamped
  const int gslow0 = max(a.stride, 1);
  float cprev = 0.0f;  // c_0: P[0][k] = 0
  float4* const slab4 = reinterpret_cast<float4*>(slab);
  const size_t rowGroups = rowFloats / 4;
  FfdlRow* gr = grows;
  for (int j = 1; j <= J; j++, gr++) {
    const unsigned long long rq = rowq[j - 1];
    stage_s();  // row j+1; its slot was last read two barriers ago
    const float4* s4 = sm.sring + ((j - 1) & 3) * sGroups;
    const float4 sA = s4[0], sBv = s4[sB];
    const int nblk = (int)(rq & 0x7fff);
    const bool slow = (rq & 0x8000) != 0;
    const int tc0 = (int)((rq >> 16) & 0xffff), tcE = (int)((rq >> 32) & 0xffff), fillE = (int)(rq >> 48);
    // c_j = max(-10000, max_g (S_j[g] + c_{j-1})) = max(-10000, (max_g S_j[g]) + c_{j-1}): rounding
    // is monotone, so the maximum commutes with the addition exactly
    const float ccur = fmaxf(-10000.0f, srmax[j - 1] + cprev);
    if (tid == 0) gr->c = ccur;
    // row J is only read at [K], from shared memory; a sparse round stores every kCkpt-th row
    // (slab row j-1, or j/kCkpt - 1 among the checkpoint rows)
    float4* gdst = (j < J && (!sparse || j % kCkpt == 0)) ? slab4 + (size_t)(sparse ? j / kCkpt - 1 : j - 1) * rowGroups
                                                          : nullptr;
    // passes of 32 tiles from the right end of the band; lanes beyond the band recompute its last
    // tile and store nothing (every lane must reach the __syncwarp in front of the stores)
    {
      const float4 c4 = make_float4(ccur, ccur, ccur, ccur);
      for (int tq = tcE + tid; tq < fillE; tq += kThreads) {  // not read by this row's tiles
        row[tq + ost] = c4;
        row[tq + 1] = c4;
      }
    }
    const int gslow1 = slow ? gr->gmax : -1;  // written by this warp during the set-up
    for (int base = tc0 + ((tcE - tc0 - 1) & ~31); base >= tc0 && tcE > tc0; base -= 32) {
      const bool act = base + tid < tcE;
      const int tq = min(base + tid, tcE - 1);
      if (!slow && nblk == 1) ffdl_tile_nb<1>(row, ost, tq, sA, sBv, gdst, act);
      else if (!slow && nblk == 2) ffdl_tile_nb<2>(row, ost, tq, sA, sBv, gdst, act);
      else ffdl_tile_any(row, ost, tq, s4, nblk, gslow0, gslow1, gdst, act);
    }
    __syncwarp();
    cprev = ccur;
    if (tid == 0) row[0] = make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, -10000.0f);  // P[j][0]
    cp_async_wait<1>();  // everything but this row's stage_s has landed: the next row's S is in its slot
    __syncthreads();
  }
  *last = row;
}

// Kernel A: sort + DP rows.  One CTA per round at a time; P rows / row records / meta go to the
// round's own slot of the scratch so that the CTA can move on to its next round immediately.
__global__ void __launch_bounds__(kThreads, VODA_FFDL_MINBLOCKS)
    ffdl_dp_kernel(AllocArgs a, int r0, int count, FfdlScratch sc, int KTcap, int SW, int Jp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FfdlSmem sm;
  {
    unsigned char* p = smem_raw;
    const int ph = pad_half(SW);
    sm.row = reinterpret_cast<float4*>(p) + ph; p += sizeof(float4) * (size_t)row_groups(SW, KTcap);
    sm.sring = reinterpret_cast<float4*>(p); p += sizeof(float) * 4 * (size_t)SW;
    sm.khi = reinterpret_cast<int64_t*>(p); p += sizeof(int64_t) * (size_t)Jp;
    sm.klo = reinterpret_cast<int32_t*>(p); p += sizeof(int32_t) * (size_t)Jp;
    sm.perm = reinterpret_cast<int*>(p); p += sizeof(int) * (size_t)Jp;
    sm.pre = reinterpret_cast<int32_t*>(p);
  }
  const int tid = threadIdx.x;
  const int ost = pad_half(SW) + KTcap + 1;  // O[i] = E[i + ost]
  for (int q = tid; q < pad_half(SW); q += kThreads) {  // written once: nothing else stores there
    const float4 ninf = make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF);
    sm.row[-1 - q] = ninf;
    sm.row[-1 - q + ost] = ninf;
  }

  for (int rr = blockIdx.x; rr < count; rr += gridDim.x) {
    const int r = r0 + rr;
    const int o = a.job_off[r] - a.job_base;
    const int n = a.job_off[r + 1] - a.job_off[r];
    const int K = a.total_gpu[r];
    FfdlMeta meta{0, VODA_ST_OK, K, 0};
    if (n == 0) {  // :33-37 returns the empty result before anything else
      if (tid == 0) sc.meta[rr] = meta;
      continue;
    }
    for (int i = tid; i < n; i += kThreads) a.out_gpus[o + i] = 0;  // :38-40
    if (n > a.max_jobs || K > a.max_total) {
      meta.st = VODA_ST_BOUND;
      if (tid == 0) sc.meta[rr] = meta;
      continue;
    }
    if (K < 0) {
      meta.st = VODA_ST_RUNTIME_PANIC;  // make(ReadyJobs, 1, K+1) with cap < len (:54)
      if (tid == 0) sc.meta[rr] = meta;
      continue;
    }
    __syncthreads();  // the previous round's shared state is dead
    // 1. stable sort by SubmitTime: on 64-bit packed keys when every time of the round fits
    //    (seconds in [0, 2^32)), by the general argsort otherwise
    {
      const int P = next_pow2(n);
      unsigned long long* ukey = reinterpret_cast<unsigned long long*>(sm.khi);
      bool fits = true;
      for (int i = tid; i < P; i += kThreads) {
        unsigned long long u = ~0ull;
        if (i < n)
          fits = pack_time_key(a.submit_sec ? a.submit_sec[o + i] : 0, a.submit_nsec ? a.submit_nsec[o + i] : 0, &u) && fits;
        ukey[i] = u;
      }
      if (__syncthreads_and(fits)) {
        block_sort_kv(ukey, sm.perm, P);
      } else {
        for (int i = tid; i < n; i += kThreads) {
          sm.khi[i] = a.submit_sec ? a.submit_sec[o + i] : 0;
          sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
        }
        __syncthreads();
        block_argsort(sm.khi, sm.klo, sm.perm, n, P);
      }
    }
    // 2. trim, 3-4. DP rows
    const int J = min(n, K);
    const int KT = (K + kCells - 1) / kCells;
    float* gs = sc.srows + (size_t)rr * (size_t)sc.Jcap * (size_t)SW;
    FfdlRow* grows = sc.rows + (size_t)rr * (size_t)sc.Jcap;
    const int gtop = ffdl_setup_rows(a, sm, o, J, K, KT, SW, gs, grows);
    const bool sparse = gtop <= kCkptGmax;
    if (a.max_gmax > 0 && gtop > a.max_gmax) {  // some job can take more GPUs than the caller declared (max_max_proc)
      meta.st = VODA_ST_BOUND;
      if (tid == 0) sc.meta[rr] = meta;
      continue;
    }
    const float4* last = nullptr;
    ffdl_rows(a, sm, J, KT, SW, ost, sparse, sc.P + (size_t)rr * sc.slabFloats, sc.rowFloats, gs, grows, &last);
    const float PJK = cell(last, ost, K);
    meta.J = J;
    meta.sparse = sparse ? 1 : 0;
    if (PJK <= 0.0f) meta.st = VODA_ST_INFEASIBLE;  // :109-114 (NaN is not <= 0, as in Go)
    if (tid == 0) sc.meta[rr] = meta;
  }
}

// Kernel B: backtrack (:116-125) + deferred validateResult (:42).
// The walk is a chain of J dependent steps (row j-1 is read at k - g_j).  ONE WARP walks one
// round.  At row j it re-evaluates the reference's inner loop (:93-103) for the single cell (j, k),
//   best = -10000, sol = 0;  for g = 1 .. min(max_j, K, k):  p = S_j[g] + P[j-1][k-g];
//   if p > best { best = p; sol = g }
// 32 candidates at a time, one per lane: the P values are adjacent floats of row j-1, and the
// strict-'>' scan in increasing g is "the first lane that holds the maximum": an order-preserving
// integer key, one REDUX.MAX and one ballot.  P[j-1][x] comes from the stored row, its saturated
// constant, or row 0 = 0; in a checkpointed round from the band recomputed in shared memory
// (ffdl_walk_sparse), otherwise straight from the slab (ffdl_walk_dense).  Then the warp
// validates its round with coalesced reads.
constexpr int kBtThreads = 128;
constexpr int kBtAhead = 3;  // rows of lookahead for the L2 prefetch of P

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Order-preserving map float -> unsigned (the usual sign flip), after folding -0 into +0 so that
// equal floats get equal keys; NaN and "no such candidate" are mapped to key 0 by the caller.
__device__ __forceinline__ unsigned f32_key(float p) {
  const unsigned u = __float_as_uint(p + 0.0f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// The walk of a SPARSE round (only every kCkpt-th row of P in scratch): in segments of kCkpt
// rows.  For the segment of walk rows jTop .. c+1 (c the checkpoint row below) the warp loads
// the stretch of row c its walk can reach — the position only drifts down, by at most the row's
// g limit per step, so row r is needed on [k - D_r, k-1], D_r = the later rows' limits summed,
// 128 cells at most — recomputes rows c+1 .. jTop-1 on their (shrinking) stretches into shared
// memory with the forward pass's own recurrence (same additions, a maximum: any order gives the
// same value), and then takes the kCkpt walk steps from shared memory.  One dependent global
// load per segment instead of one per row.
// One segment of the checkpointed walk: the NR walk rows c+NR .. c+1 (NR = kCkpt except for the
// topmost segment of a round), from position k; returns the position below the segment.
template <int NR>
__device__ __forceinline__ int ffdl_segment(const float* slab, const float* gsrows, const int4* grows, int32_t* out,
                                            int SW, int W, int rowFloats, int lane, int slane, int c, int k,
                                            float (*band)[kBandRow], float (*ssm)[32]) {
  const unsigned full = 0xffffffffu;
  const int4 none = make_int4(0, 0, 0, 0);
  const unsigned keyFloor = f32_key(-10000.0f);
  int4 rec[NR];    // rec[i]: row c+1+i
  float sreg[NR];  // S_{c+1+i}[1 + lane]
  const int4* recp = grows + c;
  const float* sp = gsrows + c * SW + slane;
#pragma unroll
  for (int i = 0; i < NR; i++) {
    rec[i] = recp[i];
    sreg[i] = sp[i * SW];
  }
  const int4 recc = c >= 1 ? recp[-1] : none;  // the checkpoint row's klim and c
  // the next segment's records and S rows do not depend on the position: pull them into L2
  if (c >= kCkpt) {
    if (lane == 0) prefetch_l2(recp - kCkpt - 1);
    if (lane >= 1 && lane <= kCkpt) prefetch_l2(gsrows + (c - kCkpt + lane - 1) * SW);
    // ... and so is the stretch of the next checkpoint row just below the current position
    if (lane >= 8 && lane < 12 && c > kCkpt)
      prefetch_l2(slab + (c / kCkpt - 2) * rowFloats + max(k - 1 - 32 * (lane - 8), 0));
  }
#pragma unroll
  for (int i = 0; i < NR; i++) {
    // S as the recurrence uses it: a g the table does not hold reads 0 (a Go map miss)
    sreg[i] = lane + 1 <= W - 1 ? sreg[i] : 0.0f;
    ssm[i][lane] = sreg[i];
  }
  int lo[NR];  // lo[i]: lowest column of row c+i the walk can read
  {
    int D = 0;
#pragma unroll
    for (int i = NR - 1; i >= 0; i--) {
      D += max(0, rec[i].y);  // (0 past the segment)
      lo[i] = max(0, k - D);
    }
  }
  const int base = lo[0];
  const int width = k - base;  // <= kBand: columns base .. k-1
  // 1. the stored row c (or row 0 = 0)
  {
    const float* crow = slab + max(c / kCkpt - 1, 0) * rowFloats;  // P[c][x] at crow[x-1]
    const float cC = __int_as_float(recc.w);
    for (int idx = lane; idx < width; idx += 32) {
      const int x = base + idx;
      const float pv = crow[max(x, 1) - 1];
      band[0][kBandPad + idx] = c == 0 ? 0.0f : (x == 0 ? -10000.0f : (x >= recc.z ? cC : pv));
    }
  }
  __syncwarp();
  // 2. rows c+1 .. c+NR-1
#pragma unroll
  for (int i = 1; i < NR; i++) {
    {
      const int gmax = rec[i - 1].y, klim = rec[i - 1].z;
      const float cR = __int_as_float(rec[i - 1].w);
      const float* sv = ssm[i - 1];
      for (int idx0 = lo[i] - base; idx0 < width; idx0 += 32) {  // uniform trip count
        const int idx = idx0 + lane;
        const int x = base + idx;
        // P[r-1][x-g] sits at prow[-g]; a column below 0 falls into the -inf pad (the stretch
        // starts at column 0 whenever that can happen), so the k-g >= 0 test needs no branch
        const float* prow = band[i - 1] + kBandPad + min(idx, width);
        float best = -10000.0f;
        int g = 1;
        for (; g + 1 <= gmax; g += 2) best = fmaxf(fmaxf(best, sv[g - 1] + prow[-g]), sv[g] + prow[-g - 1]);
        if (g <= gmax) best = fmaxf(best, sv[g - 1] + prow[-g]);
        if (x == 0) best = -10000.0f;
        else if (x >= klim) best = cR;
        if (idx < width) band[i][kBandPad + idx] = best;
      }
      __syncwarp();
    }
  }
  // 3. the walk steps c+NR .. c+1 (row c+1+i reads band row i)
#pragma unroll
  for (int i = NR - 1; i >= 0; i--) {
    {
      const int gm = min(rec[i].y, k);  // candidates g = 1 .. gm (k-g >= 0); gm <= 32 in a sparse round
      const int g = 1 + lane;
      const float sg = sreg[i];
      const float pg = g <= gm ? band[i][kBandPad + k - g - base] : 0.0f;
      const float p = sg + pg;
      const unsigned key = (g <= gm && p == p) ? f32_key(p) : 0u;
      const unsigned top = __reduce_max_sync(full, key);
      const int sol = top > keyFloor ? __ffs((int)__ballot_sync(full, key == top)) : 0;
      if (lane == 0) out[rec[i].x] = sol;
      k -= sol;
    }
  }
  __syncwarp();
  return k;
}

__device__ __forceinline__ void ffdl_walk_sparse(const AllocArgs& a, const FfdlScratch& sc, int rr, int o, int J, int K,
                                                 int lane, float (*band)[kBandRow], float (*ssm)[32]) {
  const float* slab = sc.P + (size_t)rr * sc.slabFloats;
  const float* gsrows = sc.srows + (size_t)rr * (size_t)sc.Jcap * (size_t)sc.SW;
  const int4* grows = reinterpret_cast<const int4*>(sc.rows + (size_t)rr * (size_t)sc.Jcap);  // job, gmax, klim, c
  int32_t* out = a.out_gpus + o;
  const int SW = sc.SW;
  const int W = a.stride;
  const int rowFloats = (int)sc.rowFloats;
  const int slane = min(lane, SW - 1);
  int k = K;
  // the topmost segment takes the rows above the highest multiple of kCkpt; the others are full
  int c = ((J - 1) / kCkpt) * kCkpt;
  switch (J - c) {
    case 1: k = ffdl_segment<1>(slab, gsrows, grows, out, SW, W, rowFloats, lane, slane, c, k, band, ssm); break;
    case 2: k = ffdl_segment<2>(slab, gsrows, grows, out, SW, W, rowFloats, lane, slane, c, k, band, ssm); break;
    case 3: k = ffdl_segment<3>(slab, gsrows, grows, out, SW, W, rowFloats, lane, slane, c, k, band, ssm); break;
    default: k = ffdl_segment<kCkpt>(slab, gsrows, grows, out, SW, W, rowFloats, lane, slane, c, k, band, ssm); break;
  }
  for (c -= kCkpt; c >= 0; c -= kCkpt)
    k = ffdl_segment<kCkpt>(slab, gsrows, grows, out, SW, W, rowFloats, lane, slane, c, k, band, ssm);
}

// The walk of a round that stored every row of P (some job can take more than kCkptGmax GPUs):
// row by row, 32 candidates at a time; the row record and the S values of the next row are
// fetched one row ahead and the stretch of P reachable a few rows from now is prefetched into L2.
__device__ __forceinline__ void ffdl_walk_dense(const AllocArgs& a, const FfdlScratch& sc, int rr, int o, int J, int K,
                                                int lane) {
  const unsigned full = 0xffffffffu;
  const float* slab = sc.P + (size_t)rr * sc.slabFloats;
  const float* gsrows = sc.srows + (size_t)rr * (size_t)sc.Jcap * (size_t)sc.SW;
  const int4* grows = reinterpret_cast<const int4*>(sc.rows + (size_t)rr * (size_t)sc.Jcap);  // job, gmax, klim, c
  int32_t* out = a.out_gpus + o;
  const int SW = sc.SW;
  const int W = a.stride;
  const int rowFloats = (int)sc.rowFloats;
  const int slane = min(lane, SW - 1);
  const int4 none = make_int4(0, 0, 0, 0);
  const unsigned keyFloor = f32_key(-10000.0f);
  int k = K;
  int4 rc = grows[J - 1];                  // row j
  int4 rp = J >= 2 ? grows[J - 2] : none;  // row j-1: its klim and c describe the row that is read
  float s1 = gsrows[(J - 1) * SW + slane];  // S_j[1 + lane]
  for (int j = J; j >= 1; j--) {
    // one row ahead (none of this depends on the position)
    const int4 rpp = j >= 3 ? grows[j - 3] : none;
    const float s1next = j >= 2 ? gsrows[(j - 2) * SW + slane] : 0.0f;
    // the position can only drift down, by the g's chosen on the way
    if (lane < 4 && j - 2 - kBtAhead >= 0) prefetch_l2(slab + (j - 2 - kBtAhead) * rowFloats + max(k - 1 - 32 * lane, 0));
    const int gm = min(rc.y, k);                           // candidates g = 1 .. gm (k-g >= 0)
    const float* prow = slab + max(j - 2, 0) * rowFloats;  // P[j-1][x] at prow[x-1] (j >= 2)
    const int klp = rp.z;
    const float cp = __int_as_float(rp.w);
    unsigned bestKey = keyFloor;  // best = -10000
    int sol = 0;
    for (int g0 = 1; g0 <= gm; g0 += 32) {
      const int g = g0 + lane;
      const int x = k - g;
      // unconditional loads on clamped, always-valid addresses; the cases are sorted out below
      const float pv = prow[max(x, 1) - 1];
      const float sv = g0 == 1 ? s1 : gsrows[(j - 1) * SW + min(g, SW) - 1];
      const float sg = g <= W - 1 ? sv : 0.0f;  // a g the table does not hold: Go map miss
      float pg = x <= 0 ? -10000.0f : (x >= klp ? cp : pv);
      if (j == 1) pg = 0.0f;  // row 0
      const float p = sg + pg;
      // strict '>' in increasing g == the first lane holding the maximum; a NaN never wins
      const unsigned key = (g <= gm && p == p) ? f32_key(p) : 0u;
      const unsigned top = __reduce_max_sync(full, key);
      if (top > bestKey) {
        bestKey = top;
        sol = g0 + __ffs((int)__ballot_sync(full, key == top)) - 1;
      }
    }
    if (lane == 0) out[rc.x] = sol;
    k -= sol;
    rc = rp;
    rp = rpp;
    s1 = s1next;
  }
}

__global__ void __launch_bounds__(kBtThreads, 8)
    ffdl_backtrack_kernel(AllocArgs a, int r0, int count, FfdlScratch sc) {
  __shared__ float band[kBtThreads / 32][kCkpt][kBandRow];
  __shared__ float ssm[kBtThreads / 32][kCkpt][32];
  for (int i = threadIdx.x; i < (kBtThreads / 32) * kCkpt * kBandPad; i += kBtThreads)  // the pads, written once
    band[i / (kCkpt * kBandPad)][(i / kBandPad) % kCkpt][i % kBandPad] = VODA_NEG_INF;
  __syncthreads();
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rr = blockIdx.x * (kBtThreads / 32) + warp;  // a warp per round
  if (rr >= count) return;
  const int r = r0 + rr;
  const int o = a.job_off[r] - a.job_base;
  const int n = a.job_off[r + 1] - a.job_off[r];
  const FfdlMeta meta = sc.meta[rr];
  // ---- phase 1: the walk
  if (meta.st == VODA_ST_OK && meta.J > 0) {
    if (meta.sparse) ffdl_walk_sparse(a, sc, rr, o, meta.J, meta.K, lane, band[warp], ssm[warp]);
    else ffdl_walk_dense(a, sc, rr, o, meta.J, meta.K, lane);
  }
  __syncwarp(full);
  // ---- phase 2: validateResult (utils.go:18-42); a panic it raises replaces the one in flight
  if (n == 0 || meta.st == VODA_ST_BOUND) {
    if (lane == 0) a.out_status[r] = n == 0 ? VODA_ST_OK : VODA_ST_BOUND;
    return;
  }
  int worst = 99;
  long long sum = 0;
  for (int base = 0; base < n; base += 32 * 8) {  // 24 loads in flight per lane, then the checks
    int v[8], mn[8], mx[8];
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const int i = base + 32 * u + lane;
      const bool ok = i < n;
      v[u] = ok ? a.out_gpus[o + i] : 0;
      mn[u] = ok ? a.min_proc[o + i] : 0;
      mx[u] = ok ? a.max_proc[o + i] : 0x7fffffff;
    }
#pragma unroll
    for (int u = 0; u < 8; u++) {
      int st = 99;
      if (v[u] < 0) st = VODA_ST_NEGATIVE;
      else if (v[u] > 0 && v[u] < mn[u]) st = VODA_ST_LESS_THAN_MIN;
      else if (v[u] > mx[u]) st = VODA_ST_EXCEEDED_MAX;
      worst = min(worst, st);
      sum += v[u];
    }
  }
  for (int d = 16; d > 0; d >>= 1) {
    worst = min(worst, __shfl_xor_sync(full, worst, d));
    sum += __shfl_xor_sync(full, sum, d);
  }
  if (lane == 0) {
    int vv = worst != 99 ? worst : (sum > (long long)meta.K ? VODA_ST_EXCEEDED_TOTAL : VODA_ST_OK);
    a.out_status[r] = vv ? vv : meta.st;
  }
}

}  // namespace

static void ffdl_shape(const AllocArgs& a, int* KTcap, int* SW, int* Jp, size_t* smem) {
  const int Kmax = a.max_total > 0 ? a.max_total : 1;
  *KTcap = (Kmax + kCells - 1) / kCells;  // whole tiles
  int sw = a.stride - 1 < Kmax ? a.stride - 1 : Kmax;
  if (sw < 1) sw = 1;
  *SW = (sw + 3) & ~3;
  *Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  *smem = sizeof(float4) * (size_t)row_groups(*SW, *KTcap) + sizeof(float) * 4 * (size_t)(*SW) +
          (sizeof(int64_t) + 3 * sizeof(int32_t)) * (size_t)(*Jp);
}

static size_t ffdl_jcap(const AllocArgs& a) {
  const int j = a.max_jobs < a.max_total ? a.max_jobs : a.max_total;
  return (size_t)(j > 0 ? j : 1);
}

// Rows of P a round's slab holds.  When the caller declared that no job can take more than
// kCkptGmax GPUs every round is a sparse one: only the checkpoint rows exist.
static bool ffdl_sparse_only(const AllocArgs& a) { return a.max_gmax > 0 && a.max_gmax <= kCkptGmax; }
static size_t ffdl_slab_rows(const AllocArgs& a) {
  const size_t Jcap = ffdl_jcap(a);
  return ffdl_sparse_only(a) ? (Jcap - 1) / kCkpt + 1 : Jcap;
}

// Scratch one round needs: its P slab, the sorted S rows, the row records and the meta record
// (every part a multiple of 16 bytes).
size_t ffdl_scratch_per_round(const AllocArgs& a) {
  int KTcap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &KTcap, &SW, &Jp, &smem);
  const size_t Jcap = ffdl_jcap(a);
  return ffdl_slab_rows(a) * (size_t)KTcap * 32 + Jcap * (size_t)SW * 4 + Jcap * sizeof(FfdlRow) + sizeof(FfdlMeta);
}

size_t ffdl_smem_bytes(const AllocArgs& a) {
  int KTcap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &KTcap, &SW, &Jp, &smem);
  return smem;
}

// Rounds [r0, r0+count) of `a`: DP kernel then backtrack kernel.  `scratch` must hold
// count * ffdl_scratch_per_round(a) bytes (256-byte aligned).
cudaError_t launch_ffdl(const AllocArgs& a, int r0, int count, int sms, void* scratch, cudaStream_t stream) {
  int KTcap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &KTcap, &SW, &Jp, &smem);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(ffdl_dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const size_t Jcap = ffdl_jcap(a);
  FfdlScratch sc;
  unsigned char* base = static_cast<unsigned char*>(scratch);
  sc.rowFloats = (size_t)KTcap * 8;
  sc.slabFloats = ffdl_slab_rows(a) * sc.rowFloats;
  sc.sparseOnly = ffdl_sparse_only(a) ? 1 : 0;
  sc.Jcap = (int)Jcap;
  sc.SW = SW;
  size_t off = 0;
  sc.P = reinterpret_cast<float*>(base + off);        off += (size_t)count * sc.slabFloats * 4;
  sc.srows = reinterpret_cast<float*>(base + off);    off += (size_t)count * Jcap * (size_t)SW * 4;
  sc.rows = reinterpret_cast<FfdlRow*>(base + off);   off += (size_t)count * Jcap * sizeof(FfdlRow);
  sc.meta = reinterpret_cast<FfdlMeta*>(base + off);
  int perSm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, ffdl_dp_kernel, kThreads, smem);
  if (e != cudaSuccess) return e;
  if (perSm < 1) perSm = 1;
  int grid = sms * perSm;
  if (count < grid) grid = count;
  ffdl_dp_kernel<<<grid, kThreads, smem, stream>>>(a, r0, count, sc, KTcap, SW, Jp);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int warps = kBtThreads / 32;  // a warp per round
  ffdl_backtrack_kernel<<<(count + warps - 1) / warps, kBtThreads, 0, stream>>>(a, r0, count, sc);
  return cudaGetLastError();
}

}  // namespace voda
