// ffdl.cu — FfDLOptimizer.Schedule (pkg/algorithm/ffdl_optimizer.go:32-130) for a batch of
// rounds, hand-written for sm_100a.
//
// One CTA owns one round at a time (persistent grid striding over rounds).  Per round:
//   1. stable sort of the jobs by SubmitTime (:45-47)            -> block bitonic argsort
//   2. trim to the first J = min(n, K) jobs (:54-65)
//   3. DP  P[j][k] = max_g S_j[g] + P[j-1][k-g], g in [1, max_j], k-g >= 0, strict '>' so the
//      smallest g wins ties, rows initialised to -10000 (:78-105) -> ONE ROW PER CTA STEP:
//      rows j-1 and j live in shared memory (float4 groups), each thread owns 4 adjacent
//      cells and slides an 8-register window of row j-1 over g, so a block of 4 g-steps x 4
//      cells costs one LDS.128 of P and one LDS.128 (broadcast) of S for 64 add/compare/select
//      groups.  The argmax is per cell; no cross-thread reduction is needed.
//   4. SOL[j][k] (the chosen g) is streamed to the round's slab of global scratch (uint8 when
//      the round's largest g fits, else uint16); it is written once and only J cells are read
//      back by the backtrack (:116-125), which is a SECOND kernel (one warp per round): the
//      chase is a chain of J dependent loads, and inside the DP kernel it would leave the CTA
//      idle for a fifth of its life; side by side over thousands of rounds it costs nothing.
//   5. P[J][K] <= 0 -> "infeasible" (:109-114); deferred validateResult (:42).
//
// float32 evaluation order: p = S_j[g] + P[j-1][k-g] is a single rounded add (-fmad=false;
// there is nothing to contract anyway).  Cells with k-g < 0 are fed -inf, which can never win
// the strict compare (and -inf + +inf = NaN never does either), which reproduces the
// reference's `if k-g >= 0` guard without a branch.
#include "common.cuh"
#include "kernels.h"

namespace voda {

namespace {

#ifndef VODA_FFDL_THREADS
#define VODA_FFDL_THREADS 64
#endif
#ifndef VODA_FFDL_CELLS
#define VODA_FFDL_CELLS 8
#endif
#ifndef VODA_FFDL_MINBLOCKS
#define VODA_FFDL_MINBLOCKS 12
#endif
// 64 threads x 8 cells (two tiles per thread at K = 1024) measured best on B200 among
// {32,64,128} x {4,8,16,32}; see profiles/r1_ffdl_shape_sweep.txt
constexpr int kThreads = VODA_FFDL_THREADS;

struct FfdlSmem {
  float4* rowA;      // [G4]  row buffers, group q = P[4q-3 .. 4q]
  float4* rowB;      // [G4]
  float* srow0;      // [SW]  S_j[1..] of even rows, -inf beyond the row's g limit
  float* srow1;      // [SW]  ... of odd rows
  int64_t* khi;      // [Jp]
  int32_t* klo;      // [Jp]
  int* perm;         // [Jp]
  int32_t* smax;     // [Jp]  max_proc in sorted order
  int32_t* scum;     // [Jp]  min(K, sum of clamp(max_proc, 0, K) over the first j+1 sorted jobs)
  int32_t* slo;      // [Jp]  max(0, K - sum of clamp(max_proc, 0, K) over the sorted jobs after j)
  float* srmax;      // [Jp]  max over the row's usable g of S_j[g] (-inf if none)
  long long* red;    // [2]
  int* misc;         // [4]
};

__device__ __forceinline__ float4 ld_group(const float4* row, int q) {
  // groups left of the row (k-g < 0) read as -inf
  if (q < 0) return make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF);
  return row[q];
}

// One (cell, g) step of  p = S + P[j-1][k-g]; if p > best { best = p; SOL = g }.
// Instruction mix matters here (the loop is bound by the ALU pipe): the running maximum is an
// FMNMX, the "did it win" bit an FSET (1.0f / 0.0f), and the winning g is NOT tracked with a
// select per step: the win bits are shifted into a float accumulator with an FFMA
// (acc = acc/2 + bit, i.e. on the FMA pipe; exact for up to 23 steps), so the exponent of acc
// says how many steps ago the LAST win happened, and it is decoded once per flush.
// fmaxf(best, p) equals the reference's strict-'>' select for every non-NaN input (equal
// values are interchangeable; a NaN p never replaces best in either form).
#define VODA_CELL(c, s)                              \
  {                                                  \
    const float p = sv[s] + v[3 + (c) - (s)];        \
    const float t = (p > best[c]) ? 1.0f : 0.0f;     \
    best[c] = fmaxf(best[c], p);                     \
    acc[c] = fmaf(acc[c], 0.5f, t);                  \
  }

constexpr int kCells = VODA_FFDL_CELLS;  // adjacent cells per thread
constexpr int kGroups = kCells / 4 + 1;  // float4 groups of row j-1 in the sliding window
constexpr int kFlushBlocks = 5;          // 5 blocks x 4 g-steps = 20 win bits per accumulator flush

// One DP row for column tile `tq` (cells k = kCells*tq+1 .. kCells*tq+kCells).
// Row storage: group q = P[4q-3 .. 4q].  For the block of g = 4m+1 .. 4m+4 the cells need
// P[base-3 .. base+kCells-1], base = kCells*tq - 4m, i.e. groups q0 .. q0+kCells/4 with
// q0 = (kCells/4)*tq - m; the next block slides the window down by one group (one LDS.128).
template <typename SolT>
__device__ __forceinline__ void ffdl_row_tile(const float4* __restrict__ prev, float4* __restrict__ cur,
                                              const float* __restrict__ srow, int tq, int nblk,
                                              int gslow0, int gslow1, SolT* __restrict__ solrow) {
  float best[kCells];
  float acc[kCells];
  int bg[kCells];
#pragma unroll
  for (int c = 0; c < kCells; c++) {
    best[c] = -10000.0f;
    acc[c] = 0.0f;
    bg[c] = 0;
  }
  const int qtop = (kCells / 4) * tq;
  float4 grp[kGroups];
#pragma unroll
  for (int i = 0; i < kGroups; i++) grp[i] = ld_group(prev, qtop + i);
  const float4* s4 = reinterpret_cast<const float4*>(srow);
  int sinceFlush = 0;
  for (int m = 0; m < nblk; m++) {
    float v[4 * kGroups];
#pragma unroll
    for (int i = 0; i < kGroups; i++) {
      v[4 * i + 0] = grp[i].x;
      v[4 * i + 1] = grp[i].y;
      v[4 * i + 2] = grp[i].z;
      v[4 * i + 3] = grp[i].w;
    }
    const float4 sq = s4[m];
    const float sv[4] = {sq.x, sq.y, sq.z, sq.w};
#pragma unroll
    for (int s = 0; s < 4; s++) {
#pragma unroll
      for (int c = 0; c < kCells; c++) VODA_CELL(c, s)
    }
#pragma unroll
    for (int i = kGroups - 1; i > 0; i--) grp[i] = grp[i - 1];
    grp[0] = ld_group(prev, qtop - m - 1);
    if (++sinceFlush == kFlushBlocks || m == nblk - 1) {
      // decode the last win of the flushed span: acc in [2^-d, 2^(1-d)) <=> it happened d
      // steps before the last one of the span, i.e. at g = gEnd - d
      const int gEnd = 4 * m + 4;
#pragma unroll
      for (int c = 0; c < kCells; c++) {
        const int e = __float_as_int(acc[c]) >> 23;  // biased exponent (acc >= 0)
        if (e) bg[c] = gEnd - 127 + e;
        acc[c] = 0.0f;
      }
      sinceFlush = 0;
    }
  }
  // g beyond the dense speedup table: the Go map misses, S = 0 (rare: max_j >= W)
  if (gslow1 >= gslow0) {
    const float* pf = reinterpret_cast<const float*>(prev);
#pragma unroll
    for (int c = 0; c < kCells; c++) {
      const int k = kCells * tq + 1 + c;
      for (int g = gslow0; g <= gslow1 && g <= k; g++) {
        float p = 0.0f + pf[k - g + 3];
        if (p > best[c]) {
          best[c] = p;
          bg[c] = g;
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kCells / 4; i++)
    cur[qtop + 1 + i] = make_float4(best[4 * i], best[4 * i + 1], best[4 * i + 2], best[4 * i + 3]);
  // SOL entries of the tile, packed and stored with the widest aligned vector store
  constexpr int kWords = kCells * (int)sizeof(SolT) / 4;
  uint32_t w[kWords];
  if (sizeof(SolT) == 1) {
#pragma unroll
    for (int i = 0; i < kWords; i++)
      w[i] = (uint32_t)bg[4 * i] | ((uint32_t)bg[4 * i + 1] << 8) | ((uint32_t)bg[4 * i + 2] << 16) |
             ((uint32_t)bg[4 * i + 3] << 24);
  } else {
#pragma unroll
    for (int i = 0; i < kWords; i++) w[i] = (uint32_t)bg[2 * i] | ((uint32_t)bg[2 * i + 1] << 16);
  }
  uint32_t* dst = reinterpret_cast<uint32_t*>(solrow) + (size_t)tq * kWords;
  if (kWords >= 4) {
#pragma unroll
    for (int i = 0; i < kWords; i += 4) *reinterpret_cast<uint4*>(dst + i) = make_uint4(w[i], w[i + 1], w[i + 2], w[i + 3]);
  } else {
    *reinterpret_cast<uint2*>(dst) = make_uint2(w[0], w[1]);
  }
}

// All DP rows of one round.  Returns with the last row in `*last` (shared memory).
template <typename SolT>
__device__ void ffdl_rows(const AllocArgs& a, const FfdlSmem& sm, int o, int J, int K, int KT,
                          int SW, size_t solStride, unsigned char* solSlab, const float4** last) {
  const int tid = threadIdx.x;
  const int W = a.stride;
  float4* prev = sm.rowA;
  float4* cur = sm.rowB;
  // row 0: P[0][k] = 0 for every k (:80-82)
  for (int q = tid; q <= (kCells / 4) * KT; q += kThreads)
    prev[q] = q == 0 ? make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, 0.0f) : make_float4(0.f, 0.f, 0.f, 0.f);
  // stage S of row 1
  auto s_value = [&](int j, int g) -> float {  // value of srow slot g-1 for sorted job j (1-based)
    const int mx = sm.smax[j - 1];
    const int gt = min(min(mx, K), W - 1);  // last g served from the table loop
    if (g > gt) return VODA_NEG_INF;
    const int job = o + sm.perm[j - 1];
    const size_t row = a.sp_off ? (size_t)a.sp_off[job] : (size_t)job * (size_t)W;
    return a.speedup[row + (size_t)g];
  };
  if (J >= 1)
    for (int g = tid + 1; g <= SW; g += kThreads) sm.srow1[g - 1] = s_value(1, g);
  __syncthreads();
  // Saturation: once k >= klim_j = sum of the first j jobs' max sizes, the GPU budget no longer
  // binds, so P[j][k] is one constant c_j (and SOL[j][k] one constant) for every k >= klim_j —
  // bit for bit, because each candidate S_j[g] + P[j-1][k-g] then adds the same two numbers.
  // Tiles wholly beyond klim_j are filled with c_j instead of being computed, their SOL is
  // not stored, and the backtrack reads SOL[j][min(k, klim_j)].
  float cprev = 0.0f;  // c_0: P[0][k] = 0
  for (int j = 1; j <= J; j++) {
    const float* srow = (j & 1) ? sm.srow1 : sm.srow0;
    float* snext = (j & 1) ? sm.srow0 : sm.srow1;
    // software-pipelined prefetch of the next row's S values (global -> register -> smem)
    float pre[2] = {VODA_NEG_INF, VODA_NEG_INF};
    if (j < J) {
#pragma unroll
      for (int u = 0; u < 2; u++) {
        int g = tid + 1 + u * kThreads;
        if (g <= SW) pre[u] = s_value(j + 1, g);
      }
    }
    const int mx = sm.smax[j - 1];
    const int gmax = min(mx, K);
    const int gt = min(gmax, W - 1);
    const int nblk = gt > 0 ? (gt + 3) >> 2 : 0;
    SolT* solrow = reinterpret_cast<SolT*>(solSlab + (size_t)(j - 1) * solStride);
    const int klim = sm.scum[j - 1];
    // c_j = max(-10000, max_g (S_j[g] + c_{j-1})) = max(-10000, (max_g S_j[g]) + c_{j-1}): rounding
    // is monotone, so the maximum commutes with the addition exactly
    const float ccur = fmaxf(-10000.0f, sm.srmax[j - 1] + cprev);
    // The other side of the budget: the backtrack starts at k = K and gives up at most max_i
    // per row, so row j is only ever consulted at k >= K - (sum of the later jobs' max sizes),
    // or at its saturation point klim_j when that is lower; klo_j is the smaller of the two,
    // and a cell at k >= klo_j reads row j-1 at k-g >= klo_j - max_j >= klo_{j-1}.  Tiles
    // wholly below klo_j feed neither P[J][K] nor any SOL entry the backtrack can read: they
    // are not computed (what the row buffer holds there is never read by a cell that matters).
    const int klo = sm.slo[j - 1];
    for (int tq = tid; tq < KT; tq += kThreads) {
      if (kCells * tq + kCells < klo) continue;
      if (kCells * tq + 1 > klim) {
        const float4 c4 = make_float4(ccur, ccur, ccur, ccur);
#pragma unroll
        for (int i = 0; i < kCells / 4; i++) cur[(kCells / 4) * tq + 1 + i] = c4;
      } else {
        ffdl_row_tile<SolT>(prev, cur, srow, tq, nblk, max(W, 1), gmax, solrow);
      }
    }
    cprev = ccur;
    if (tid == 0) cur[0] = make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, -10000.0f);  // P[j][0]
    if (j < J) {
#pragma unroll
      for (int u = 0; u < 2; u++) {
        int g = tid + 1 + u * kThreads;
        if (g <= SW) snext[g - 1] = pre[u];
      }
      for (int g = tid + 1 + 2 * kThreads; g <= SW; g += kThreads) snext[g - 1] = s_value(j + 1, g);
    }
    __syncthreads();
    float4* t = prev;
    prev = cur;
    cur = t;
  }
  *last = prev;
}

// Per-round records the DP kernel leaves for the backtrack kernel.
struct FfdlMeta {
  int32_t J;     // rows of the DP (jobs kept after the trim)
  int32_t st;    // VODA_ST_OK, or the panic raised before / by the DP
  int32_t elem;  // bytes per SOL entry (1 or 2)
  int32_t K;
};

struct FfdlScratch {
  FfdlMeta* meta;      // [count]
  int32_t* perm;       // [count][Jcap] sorted position -> job index within the round
  int32_t* klim;       // [count][Jcap] saturation point of each DP row
  unsigned char* sol;  // [count][Jcap][solStride]
  size_t slabBytes;    // Jcap * solStrideMax
  int Jcap;
};

// Kernel A: sort + DP rows.  One CTA per round at a time; SOL / perm / meta go to the round's
// own slot of the scratch so that the CTA can move on to its next round immediately.
__global__ void __launch_bounds__(kThreads, VODA_FFDL_MINBLOCKS) ffdl_dp_kernel(AllocArgs a, int r0, int count, FfdlScratch sc,
                                                            int K4cap, int SW, int Jp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FfdlSmem sm;
  {
    unsigned char* p = smem_raw;
    sm.rowA = reinterpret_cast<float4*>(p); p += sizeof(float4) * (size_t)(K4cap + 1);
    sm.rowB = reinterpret_cast<float4*>(p); p += sizeof(float4) * (size_t)(K4cap + 1);
    sm.srow0 = reinterpret_cast<float*>(p); p += sizeof(float) * (size_t)SW;
    sm.srow1 = reinterpret_cast<float*>(p); p += sizeof(float) * (size_t)SW;
    sm.khi = reinterpret_cast<int64_t*>(p); p += sizeof(int64_t) * (size_t)Jp;
    sm.red = reinterpret_cast<long long*>(p); p += sizeof(long long) * 2;
    sm.klo = reinterpret_cast<int32_t*>(p); p += sizeof(int32_t) * (size_t)Jp;
    sm.perm = reinterpret_cast<int*>(p); p += sizeof(int) * (size_t)Jp;
    sm.smax = reinterpret_cast<int32_t*>(p); p += sizeof(int32_t) * (size_t)Jp;
    sm.scum = reinterpret_cast<int32_t*>(p); p += sizeof(int32_t) * (size_t)Jp;
    sm.slo = reinterpret_cast<int32_t*>(p); p += sizeof(int32_t) * (size_t)Jp;
    sm.srmax = reinterpret_cast<float*>(p); p += sizeof(float) * (size_t)Jp;
    sm.misc = reinterpret_cast<int*>(p);
  }
  const int tid = threadIdx.x;

  for (int rr = blockIdx.x; rr < count; rr += gridDim.x) {
    const int r = r0 + rr;
    const int o = a.job_off[r] - a.job_base;
    const int n = a.job_off[r + 1] - a.job_off[r];
    const int K = a.total_gpu[r];
    FfdlMeta meta{0, VODA_ST_OK, 1, K};
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
    // 1. stable sort by SubmitTime
    for (int i = tid; i < n; i += kThreads) {
      sm.khi[i] = a.submit_sec ? a.submit_sec[o + i] : 0;
      sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
    }
    __syncthreads();
    block_argsort(sm.khi, sm.klo, sm.perm, n, next_pow2(n));
    // 2. trim
    const int J = min(n, K);
    int lmax = 0;
    int32_t* gperm = sc.perm + (size_t)rr * sc.Jcap;
    for (int j = tid; j < J; j += kThreads) {
      const int pj = sm.perm[j];
      gperm[j] = pj;
      const int mx = a.max_proc[o + pj];
      sm.smax[j] = mx;
      lmax = max(lmax, mx);
      // largest speedup value the row can use (g beyond the table reads 0: a Go map miss)
      const int gmx = min(mx, K);
      float rm = gmx >= max(a.stride, 1) ? 0.0f : VODA_NEG_INF;
      const float* sj = a.speedup + (a.sp_off ? (size_t)a.sp_off[o + pj] : (size_t)(o + pj) * (size_t)a.stride);
      for (int g = 1; g <= min(gmx, a.stride - 1); g++) rm = fmaxf(rm, sj[g]);
      sm.srmax[j] = rm;
    }
    if (tid == 0) sm.misc[0] = 0;
    __syncthreads();
    for (int d = 16; d > 0; d >>= 1) lmax = max(lmax, __shfl_xor_sync(0xffffffffu, lmax, d));
    if ((tid & 31) == 0) atomicMax(&sm.misc[0], lmax);
    if (tid == 0) {  // running sum of the usable sizes, clipped to K (a short serial scan)
      long long acc = 0;
      int32_t* gklim = sc.klim + (size_t)rr * sc.Jcap;
      for (int j = 0; j < J; j++) {
        acc += max(0, min(sm.smax[j], K));
        const int kl = (int)min(acc, (long long)K);
        sm.scum[j] = kl;
        gklim[j] = kl;
      }
      acc = 0;
      for (int j = J - 1; j >= 0; j--) {
        // (the backtrack reads SOL[j] at min(k, klim_j), so the saturation point itself must
        // stay inside the computed band when it lies below the cone)
        sm.slo[j] = min((int)max(0LL, (long long)K - acc), sm.scum[j]);
        acc += max(0, min(sm.smax[j], K));
      }
    }
    __syncthreads();
    const int gmaxRound = min(sm.misc[0], K);
    // 3-4. DP rows
    const int KT = (K + kCells - 1) / kCells;
    unsigned char* slab = sc.sol + (size_t)rr * sc.slabBytes;
    const float4* last = nullptr;
    if (gmaxRound <= 255) {
      ffdl_rows<uint8_t>(a, sm, o, J, K, KT, SW, (size_t)K4cap * 4, slab, &last);
      meta.elem = 1;
    } else {
      ffdl_rows<uint16_t>(a, sm, o, J, K, KT, SW, (size_t)K4cap * 8, slab, &last);
      meta.elem = 2;
    }
    const float PJK = reinterpret_cast<const float*>(last)[K + 3];
    meta.J = J;
    if (PJK <= 0.0f) meta.st = VODA_ST_INFEASIBLE;  // :109-114 (NaN is not <= 0, as in Go)
    if (tid == 0) sc.meta[rr] = meta;
  }
}

// Kernel B: backtrack (:116-125) + deferred validateResult (:42).
// The chase through SOL is a chain of J dependent loads (row j-1 is read at k - g_j), so it is
// all latency.  Each LANE chases one round (32 independent chains per warp, every round of the
// batch in flight at once), eight rows at a time: the row constants (perm, klim) of the block
// are loaded up front, and the SOL lines the NEXT block will touch are prefetched into L2 —
// the position can only drift down by the g's chosen on the way, a few bytes per row — so the
// dependent loads hit L2 instead of HBM.  Then the warp validates its 32 rounds one after the
// other with coalesced reads.
constexpr int kBtThreads = 128;
constexpr int kBtBlock = 8;  // rows per chase block

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__global__ void __launch_bounds__(kBtThreads) ffdl_backtrack_kernel(AllocArgs a, int r0, int count, FfdlScratch sc,
                                                                    int K4cap) {
  const int lane = threadIdx.x & 31;
  const int rr0 = (blockIdx.x * (kBtThreads / 32) + (threadIdx.x >> 5)) * 32;
  if (rr0 >= count) return;
  // ---- phase 1: one chase per lane
  {
    const int rr = rr0 + lane;
    if (rr < count) {
      const int r = r0 + rr;
      const int o = a.job_off[r] - a.job_base;
      const FfdlMeta meta = sc.meta[rr];
      if (meta.st == VODA_ST_OK && meta.J > 0) {
        const unsigned char* slab = sc.sol + (size_t)rr * sc.slabBytes;
        const int32_t* gperm = sc.perm + (size_t)rr * sc.Jcap;
        const int32_t* gklim = sc.klim + (size_t)rr * sc.Jcap;
        const size_t stride = (size_t)K4cap * 4 * (size_t)meta.elem;
        const int es = meta.elem;  // bytes per SOL entry
        int k = meta.K;
        auto prefetch_rows = [&](int jtop) {  // rows jtop, jtop-1, ... (kBtBlock of them), around position k
          for (int d = 0; d < kBtBlock && jtop - d > 0; d++) {
            const int j = jtop - d;
            const int kk = min(k, gklim[j - 1]);
            if (kk > 0) {
              const unsigned char* p = slab + (size_t)(j - 1) * stride + (size_t)(kk - 1) * es;
              prefetch_l2(p);
              if (kk * es > 96) prefetch_l2(p - 96);
            }
          }
        };
        prefetch_rows(meta.J);
        for (int jb = meta.J; jb > 0; jb -= kBtBlock) {
          int pj[kBtBlock], kl[kBtBlock];
#pragma unroll
          for (int d = 0; d < kBtBlock; d++) {
            const int j = jb - d;
            pj[d] = j > 0 ? gperm[j - 1] : 0;
            kl[d] = j > 0 ? gklim[j - 1] : 0;
          }
          prefetch_rows(jb - kBtBlock);
#pragma unroll
          for (int d = 0; d < kBtBlock; d++) {
            const int j = jb - d;
            if (j <= 0) break;
            int g = 0;  // SOL[j][0] == 0
            const int kk = min(k, kl[d]);  // SOL[j][k] is constant from the row's saturation point on
            if (kk > 0) {
              const unsigned char* row = slab + (size_t)(j - 1) * stride;
              g = es == 1 ? (int)row[kk - 1] : (int)reinterpret_cast<const uint16_t*>(row)[kk - 1];
            }
            a.out_gpus[o + pj[d]] = g;
            k -= g;
          }
        }
      }
    }
  }
  __syncwarp();
  // ---- phase 2: validateResult (utils.go:18-42) per round; a panic it raises replaces the one in flight
  for (int q = 0; q < 32 && rr0 + q < count; q++) {
    const int rr = rr0 + q;
    const int r = r0 + rr;
    const int o = a.job_off[r] - a.job_base;
    const int n = a.job_off[r + 1] - a.job_off[r];
    const FfdlMeta meta = sc.meta[rr];
    if (n == 0 || meta.st == VODA_ST_BOUND) {
      if (lane == 0) a.out_status[r] = n == 0 ? VODA_ST_OK : VODA_ST_BOUND;
      continue;
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
      worst = min(worst, __shfl_xor_sync(0xffffffffu, worst, d));
      sum += __shfl_xor_sync(0xffffffffu, sum, d);
    }
    if (lane == 0) {
      int vv = worst != 99 ? worst : (sum > (long long)meta.K ? VODA_ST_EXCEEDED_TOTAL : VODA_ST_OK);
      a.out_status[r] = vv ? vv : meta.st;
    }
  }
}

}  // namespace

static void ffdl_shape(const AllocArgs& a, int* K4cap, int* SW, int* Jp, size_t* smem) {
  const int Kmax = a.max_total > 0 ? a.max_total : 1;
  *K4cap = (kCells / 4) * ((Kmax + kCells - 1) / kCells);  // row groups of 4 cells, whole tiles
  int sw = a.stride - 1 < Kmax ? a.stride - 1 : Kmax;
  if (sw < 1) sw = 1;
  *SW = (sw + 3) & ~3;
  *Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  *smem = sizeof(float4) * 2 * (size_t)(*K4cap + 1) + sizeof(float) * 2 * (size_t)(*SW) +
          (sizeof(int64_t) + 6 * sizeof(int32_t)) * (size_t)(*Jp) + 2 * sizeof(long long) + 4 * sizeof(int);
}

static size_t ffdl_jcap(const AllocArgs& a) {
  const int j = a.max_jobs < a.max_total ? a.max_jobs : a.max_total;
  return (size_t)(j > 0 ? j : 1);
}

// Scratch one round needs: its SOL slab (sized for 16-bit entries), perm row and meta record.
size_t ffdl_scratch_per_round(const AllocArgs& a) {
  int K4cap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &K4cap, &SW, &Jp, &smem);
  const size_t Jcap = ffdl_jcap(a);
  return Jcap * (size_t)K4cap * 8 + 2 * ((Jcap * 4 + 15) & ~(size_t)15) + sizeof(FfdlMeta);
}

size_t ffdl_smem_bytes(const AllocArgs& a) {
  int K4cap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &K4cap, &SW, &Jp, &smem);
  return smem;
}

// Rounds [r0, r0+count) of `a`: DP kernel then backtrack kernel.  `scratch` must hold
// count * ffdl_scratch_per_round(a) bytes (256-byte aligned).
cudaError_t launch_ffdl(const AllocArgs& a, int r0, int count, int sms, void* scratch, cudaStream_t stream) {
  int K4cap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &K4cap, &SW, &Jp, &smem);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(ffdl_dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const size_t Jcap = ffdl_jcap(a);
  const size_t permRow = (Jcap * 4 + 15) & ~(size_t)15;
  FfdlScratch sc;
  unsigned char* base = static_cast<unsigned char*>(scratch);
  sc.slabBytes = Jcap * (size_t)K4cap * 8;
  sc.Jcap = (int)(permRow / 4);
  sc.sol = base;
  sc.perm = reinterpret_cast<int32_t*>(base + (size_t)count * sc.slabBytes);
  sc.klim = reinterpret_cast<int32_t*>(base + (size_t)count * (sc.slabBytes + permRow));
  sc.meta = reinterpret_cast<FfdlMeta*>(base + (size_t)count * (sc.slabBytes + 2 * permRow));
  int perSm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, ffdl_dp_kernel, kThreads, smem);
  if (e != cudaSuccess) return e;
  if (perSm < 1) perSm = 1;
  int grid = sms * perSm;
  if (count < grid) grid = count;
  ffdl_dp_kernel<<<grid, kThreads, smem, stream>>>(a, r0, count, sc, K4cap, SW, Jp);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int perBlock = (kBtThreads / 32) * 32;  // rounds per block: one lane per round
  ffdl_backtrack_kernel<<<(count + perBlock - 1) / perBlock, kBtThreads, 0, stream>>>(a, r0, count, sc, K4cap);
  return cudaGetLastError();
}

}  // namespace voda
