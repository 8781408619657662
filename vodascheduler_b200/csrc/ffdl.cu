// ffdl.cu — FfDLOptimizer.Schedule (pkg/algorithm/ffdl_optimizer.go:32-130) for a batch of
// rounds, hand-written for sm_100a.
//
// One CTA owns one round at a time (persistent grid striding over rounds).  Per round:
//   1. stable sort of the jobs by SubmitTime (:45-47)            -> block bitonic argsort
//   2. trim to the first J = min(n, K) jobs (:54-65)
//   3. DP  P[j][k] = max_g S_j[g] + P[j-1][k-g], g in [1, max_j], k-g >= 0, rows initialised
//      to -10000 (:78-105) -> ONE ROW PER CTA STEP: rows j-1 and j live in shared memory
//      (float4 groups), each thread owns kCells adjacent cells and slides a register window
//      of row j-1 over g, so a block of 4 g-steps costs one LDS.128 of P and one LDS.128
//      (broadcast) of S for 4*kCells add+max pairs.  The forward pass keeps ONLY the values:
//      per candidate it is one FADD and half a three-input FMNMX3.
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

#ifndef VODA_FFDL_THREADS
#define VODA_FFDL_THREADS 64
#endif
#ifndef VODA_FFDL_CELLS
#define VODA_FFDL_CELLS 8
#endif
#ifndef VODA_FFDL_MINBLOCKS
#define VODA_FFDL_MINBLOCKS 12
#endif
constexpr int kThreads = VODA_FFDL_THREADS;
constexpr int kCells = VODA_FFDL_CELLS;  // adjacent cells per thread
constexpr int kGroups = kCells / 4 + 1;  // float4 groups of row j-1 in the sliding window
static_assert(kCells % 4 == 0, "cells per thread come in float4 groups");

struct FfdlSmem {
  float4* rowA;      // [G4]  row buffers, group q = P[4q-3 .. 4q]
  float4* rowB;      // [G4]
  float* srow0;      // [SW]  S_j[1..] of even rows, -inf beyond the row's g limit
  float* srow1;      // [SW]  ... of odd rows
  int64_t* khi;      // [Jp]
  int32_t* klo;      // [Jp]
  int* perm;         // [Jp]
  int32_t* smax;     // [Jp]  max_proc in sorted order
  int32_t* scum;     // [Jp]  klim: min(K, sum of clamp(max_proc, 0, K) over the first j+1 sorted jobs)
  int32_t* slo;      // [Jp]  cone: max(0, K - sum of clamp(max_proc, 0, K) over the sorted jobs after j)
  float* srmax;      // [Jp]  max over the row's usable g of S_j[g] (-inf if none)
  long long* red;    // [2]
  int* misc;         // [4]
};

// What the DP kernel leaves per round for the backtrack kernel ...
struct FfdlMeta {
  int32_t J;   // rows of the DP (jobs kept after the trim)
  int32_t st;  // VODA_ST_OK, or the panic raised before / by the DP
  int32_t K;
  int32_t pad;
};
// ... and per DP row j (1-based; record j-1), 16 bytes so that one load fetches it
struct FfdlRow {
  int32_t job;   // index of the row's job within the round (input order)
  int32_t gmax;  // min(max_proc, K): candidates are g = 1 .. min(gmax, k)
  int32_t klim;  // saturation point of THIS row: P[j][k] == c for k >= klim
  float c;       // the saturated value c_j
};

struct FfdlScratch {
  FfdlMeta* meta;     // [count]
  FfdlRow* rows;      // [count][Jcap]
  float* srows;       // [count][Jcap][SW]  S rows in sorted order, -inf beyond the table limit
  float* P;           // [count][Jcap][rowFloats]  P[j][1..K]: row j-1, float k-1
  size_t rowFloats;   // 4 * K4cap
  size_t slabFloats;  // Jcap * rowFloats
  int Jcap;
  int SW;
};

// Row buffers carry kPadGroups(SW) groups of -inf in front of group 0, so the window can slide
// left of the row (k-g < 0) without a branch.
__host__ __device__ constexpr int pad_groups(int SW) { return SW / 4 + 1; }

// 4 g-steps (g = 4m+1 .. 4m+4) for the kCells cells of a tile: cell c at step s reads element
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

// One DP row for column tile `tq` (cells k = kCells*tq+1 .. kCells*tq+kCells).
// Row storage: group q = P[4q-3 .. 4q].  For the block of g = 4m+1 .. 4m+4 the cells need
// P[base-3 .. base+kCells-1], base = kCells*tq - 4m, i.e. groups q0 .. q0+kCells/4 with
// q0 = (kCells/4)*tq - m; the next block slides the window down by one group (one LDS.128).
__device__ __forceinline__ void ffdl_row_tile(const float4* __restrict__ prev, float4* __restrict__ cur,
                                              const float* __restrict__ srow, int tq, int nblk, int gslow0,
                                              int gslow1) {
  float best[kCells];
#pragma unroll
  for (int c = 0; c < kCells; c++) best[c] = -10000.0f;
  const int qtop = (kCells / 4) * tq;
  const float4* s4 = reinterpret_cast<const float4*>(srow);
  if constexpr (kCells == 8) {
    // the window is three groups; instead of shifting registers after every block the three
    // roles rotate through three unrolled copies of the block
    float4 g0 = prev[qtop], g1 = prev[qtop + 1], g2 = prev[qtop + 2];
    int m = 0;
    for (; m + 3 <= nblk; m += 3) {
      ffdl_block8(best, s4[m], g0, g1, g2);
      g2 = prev[qtop - m - 1];
      ffdl_block8(best, s4[m + 1], g2, g0, g1);
      g1 = prev[qtop - m - 2];
      ffdl_block8(best, s4[m + 2], g1, g2, g0);
      g0 = prev[qtop - m - 3];
    }
    if (m < nblk) {
      ffdl_block8(best, s4[m], g0, g1, g2);
      if (m + 1 < nblk) {
        g2 = prev[qtop - m - 1];
        ffdl_block8(best, s4[m + 1], g2, g0, g1);
      }
    }
  } else {
    float4 grp[kGroups];
#pragma unroll
    for (int i = 0; i < kGroups; i++) grp[i] = prev[qtop + i];
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
#pragma unroll
      for (int c = 0; c < kCells; c++) {
        best[c] = fmaxf(fmaxf(best[c], sq.x + v[3 + c]), sq.y + v[2 + c]);
        best[c] = fmaxf(fmaxf(best[c], sq.z + v[1 + c]), sq.w + v[c]);
      }
#pragma unroll
      for (int i = kGroups - 1; i > 0; i--) grp[i] = grp[i - 1];
      grp[0] = prev[qtop - m - 1];
    }
  }
  // g beyond the dense speedup table: the Go map misses, S = 0 (rare: max_j >= W)
  if (gslow1 >= gslow0) {
    const float* pf = reinterpret_cast<const float*>(prev);
#pragma unroll
    for (int c = 0; c < kCells; c++) {
      const int k = kCells * tq + 1 + c;
      for (int g = gslow0; g <= gslow1 && g <= k; g++) best[c] = fmaxf(best[c], 0.0f + pf[k - g + 3]);
    }
  }
#pragma unroll
  for (int i = 0; i < kCells / 4; i++) {
    const float4 b4 = make_float4(best[4 * i], best[4 * i + 1], best[4 * i + 2], best[4 * i + 3]);
    cur[qtop + 1 + i] = b4;
  }
}

// Bulk asynchronous copy shared -> global (cp.async.bulk, the TMA engine without a tensor map).
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void bulk_store(float* dst, const float4* src, uint32_t bytes) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(src);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(s), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// All DP rows of one round.  Returns with the last row in `*last` (shared memory).
__device__ void ffdl_rows(const AllocArgs& a, const FfdlSmem& sm, int o, int J, int K, int KT, int SW,
                          const FfdlScratch& sc, int rr, const float4** last) {
  const int tid = threadIdx.x;
  const int W = a.stride;
  float4* prev = sm.rowA;
  float4* cur = sm.rowB;
  float* slab = sc.P + (size_t)rr * sc.slabFloats;
  float* gsrows = sc.srows + (size_t)rr * (size_t)sc.Jcap * (size_t)SW;
  FfdlRow* grows = sc.rows + (size_t)rr * (size_t)sc.Jcap;
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
    for (int g = tid + 1; g <= SW; g += kThreads) {
      const float s1 = s_value(1, g);
      sm.srow1[g - 1] = s1;
      gsrows[g - 1] = s1;
    }
  __syncthreads();
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
    const int klim = sm.scum[j - 1];
    const int klo = sm.slo[j - 1];
    // c_j = max(-10000, max_g (S_j[g] + c_{j-1})) = max(-10000, (max_g S_j[g]) + c_{j-1}): rounding
    // is monotone, so the maximum commutes with the addition exactly
    const float ccur = fmaxf(-10000.0f, sm.srmax[j - 1] + cprev);
    if (tid == 0) grows[j - 1] = FfdlRow{sm.perm[j - 1], gmax, klim, ccur};
    for (int tq = tid; tq < KT; tq += kThreads) {
      if (kCells * tq + kCells < klo) continue;
      if (kCells * tq + 1 > klim) {
        const float4 c4 = make_float4(ccur, ccur, ccur, ccur);
#pragma unroll
        for (int i = 0; i < kCells / 4; i++) cur[(kCells / 4) * tq + 1 + i] = c4;
      } else {
        ffdl_row_tile(prev, cur, srow, tq, nblk, max(W, 1), gmax);
      }
    }
    cprev = ccur;
    if (tid == 0) cur[0] = make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, -10000.0f);  // P[j][0]
    if (j < J) {
      float* gs = gsrows + (size_t)j * (size_t)SW;
#pragma unroll
      for (int u = 0; u < 2; u++) {
        int g = tid + 1 + u * kThreads;
        if (g <= SW) {
          snext[g - 1] = pre[u];
          gs[g - 1] = pre[u];
        }
      }
      for (int g = tid + 1 + 2 * kThreads; g <= SW; g += kThreads) {
        const float s2 = s_value(j + 1, g);
        snext[g - 1] = s2;
        gs[g - 1] = s2;
      }
    }
    // Row j goes to the round's slab as ONE bulk copy (shared -> global, issued by thread 0 and
    // carried out by the copy engine while the CTA computes row j+1 from the same buffer), so the
    // store costs the threads neither issue slots nor LSU cycles.  The copy of row j-1 has had a
    // whole row of compute to finish reading its buffer, which row j+1 is about to overwrite.
    fence_proxy_async_smem();
    if (tid == 0) bulk_wait_read();
    __syncthreads();
    if (tid == 0 && j < J) {
      const int tlo = max(0, (klo + kCells - 1) / kCells - 1);  // first tile with a cell >= klo
      const int thi = min(KT - 1, (klim - 1) / kCells);         // last tile that starts at or below klim
      if (klim >= 1 && thi >= tlo) {
        const int q0 = (kCells / 4) * tlo;
        bulk_store(slab + (size_t)(j - 1) * sc.rowFloats + (size_t)q0 * 4, cur + q0 + 1,
                   (uint32_t)(thi - tlo + 1) * kCells * 4u);
      }
    }
    float4* t = prev;
    prev = cur;
    cur = t;
  }
  if (tid == 0) bulk_wait_all();
  *last = prev;
}

// Kernel A: sort + DP rows.  One CTA per round at a time; P rows / row records / meta go to the
// round's own slot of the scratch so that the CTA can move on to its next round immediately.
__global__ void __launch_bounds__(kThreads, VODA_FFDL_MINBLOCKS)
    ffdl_dp_kernel(AllocArgs a, int r0, int count, FfdlScratch sc, int K4cap, int SW, int Jp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FfdlSmem sm;
  {
    unsigned char* p = smem_raw;
    const int pad = pad_groups(SW);
    sm.rowA = reinterpret_cast<float4*>(p) + pad; p += sizeof(float4) * (size_t)(pad + K4cap + 1);
    sm.rowB = reinterpret_cast<float4*>(p) + pad; p += sizeof(float4) * (size_t)(pad + K4cap + 1);
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
  for (int q = tid; q < pad_groups(SW); q += kThreads) {  // written once: nothing else stores there
    const float4 ninf = make_float4(VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF, VODA_NEG_INF);
    sm.rowA[-1 - q] = ninf;
    sm.rowB[-1 - q] = ninf;
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
    // 1. stable sort by SubmitTime
    for (int i = tid; i < n; i += kThreads) {
      sm.khi[i] = a.submit_sec ? a.submit_sec[o + i] : 0;
      sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
    }
    __syncthreads();
    block_argsort(sm.khi, sm.klo, sm.perm, n, next_pow2(n));
    // 2. trim
    const int J = min(n, K);
    for (int j = tid; j < J; j += kThreads) {
      const int pj = sm.perm[j];
      const int mx = a.max_proc[o + pj];
      sm.smax[j] = mx;
      // largest speedup value the row can use (g beyond the table reads 0: a Go map miss)
      const int gmx = min(mx, K);
      float rm = gmx >= max(a.stride, 1) ? 0.0f : VODA_NEG_INF;
      const float* sj = a.speedup + (a.sp_off ? (size_t)a.sp_off[o + pj] : (size_t)(o + pj) * (size_t)a.stride);
      for (int g = 1; g <= min(gmx, a.stride - 1); g++) rm = fmaxf(rm, sj[g]);
      sm.srmax[j] = rm;
    }
    __syncthreads();
    if (tid == 0) {  // the two running sums of the usable sizes (short serial scans)
      long long acc = 0;
      for (int j = 0; j < J; j++) {
        acc += max(0, min(sm.smax[j], K));
        sm.scum[j] = (int)min(acc, (long long)K);
      }
      acc = 0;
      for (int j = J - 1; j >= 0; j--) {
        sm.slo[j] = (int)max(0LL, (long long)K - acc);
        acc += max(0, min(sm.smax[j], K));
      }
    }
    __syncthreads();
    // 3-4. DP rows
    const int KT = (K + kCells - 1) / kCells;
    const float4* last = nullptr;
    ffdl_rows(a, sm, o, J, K, KT, SW, sc, rr, &last);
    const float PJK = reinterpret_cast<const float*>(last)[K + 3];
    meta.J = J;
    if (PJK <= 0.0f) meta.st = VODA_ST_INFEASIBLE;  // :109-114 (NaN is not <= 0, as in Go)
    if (tid == 0) sc.meta[rr] = meta;
  }
}

// Kernel B: backtrack (:116-125) + deferred validateResult (:42).
// The walk is a chain of J dependent steps (row j-1 is read at k - g_j), so it is all latency,
// and what it reads is scattered over the batch's slabs.  EIGHT LANES walk one round (four
// rounds per warp, every round of the batch in flight at once): at row j they re-evaluate the
// reference's inner loop (:93-103) for the single cell (j, k),
//   best = -10000, sol = 0;  for g = 1 .. min(max_j, K, k):  p = S_j[g] + P[j-1][k-g];
//   if p > best { best = p; sol = g }
// eight candidates at a time, one per lane: the eight P values are adjacent floats of the stored
// row (one or two sectors per step instead of eight scattered requests), and the strict-'>'
// scan becomes a three-step butterfly that keeps the larger value and, among equal values, the
// smaller g.  P[j-1][x] comes from the stored row, its saturated constant, or row 0 = 0.
// Only that load depends on the position: the row records and the S values of the next row are
// fetched one row ahead.  Then the warp validates its four rounds with coalesced reads.
constexpr int kBtThreads = 128;
constexpr int kBtLanes = 8;                  // lanes per round
constexpr int kBtRounds = 32 / kBtLanes;     // rounds per warp

__global__ void __launch_bounds__(kBtThreads)
    ffdl_backtrack_kernel(AllocArgs a, int r0, int count, FfdlScratch sc) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & (kBtLanes - 1);
  const int rr0 = (blockIdx.x * (kBtThreads / 32) + (threadIdx.x >> 5)) * kBtRounds;
  if (rr0 >= count) return;
  // ---- phase 1: one walk per group of eight lanes
  {
    const int rr = rr0 + lane / kBtLanes;
    const unsigned gmask = ((1u << kBtLanes) - 1u) << (lane & ~(kBtLanes - 1));
    if (rr < count) {
      const int r = r0 + rr;
      const int o = a.job_off[r] - a.job_base;
      const FfdlMeta meta = sc.meta[rr];
      if (meta.st == VODA_ST_OK && meta.J > 0) {
        const float* slab = sc.P + (size_t)rr * sc.slabFloats;
        const float* gsrows = sc.srows + (size_t)rr * (size_t)sc.Jcap * (size_t)sc.SW;
        const FfdlRow* grows = sc.rows + (size_t)rr * (size_t)sc.Jcap;
        const int SW = sc.SW;
        const int W = a.stride;
        const int ssub = min(sub, SW - 1);
        const FfdlRow none{0, 0, 0, 0.0f};
        int k = meta.K;
        FfdlRow rc = grows[meta.J - 1];                      // row j
        FfdlRow rp = meta.J >= 2 ? grows[meta.J - 2] : none;  // row j-1: its klim and c describe the row read
        float s1 = gsrows[(size_t)(meta.J - 1) * SW + ssub];  // S_j[1 + sub]
        for (int j = meta.J; j >= 1; j--) {
          // one row ahead (none of this depends on the position)
          const FfdlRow rpp = j >= 3 ? grows[j - 3] : none;
          const float s1next = j >= 2 ? gsrows[(size_t)(j - 2) * SW + ssub] : 0.0f;
          const int gm = min(rc.gmax, k);  // k-g >= 0
          const float* sj = gsrows + (size_t)(j - 1) * SW;              // sj[g-1] = S_j[g]
          const float* prow = slab + (size_t)max(j - 2, 0) * sc.rowFloats;  // P[j-1][x] at prow[x-1] (j >= 2)
          float best = -10000.0f;
          int sol = 0;
          for (int g0 = 1; g0 <= gm; g0 += kBtLanes) {
            const int g = g0 + sub;
            const int x = k - g;
            // unconditional loads on clamped, always-valid addresses; the cases are sorted out below
            const float pv = prow[max(x, 1) - 1];
            const float sv = g0 == 1 ? s1 : sj[min(g, SW) - 1];
            const float sg = g <= W - 1 ? sv : 0.0f;  // a g the table does not hold: Go map miss
            float pg = x <= 0 ? -10000.0f : (x >= rp.klim ? rp.c : pv);
            if (j == 1) pg = 0.0f;  // row 0
            float p = sg + pg;
            if (g > gm || p != p) p = VODA_NEG_INF;  // no such candidate / NaN: can never win
            int gg = g;
#pragma unroll
            for (int d = 1; d < kBtLanes; d <<= 1) {
              const float po = __shfl_xor_sync(gmask, p, d);
              const int go = __shfl_xor_sync(gmask, gg, d);
              if (po > p || (po == p && go < gg)) {
                p = po;
                gg = go;
              }
            }
            if (p > best) {
              best = p;
              sol = gg;
            }
          }
          if (sub == 0) a.out_gpus[o + rc.job] = sol;
          k -= sol;
          rc = rp;
          rp = rpp;
          s1 = s1next;
        }
      }
    }
  }
  __syncwarp();
  // ---- phase 2: validateResult (utils.go:18-42) per round; a panic it raises replaces the one in flight
  for (int q = 0; q < kBtRounds && rr0 + q < count; q++) {
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
  *smem = sizeof(float4) * 2 * (size_t)(pad_groups(*SW) + *K4cap + 1) + sizeof(float) * 2 * (size_t)(*SW) +
          (sizeof(int64_t) + 6 * sizeof(int32_t)) * (size_t)(*Jp) + 2 * sizeof(long long) + 4 * sizeof(int);
}

static size_t ffdl_jcap(const AllocArgs& a) {
  const int j = a.max_jobs < a.max_total ? a.max_jobs : a.max_total;
  return (size_t)(j > 0 ? j : 1);
}

// Scratch one round needs: its P slab, the sorted S rows, the row records and the meta record
// (every part a multiple of 16 bytes).
size_t ffdl_scratch_per_round(const AllocArgs& a) {
  int K4cap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &K4cap, &SW, &Jp, &smem);
  const size_t Jcap = ffdl_jcap(a);
  return Jcap * (size_t)K4cap * 16 + Jcap * (size_t)SW * 4 + Jcap * sizeof(FfdlRow) + sizeof(FfdlMeta);
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
  FfdlScratch sc;
  unsigned char* base = static_cast<unsigned char*>(scratch);
  sc.rowFloats = (size_t)K4cap * 4;
  sc.slabFloats = Jcap * sc.rowFloats;
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
  ffdl_dp_kernel<<<grid, kThreads, smem, stream>>>(a, r0, count, sc, K4cap, SW, Jp);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int perBlock = (kBtThreads / 32) * kBtRounds;  // rounds per block: eight lanes per round
  ffdl_backtrack_kernel<<<(count + perBlock - 1) / perBlock, kBtThreads, 0, stream>>>(a, r0, count, sc);
  return cudaGetLastError();
}

}  // namespace voda
