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
//   4. SOL[j][k] (the chosen g) is streamed to a per-CTA slab of global scratch (uint8 when
//      the round's largest g fits, else uint16); it is written once and only J cells are read
//      back by the backtrack (:116-125).
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

constexpr int kThreads = 64;

struct FfdlSmem {
  float4* rowA;      // [G4]  row buffers, group q = P[4q-3 .. 4q]
  float4* rowB;      // [G4]
  float* srow0;      // [SW]  S_j[1..] of even rows, -inf beyond the row's g limit
  float* srow1;      // [SW]  ... of odd rows
  int64_t* khi;      // [Jp]
  int32_t* klo;      // [Jp]
  int* perm;         // [Jp]
  int32_t* smax;     // [Jp]  max_proc in sorted order
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
// (acc = acc*2 + bit, exact below 2^24), i.e. on the FMA pipe, and the last win is decoded
// once per flush.  fmaxf(best, p) equals the reference's strict-'>' select for every non-NaN
// input (equal values are interchangeable; a NaN p never replaces best in either form).
#define VODA_CELL(c, s)                              \
  {                                                  \
    const float p = sv[s] + v[3 + (c) - (s)];        \
    const float t = (p > best[c]) ? 1.0f : 0.0f;     \
    best[c] = fmaxf(best[c], p);                     \
    acc[c] = fmaf(acc[c], 2.0f, t);                  \
  }

constexpr int kCells = 16;              // adjacent cells per thread
constexpr int kGroups = kCells / 4 + 1;  // float4 groups of row j-1 in the sliding window
constexpr int kFlushBlocks = 6;          // 6 blocks x 4 g-steps = 24 win bits per accumulator flush

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
      // decode the last win of the flushed span: bit b (from the LSB) is g = gEnd - b
      const int gEnd = 4 * m + 4;
#pragma unroll
      for (int c = 0; c < kCells; c++) {
        const int bits = __float2int_rz(acc[c]);
        if (bits) bg[c] = gEnd - (__ffs(bits) - 1);
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
  if (sizeof(SolT) == 1) {
    uint32_t w[kCells / 4];
#pragma unroll
    for (int i = 0; i < kCells / 4; i++)
      w[i] = (uint32_t)bg[4 * i] | ((uint32_t)bg[4 * i + 1] << 8) | ((uint32_t)bg[4 * i + 2] << 16) |
             ((uint32_t)bg[4 * i + 3] << 24);
    reinterpret_cast<uint4*>(solrow)[tq] = make_uint4(w[0], w[1], w[2], w[3]);
  } else {
    uint32_t w[kCells / 2];
#pragma unroll
    for (int i = 0; i < kCells / 2; i++) w[i] = (uint32_t)bg[2 * i] | ((uint32_t)bg[2 * i + 1] << 16);
    reinterpret_cast<uint4*>(solrow)[2 * tq] = make_uint4(w[0], w[1], w[2], w[3]);
    reinterpret_cast<uint4*>(solrow)[2 * tq + 1] = make_uint4(w[4], w[5], w[6], w[7]);
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
    return a.speedup[(size_t)(o + sm.perm[j - 1]) * (size_t)W + (size_t)g];
  };
  if (J >= 1)
    for (int g = tid + 1; g <= SW; g += kThreads) sm.srow1[g - 1] = s_value(1, g);
  __syncthreads();
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
    for (int tq = tid; tq < KT; tq += kThreads)
      ffdl_row_tile<SolT>(prev, cur, srow, tq, nblk, max(W, 1), gmax, solrow);
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

__global__ void __launch_bounds__(kThreads) ffdl_kernel(AllocArgs a, unsigned char* scratch, size_t slabBytes,
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
    sm.misc = reinterpret_cast<int*>(p);
  }
  const int tid = threadIdx.x;
  unsigned char* slab = scratch + (size_t)blockIdx.x * slabBytes;

  for (int r = blockIdx.x; r < a.rounds; r += gridDim.x) {
    const int o = a.job_off[r] - a.job_base;
    const int n = a.job_off[r + 1] - a.job_off[r];
    const int K = a.total_gpu[r];
    if (n == 0) {  // :33-37 returns the empty result before anything else
      if (tid == 0) a.out_status[r] = VODA_ST_OK;
      continue;
    }
    if (n > a.max_jobs || K > a.max_total) {
      for (int i = tid; i < n; i += kThreads) a.out_gpus[o + i] = 0;
      if (tid == 0) a.out_status[r] = VODA_ST_BOUND;
      continue;
    }
    for (int i = tid; i < n; i += kThreads) a.out_gpus[o + i] = 0;  // :38-40
    int st = VODA_ST_OK;
    if (K < 0) {
      st = VODA_ST_RUNTIME_PANIC;  // make(ReadyJobs, 1, K+1) with cap < len (:54)
    } else {
      // 1. stable sort by SubmitTime
      const int P2 = next_pow2(n);
      for (int i = tid; i < n; i += kThreads) {
        sm.khi[i] = a.submit_sec ? a.submit_sec[o + i] : 0;
        sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
      }
      __syncthreads();
      block_argsort(sm.khi, sm.klo, sm.perm, n, P2);
      // 2. trim
      const int J = min(n, K);
      int lmax = 0;
      for (int j = tid; j < J; j += kThreads) {
        int mx = a.max_proc[o + sm.perm[j]];
        sm.smax[j] = mx;
        lmax = max(lmax, mx);
      }
      if (tid == 0) sm.misc[0] = 0;
      __syncthreads();
      for (int d = 16; d > 0; d >>= 1) lmax = max(lmax, __shfl_xor_sync(0xffffffffu, lmax, d));
      if ((tid & 31) == 0) atomicMax(&sm.misc[0], lmax);
      __syncthreads();
      const int gmaxRound = min(sm.misc[0], K);
      // 3-4. DP rows
      const int KT = (K + kCells - 1) / kCells;
      const float4* last = nullptr;
      if (gmaxRound <= 255) {
        const size_t solStride = (size_t)K4cap * 4;
        ffdl_rows<uint8_t>(a, sm, o, J, K, KT, SW, solStride, slab, &last);
        const float PJK = reinterpret_cast<const float*>(last)[K + 3];
        if (PJK <= 0.0f) {
          st = VODA_ST_INFEASIBLE;
        } else if (tid == 0) {
          int k = K;
          for (int j = J; j > 0; j--) {
            int g = k > 0 ? slab[(size_t)(j - 1) * solStride + (size_t)(k - 1)] : 0;  // SOL[j][0] == 0
            a.out_gpus[o + sm.perm[j - 1]] = g;
            k -= g;
          }
        }
      } else {
        const size_t solStride = (size_t)K4cap * 8;
        ffdl_rows<uint16_t>(a, sm, o, J, K, KT, SW, solStride, slab, &last);
        const float PJK = reinterpret_cast<const float*>(last)[K + 3];
        if (PJK <= 0.0f) {
          st = VODA_ST_INFEASIBLE;
        } else if (tid == 0) {
          int k = K;
          for (int j = J; j > 0; j--) {
            int g = k > 0 ? reinterpret_cast<const uint16_t*>(slab + (size_t)(j - 1) * solStride)[k - 1] : 0;
            a.out_gpus[o + sm.perm[j - 1]] = g;
            k -= g;
          }
        }
      }
    }
    __syncthreads();  // out_gpus of this round written by thread 0 / zeroing threads
    // 5. deferred validateResult: a panic it raises replaces the one in flight
    int v = block_validate(a, o, n, K, a.out_gpus + o, nullptr, sm.red);
    if (tid == 0) a.out_status[r] = v ? v : st;
  }
}

}  // namespace

int ffdl_grid(int sms, int rounds) {
  int g = sms * 8;  // resident CTAs per SM (64 threads each)
  return rounds < g ? (rounds > 0 ? rounds : 1) : g;
}

static void ffdl_shape(const AllocArgs& a, int* K4cap, int* SW, int* Jp, size_t* smem) {
  const int Kmax = a.max_total > 0 ? a.max_total : 1;
  *K4cap = (kCells / 4) * ((Kmax + kCells - 1) / kCells);  // row groups of 4 cells, whole tiles
  int sw = a.stride - 1 < Kmax ? a.stride - 1 : Kmax;
  if (sw < 1) sw = 1;
  *SW = (sw + 3) & ~3;
  *Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  *smem = sizeof(float4) * 2 * (size_t)(*K4cap + 1) + sizeof(float) * 2 * (size_t)(*SW) +
          (sizeof(int64_t) + 3 * sizeof(int32_t)) * (size_t)(*Jp) + 2 * sizeof(long long) + 4 * sizeof(int);
}

size_t ffdl_scratch_bytes(const AllocArgs& a, int grid) {
  int K4cap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &K4cap, &SW, &Jp, &smem);
  const size_t Jcap = (size_t)(a.max_jobs < a.max_total ? a.max_jobs : a.max_total);
  return (size_t)grid * (Jcap > 0 ? Jcap : 1) * (size_t)K4cap * 8;
}

size_t ffdl_smem_bytes(const AllocArgs& a) {
  int K4cap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &K4cap, &SW, &Jp, &smem);
  return smem;
}

cudaError_t launch_ffdl(const AllocArgs& a, int grid, void* scratch, size_t scratch_bytes,
                        cudaStream_t stream) {
  int K4cap, SW, Jp;
  size_t smem;
  ffdl_shape(a, &K4cap, &SW, &Jp, &smem);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(ffdl_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const size_t slab = scratch_bytes / (size_t)grid;
  ffdl_kernel<<<grid, kThreads, smem, stream>>>(a, static_cast<unsigned char*>(scratch), slab, K4cap, SW, Jp);
  return cudaGetLastError();
}

}  // namespace voda
