// common.cuh — shared device helpers for the allocation kernels (sm_100a).
//
// Every kernel here follows the same shape: one CTA owns one scheduling round at a time
// (rounds are independent; a persistent grid strides over them), the round's per-job state
// lives in shared memory, and the float32 arithmetic is written operation by operation in the
// reference's evaluation order.  The library is compiled with -fmad=false and without
// fast-math so nvcc neither contracts a*b+c nor relaxes IEEE division.
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/voda_accel.h"

namespace voda {

// Kernel-side view of a voda_alloc_batch: device pointers only, plus the outputs.
struct AllocArgs {
  int32_t algo, rounds;
  int32_t job_base;  // job_off values are absolute; the job arrays start at this job
  const int32_t* job_off;
  const int32_t* total_gpu;
  const int64_t* submit_sec;
  const int32_t* submit_nsec;
  const int64_t* first_sec;
  const int32_t* first_nsec;
  const int32_t* priority;
  const int32_t* num_proc;
  const int32_t* min_proc;
  const int32_t* max_proc;
  const float* remaining_sec;
  const float* speedup;
  const int32_t* sp_off;  // null: dense table [job * stride + g].  Else PACKED rows (host path of
                          // FfDL): entry g of job j is speedup[sp_off[j] + g], only the g the
                          // policy can read are present
  int32_t stride;      // speedup table width W
  int32_t queue_num;   // TiresiasQueueNum
  int32_t compaction;  // ElasticTiresiasCompactionThreshold
  int32_t max_jobs;    // declared upper bound of jobs per round
  int32_t max_total;   // declared upper bound of total_gpu
  int32_t max_gmax;    // declared upper bound of min(MaxNumProc, total_gpu) over the batch; 0: not declared
  int32_t* out_gpus;
  int32_t* out_status;
};

// Kernel-side view of a voda_place_batch + voda_place_out (device pointers).
struct PlaceArgs {
  int32_t rounds;
  const int32_t* node_off;
  const int32_t* node_slots;
  const int32_t* req_off;
  const int32_t* req_workers;
  const int32_t* prev_off;
  const int32_t* prev_job;
  const int32_t* prev_workers;
  int32_t max_nodes, max_requests;
  const int32_t* pods_off;  // optional pod -> node table (run-length per request), see voda_accel.h
  const int32_t* pods_node;
  const int32_t* pods_slots;
  int32_t* migrated_workers;  // optional outputs
  int32_t* migrated_launchers;
  int32_t* phys_of_logical;
  int32_t* free_slots;
  int32_t* lay_cnt;
  int32_t* lay_node;
  int32_t* lay_job;
  int32_t* lay_workers;
  long long* score;
  int32_t* cross;
  int32_t* status;
};

struct TickArgs {  // voda_tick_batch with the thresholds by value
  int jobs, tiresias, queue_num, promote_knob;
  long long now_ns;
  double thr[VODA_TICK_MAX_QUEUES];
  const int32_t* status;
  const int32_t* num_gpu;
  int32_t* priority;
  long long *last_update_ns, *running_ns, *gpu_ns, *total_ns, *waiting_ns, *last_running_ns, *last_gpu_ns,
      *last_waiting_ns;
  int32_t* changed;
};

__host__ __device__ inline int next_pow2(int x) {
  int p = 1;
  while (p < x) p <<= 1;
  return p;
}

#define VODA_NEG_INF (__int_as_float(0xff800000))

// Order-preserving image of a float32 sort key under Go's `<` (-0 == +0; NaN out of domain).
__device__ __forceinline__ int64_t f32_order_key(float x) {
  int32_t b = __float_as_int(x);
  int32_t mag = b & 0x7fffffff;  // -0 and +0 both map to 0
  return b >= 0 ? (int64_t)mag : -(int64_t)mag;
}

// Strict total order on job slots: (hi, lo, slot) with slots >= n acting as +infinity.
// Because the slot index is the last component, sorting with it yields exactly the result
// of a STABLE sort by (hi, lo) — what sort.SliceStable produces in the reference.
__device__ __forceinline__ bool key_less(const int64_t* hi, const int32_t* lo, int n, int a, int b) {
  if (b >= n) return a < n ? true : a < b;
  if (a >= n) return false;
  int64_t ha = hi[a], hb = hi[b];
  if (ha != hb) return ha < hb;
  int32_t la = lo[a], lb = lo[b];
  if (la != lb) return la < lb;
  return a < b;
}

// Block-wide bitonic argsort: perm[0..P) <- slots ordered by the strict total order `less`
// (slots >= n must sort last); P = pow2 >= n.  All threads of the CTA must call it.
template <typename Less>
__device__ inline void block_argsort_by(int* perm, int P, Less less) {
  for (int i = threadIdx.x; i < P; i += blockDim.x) perm[i] = i;
  __syncthreads();
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < P; i += blockDim.x) {
        int ixj = i ^ j;
        if (ixj > i) {
          int a = perm[i], b = perm[ixj];
          bool up = (i & k) == 0;
          bool a_after_b = less(b, a);
          if (a_after_b == up) {
            perm[i] = b;
            perm[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
}

// Rank sort for small rounds: perm[rank(i)] = i where rank(i) = #{j : less(j, i)}; `less` must be
// a strict total order on [0, n).  One pass over the n keys per thread (they are read as
// shared-memory broadcasts) and a single barrier, against log^2(n) barriers for the bitonic
// network: cheaper up to n of about a hundred.  All threads of the CTA must call it.
template <typename Less>
__device__ inline void block_ranksort_by(int* perm, int n, Less less) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    int rank = 0;
    for (int j = 0; j < n; j++) rank += less(j, i) ? 1 : 0;
    perm[rank] = i;
  }
  __syncthreads();
}

// The same sort when every key of the round packs, order preserved, into ONE unsigned 64-bit
// word: a bitonic network over (key, slot) pairs, slot = input index breaking ties (the result
// of a stable sort by key).  A compare-exchange is two LDS.64, two LDS.32, one 64-bit compare and
// the conditional stores, against a dozen dependent shared loads in the indirect argsort; each
// thread takes whole pairs, so no thread idles on the "upper" element.  The caller fills
// key[0..P) (slots >= n: all ones, which sorts them last in slot order) before the call; perm is
// written here.  All threads of the CTA must call it.
__device__ inline void block_sort_kv(unsigned long long* key, int* perm, int P) {
  for (int i = threadIdx.x; i < P; i += blockDim.x) perm[i] = i;
  __syncthreads();
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (P >> 1); t += blockDim.x) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // the pair's lower slot (bit j clear)
        const int ixj = i | j;
        const unsigned long long ka = key[i], kb = key[ixj];
        const int pa = perm[i], pb = perm[ixj];
        const bool up = (i & k) == 0;
        const bool a_after_b = kb < ka || (kb == ka && pb < pa);
        if (a_after_b == up) {
          key[i] = kb;
          key[ixj] = ka;
          perm[i] = pb;
          perm[ixj] = pa;
        }
      }
      __syncthreads();
    }
  }
}

// (seconds, nanoseconds) -> one word, when it fits: seconds in [0, 2^32), nanoseconds in [0, 2^30).
__device__ __forceinline__ bool pack_time_key(long long sec, int nsec, unsigned long long* out) {
  *out = ((unsigned long long)sec << 30) | (unsigned long long)(unsigned)nsec;
  return sec >= 0 && sec < (1ll << 32) && nsec >= 0 && nsec < (1 << 30);
}

__device__ inline void block_argsort(const int64_t* hi, const int32_t* lo, int* perm, int n, int P) {
  block_argsort_by(perm, P, [=](int a, int b) { return key_less(hi, lo, n, a, b); });
}

// "Give need[i] while it fits", in order i = 0..n-1:  if free >= need[i] { take; free -= need[i] }.
// The dependence on `free` is serial (a job that does not fit is skipped and a later, smaller
// one may fit), so ONE WARP evaluates it, 32 jobs at a time: prefix sums find the first job
// that fails given that all pending ones before it were accepted; it is rejected and the rest
// of the chunk is re-evaluated.  Exact for any sign of need[].  Call with a full warp; writes
// acc[i] in {0,1} and returns the remaining free (same value in every lane).
__device__ inline long long warp_skip_scan(const int32_t* need, int n, long long free, unsigned char* acc) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    const bool valid = i < n;
    const long long m = valid ? need[i] : 0;
    bool pending = valid;
    bool accepted = false;
    const bool anyNeg = __any_sync(full, valid && m < 0);
    while (true) {
      unsigned pend = __ballot_sync(full, pending);
      if (!pend) break;
      if (!anyNeg) {
        // free never grows inside the chunk: jobs larger than it can be rejected at once
        const bool hopeless = pending && m > free;
        const unsigned h = __ballot_sync(full, hopeless);
        if (h) {
          if (hopeless) pending = false;
          pend &= ~h;
          if (!pend) break;
        }
      }
      long long incl = pending ? m : 0;  // inclusive prefix of need over pending lanes
      for (int d = 1; d < 32; d <<= 1) {
        long long up = __shfl_up_sync(full, incl, d);
        if (lane >= d) incl += up;
      }
      const long long excl = incl - (pending ? m : 0);
      const bool fits = pending && (free - excl >= m);
      const unsigned failMask = __ballot_sync(full, pending && !fits);
      if (!failMask) {  // everything pending is accepted
        if (pending) accepted = true;
        free -= __shfl_sync(full, incl, 31 - __clz(pend));
        break;
      }
      const int f = __ffs(failMask) - 1;  // first failing lane: rejected; pending lanes before it accepted
      if (pending && lane < f) {
        accepted = true;
        pending = false;
      }
      if (lane == f) pending = false;
      free -= __shfl_sync(full, excl, f);
    }
    if (valid) acc[i] = accepted ? 1 : 0;
  }
  return free;
}

// validateResult (pkg/algorithm/utils.go:18-42) over one round; all threads call.
// `skip` (may be null) marks jobs absent from the reference's result map.
// s_red: two 64-bit words of shared memory.  Returns the round status (tie rule: smallest
// code among the offending jobs, then the total check).
__device__ inline int block_validate(const AllocArgs& a, int o, int n, int total, const int32_t* r,
                                     const unsigned char* skip, long long* s_red) {
  if (threadIdx.x == 0) {
    s_red[0] = 99;
    s_red[1] = 0;
  }
  __syncthreads();
  int worst = 99;
  long long sum = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (skip && skip[i]) continue;
    int v = r[i];
    int st = 99;
    if (v < 0) st = VODA_ST_NEGATIVE;
    else if (v > 0 && v < a.min_proc[o + i]) st = VODA_ST_LESS_THAN_MIN;
    else if (v > a.max_proc[o + i]) st = VODA_ST_EXCEEDED_MAX;
    worst = min(worst, st);
    sum += v;
  }
  for (int d = 16; d > 0; d >>= 1) {
    worst = min(worst, __shfl_xor_sync(0xffffffffu, worst, d));
    sum += __shfl_xor_sync(0xffffffffu, sum, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(reinterpret_cast<int*>(&s_red[0]), worst);
    atomicAdd(reinterpret_cast<unsigned long long*>(&s_red[1]), (unsigned long long)sum);
  }
  __syncthreads();
  int st = *reinterpret_cast<int*>(&s_red[0]);
  long long tot = s_red[1];
  __syncthreads();
  if (st != 99) return st;
  return tot > (long long)total ? VODA_ST_EXCEEDED_TOTAL : VODA_ST_OK;
}

}  // namespace voda
