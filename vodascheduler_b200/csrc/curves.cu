// curves.cu — the speedup-curve builder for a batch of jobs (sm_100a): the arithmetic of
// python/metrics_collector/metrics_collector.py:131-167.  Per job, from its
// (workers, epoch_time_sec) records and the linear defaults of CreateBaseJobInfo
// (pkg/common/mongo/mongo.go:76-100):
//   epoch_time[w] = mean of the epoch times recorded with w workers      :139-141 (pandas: sum / count)
//   speedup[k]    = epoch_time[1] / epoch_time[k],  k >= 1               :150-154
//   efficiency[k] = speedup[k] / float(k)                                :163-167
// All float64, and bit-identical: the sum is numpy's PAIRWISE summation over the job's records
// of that worker count in record order (blocks of up to 128 elements with eight running
// accumulators, halves rounded down to a multiple of eight above that), restated here as a
// recursion over a cursor that walks the job's records once, left to right.
// One warp per job: a coalesced pass counts the records of every worker count, then one lane per
// worker count that occurs walks the records and sums its own.
#include "common.cuh"
#include "kernels.h"

namespace voda {
namespace {

constexpr int kCntCap = 64;  // worker counts with a shared-memory counter per warp

constexpr int kStageCap = 256;  // records per job that fit the warp's staging buffer

struct Cursor {  // the records of one job with workers == w, in record order
  const double* run;  // non-null: the values already gathered into one contiguous run
  const int32_t* workers;
  const double* t;
  int pos, end, w;
  __device__ double next() {
    if (run) return run[pos++];
    while (pos < end && workers[pos] != w) pos++;
    return pos < end ? t[pos++] : 0.0;
  }
};

// numpy @TYPE@_pairwise_sum (loops_utils.h.src) for the next n values of the cursor
__device__ double pairwise_sum(Cursor& c, int n) {
  if (n < 8) {
    double res = 0.0;
    for (int i = 0; i < n; i++) res += c.next();
    return res;
  }
  if (n <= 128) {
    double r[8];
#pragma unroll
    for (int u = 0; u < 8; u++) r[u] = c.next();
    int i = 8;
    for (; i < n - (n % 8); i += 8) {
#pragma unroll
      for (int u = 0; u < 8; u++) r[u] += c.next();
    }
    double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; i++) res += c.next();
    return res;
  }
  int n2 = n / 2;
  n2 -= n2 % 8;
  const double a = pairwise_sum(c, n2);
  return a + pairwise_sum(c, n - n2);
}

__global__ void __launch_bounds__(128) curves_kernel(int jobs, const int32_t* __restrict__ rec_off,
                                                     const int32_t* __restrict__ workers,
                                                     const double* __restrict__ t, int W, double* out_time,
                                                     double* out_speedup, double* out_eff, int32_t* out_status) {
  const int lane = threadIdx.x & 31;
  const int job = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (job >= jobs) return;
  const int r0 = rec_off[job], r1 = rec_off[job + 1];
  // a worker count outside [0, W) has no slot in the dense table
  bool bad = false;
  for (int i = r0 + lane; i < r1; i += 32) bad |= workers[i] < 0 || workers[i] >= W;
  bad = __any_sync(0xffffffffu, bad);
  double* et = out_time + (size_t)job * W;
  double* sp = out_speedup + (size_t)job * W;
  double* ef = out_eff + (size_t)job * W;
  // how many records each worker count has: one coalesced pass with shared-memory counters
  // when the table is small enough for them (kCntCap slots per warp), a scan per lane otherwise
  __shared__ int cntS[4][kCntCap];
  int* cnts = cntS[threadIdx.x >> 5];
  const bool counted = W <= kCntCap;
  if (counted) {
    for (int w = lane; w < W; w += 32) cnts[w] = 0;
    __syncwarp();
    for (int i = r0 + lane; i < r1; i += 32) {
      const int w = workers[i];
      if (w >= 0 && w < W) atomicAdd(&cnts[w], 1);
    }
    __syncwarp();
  }
  // small jobs: a second coalesced pass gathers the epoch times of every worker count into its
  // own contiguous run of the staging buffer, in record order (stable: a record's slot is its
  // count's base + the number of earlier records with the same count)
  __shared__ double stageS[4][kStageCap];
  __shared__ int baseS[4][kCntCap];
  double* stage = stageS[threadIdx.x >> 5];
  int* base = baseS[threadIdx.x >> 5];
  const bool staged = counted && (r1 - r0) <= kStageCap;
  if (staged) {
    if (lane == 0) {
      int acc = 0;
      for (int w = 0; w < W; w++) {
        base[w] = acc;
        acc += cnts[w];
      }
    }
    __syncwarp();
    for (int i0 = r0; i0 < r1; i0 += 32) {
      const int i = i0 + lane;
      const int w = i < r1 ? workers[i] : -1;
      const bool ok = w >= 0 && w < W;
      const unsigned same = __match_any_sync(0xffffffffu, ok ? w : -1 - lane);
      if (ok) {
        const int rank = __popc(same & ((1u << lane) - 1u));
        stage[base[w] + rank] = t[i];
      }
      __syncwarp();
      if (ok && (same & ((1u << lane) - 1u)) == 0) base[w] += __popc(same);  // the run's first lane
      __syncwarp();
    }
    // base[w] now points one past the run of w
  }
  for (int w = lane; w < W; w += 32) {
    double v = w == 0 ? 0.0 : 1.0;  // CreateBaseJobInfo
    int cnt = 0;
    if (counted) cnt = cnts[w];
    else
      for (int i = r0; i < r1; i++) cnt += workers[i] == w;
    if (cnt > 0) {
      Cursor c{staged ? stage + (base[w] - cnt) : nullptr, workers, t, staged ? 0 : r0, r1, w};
      v = pairwise_sum(c, cnt) / (double)cnt;
    }
    et[w] = v;
  }
  __syncwarp();
  const double t1 = W > 1 ? et[1] : 0.0;
  for (int w = lane; w < W; w += 32) {
    const double s = w == 0 ? 0.0 : t1 / et[w];
    sp[w] = s;
    ef[w] = w == 0 ? 0.0 : s / (double)w;
  }
  if (lane == 0) out_status[job] = bad ? VODA_ST_BOUND : VODA_ST_OK;
}

}  // namespace

cudaError_t launch_curves(int jobs, const int32_t* rec_off, const int32_t* workers, const double* t, int W,
                          double* out_time, double* out_speedup, double* out_eff, int32_t* out_status,
                          cudaStream_t stream) {
  if (jobs <= 0) return cudaSuccess;
  const int perBlock = 4;  // a warp per job
  curves_kernel<<<(jobs + perBlock - 1) / perBlock, 32 * perBlock, 0, stream>>>(jobs, rec_off, workers, t, W, out_time,
                                                                                out_speedup, out_eff, out_status);
  return cudaGetLastError();
}

}  // namespace voda
