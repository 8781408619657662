// tick.cu — the scheduler's per-job bookkeeping between rounds, one thread per
// job, for all the jobs of many replicas at once:
//   voda_tick             updateTimeMetrics, pkg/scheduler/scheduler/scheduler.go:769-803: duration
//                         accounting by status, then Tiresias' demote / promote rules
//                         (pkg/algorithm/tiresias.go:27-36,109-119)
//   voda_compare_results  compareResults, scheduler.go:448-480: old vs new allocation -> the action
//                         the scheduler takes for the job
// Durations are time.Duration (int64 nanoseconds, wrapping arithmetic as in Go);
// Duration.Seconds() is float64(d / 1e9) + float64(d % 1e9) / 1e9 (time/time.go).
#include "common.cuh"
#include "kernels.h"

namespace voda {
namespace {

__device__ __forceinline__ double duration_seconds(long long d) {
  const long long sec = d / 1000000000ll, nsec = d % 1000000000ll;
  return (double)sec + (double)nsec / 1e9;
}

__global__ void tick_kernel(TickArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.jobs) return;
  const int status = a.status[i];
  const long long elapsed = a.now_ns - a.last_update_ns[i];  // time.Since(LastUpdateTime)
  const long long gpus = (long long)a.num_gpu[i];
  long long lastGpu = a.last_gpu_ns[i], lastRun = a.last_running_ns[i], lastWait = a.last_waiting_ns[i];
  if (status == VODA_JOB_RUNNING) {
    a.running_ns[i] += elapsed;
    a.gpu_ns[i] += elapsed * gpus;
    a.total_ns[i] += elapsed;
    lastRun += elapsed;
    lastGpu += elapsed * gpus;
    a.last_running_ns[i] = lastRun;
    a.last_gpu_ns[i] = lastGpu;
  } else if (status == VODA_JOB_WAITING) {
    a.waiting_ns[i] += elapsed;
    a.total_ns[i] += elapsed;
    lastWait += elapsed;
    a.last_waiting_ns[i] = lastWait;
  }
  a.last_update_ns[i] = a.now_ns;
  int changed = 0;
  if (a.tiresias && (status == VODA_JOB_RUNNING || status == VODA_JOB_WAITING)) {
    const int pr = a.priority[i];
    // TiresiasThresholdsSec[priority]: a Go map, a missing key reads 0
    const double thr = (pr >= 0 && pr < a.queue_num) ? a.thr[pr] : 0.0;
    if (duration_seconds(lastGpu) > thr) {
      a.priority[i] = pr < a.queue_num - 1 ? pr + 1 : pr;  // TiresiasDemotePriority
      changed = 1;  // the reference sets the flag even when the priority stays
    } else if (lastWait >= lastRun * (long long)a.promote_knob && pr > 0) {
      a.priority[i] = 0;  // TiresiasPromotePriority
      changed = 1;
    }
  }
  a.changed[i] = changed;
}

__global__ void compare_kernel(int jobs, const int32_t* __restrict__ old_gpu, const int32_t* __restrict__ new_gpu,
                               const int32_t* __restrict__ status, int32_t* __restrict__ action) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= jobs) return;
  const int n = old_gpu[i], m = new_gpu[i];
  int act = VODA_ACT_NONE;
  if (n > m) {
    if (m == 0) {  // don't delete a completed or failed job
      const int st = status[i];
      act = (st != VODA_JOB_COMPLETED && st != VODA_JOB_FAILED) ? VODA_ACT_HALT : VODA_ACT_NONE;
    } else {
      act = VODA_ACT_SCALE_IN;
    }
  } else if (n < m) {
    act = n == 0 ? VODA_ACT_START : VODA_ACT_SCALE_OUT;
  }
  action[i] = act;
}

}  // namespace

cudaError_t launch_tick(const TickArgs& a, cudaStream_t stream) {
  if (a.jobs <= 0) return cudaSuccess;
  tick_kernel<<<(a.jobs + 255) / 256, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_compare(int jobs, const int32_t* old_gpu, const int32_t* new_gpu, const int32_t* status,
                           int32_t* action, cudaStream_t stream) {
  if (jobs <= 0) return cudaSuccess;
  compare_kernel<<<(jobs + 255) / 256, 256, 0, stream>>>(jobs, old_gpu, new_gpu, status, action);
  return cudaGetLastError();
}

}  // namespace voda
