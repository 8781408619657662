// noalloc.cu — voda_ffdl_allocate_static: the FfDL Optimizer batch kernel for callers that
// must not allocate device memory.
//
// Who needs this.  voda_allocate_dev sizes its scratch (the rows of the DP table the forward
// pass stores and the backtrack reads back) with cudaMalloc on first use and again whenever a bigger
// batch arrives.  Some hosts of this library cannot tolerate that:
//   * a scheduler that shares the GPU with a training framework whose caching allocator has
//     already claimed the HBM — a late cudaMalloc there simply fails;
//   * a control loop with a latency budget — cudaMalloc / cudaFree synchronise the whole
//     device and take an unbounded time, which stalls every other stream of the process;
//   * code that runs where the CUDA allocator must not be entered at all (inside another
//     library's callback, or while the process is deliberately kept allocation-free after
//     start-up so its memory footprint is fixed and auditable).
// For them the scratch lives in the fixed array below: it is part of the loaded module, so
// every GPU that loads this library holds VODA_STATIC_SCRATCH_BYTES of it until the process
// exits, and no call here ever reaches cudaMalloc, cudaFree or any other allocating routine
// (no torch, no std::string, no staging buffers: inputs and outputs are the caller's device
// pointers).
//
// Limits and sharing.  The array is 32 MiB.  One round needs a slab of about
// min(max_jobs, max_total_gpu) * (ceil(max_total_gpu / 8) * 32 + 4 * W + 16) bytes; the batch is walked in
// passes of as many rounds as fit (so it is slower than voda_allocate_dev on big batches),
// and a batch whose single round does not fit is rejected with VODA_E_TOO_LARGE.  All calls on
// a GPU share the one array and it keeps whatever the previous call left in it.  Nothing has
// to start at zero: the backtrack only uses cells, row records and S entries the forward pass of
// the same round has written (what it loads beyond them is masked), so no clearing is done.  Two calls that overlap on the GPU (different streams, or
// different host threads) would write the same slabs and race — serialise them.
//
// Set-up.  In host code the name `g_static_scratch` is only a stand-in for the device array,
// not an address a kernel can use, so the first call on each device (with that device
// current) asks the runtime for the real device address with cudaGetSymbolAddress, checks
// the result and remembers it per device.  Make that first call before starting any stream
// capture on the device.
#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "kernels.h"

#define VODA_STATIC_SCRATCH_BYTES (32u << 20)

static __device__ __align__(256) unsigned char g_static_scratch[VODA_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
void* g_scratch_addr[kMaxDevices];  // device address of g_static_scratch, per device
int g_sm_count[kMaxDevices];

int report(char* errbuf, size_t errlen, int code, const char* msg, const char* detail = nullptr) {
  if (errbuf && errlen) std::snprintf(errbuf, errlen, "%s%s%s", msg, detail ? ": " : "", detail ? detail : "");
  return code;
}

}  // namespace

extern "C" int voda_ffdl_allocate_static(void* stream, const voda_alloc_batch* b, int32_t* d_out_gpus,
                                         int32_t* d_out_status, char* errbuf, size_t errbuf_len) {
  using namespace voda;
  if (!b || !d_out_gpus || !d_out_status) return report(errbuf, errbuf_len, VODA_E_INVALID, "null argument");
  if (b->rounds < 0) return report(errbuf, errbuf_len, VODA_E_INVALID, "negative round count");
  if (b->rounds == 0) return VODA_OK;
  if (!b->job_off || !b->total_gpu || !b->submit_sec || !b->min_proc || !b->max_proc || !b->speedup ||
      b->speedup_stride < 1)
    return report(errbuf, errbuf_len, VODA_E_INVALID, "FfDL needs job_off, total_gpu, submit_sec, min/max_proc and the speedup table");
  if (b->max_jobs <= 0 || b->max_total_gpu <= 0)
    return report(errbuf, errbuf_len, VODA_E_INVALID, "max_jobs and max_total_gpu must be declared (device-resident batch)");

  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return report(errbuf, errbuf_len, VODA_E_CUDA, "cudaGetDevice", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return report(errbuf, errbuf_len, VODA_E_UNSUPPORTED, "device index beyond the per-device table");
  void* scratch = nullptr;
  int sms = 0;
  {
    // once per device, with that device current: resolve the array's device address
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_addr[dev]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_static_scratch);
      if (e != cudaSuccess || !p)
        return report(errbuf, errbuf_len, VODA_E_CUDA, "cudaGetSymbolAddress(g_static_scratch)", cudaGetErrorString(e));
      int n = 0;
      e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
      if (e != cudaSuccess) return report(errbuf, errbuf_len, VODA_E_CUDA, "cudaDeviceGetAttribute", cudaGetErrorString(e));
      g_sm_count[dev] = n;
      g_scratch_addr[dev] = p;
    }
    scratch = g_scratch_addr[dev];
    sms = g_sm_count[dev];
  }

  AllocArgs a{};
  a.algo = VODA_ALGO_FFDL_OPTIMIZER;
  a.rounds = b->rounds;
  a.job_base = 0;
  a.job_off = b->job_off;
  a.total_gpu = b->total_gpu;
  a.submit_sec = b->submit_sec;
  a.submit_nsec = b->submit_nsec;
  a.min_proc = b->min_proc;
  a.max_proc = b->max_proc;
  a.speedup = b->speedup;
  a.sp_off = b->speedup_off;  // a pool of curves, if the caller uses one
  a.stride = b->speedup_stride;
  a.queue_num = 2;
  a.compaction = 10;
  a.max_jobs = b->max_jobs;
  a.max_total = b->max_total_gpu;
  a.max_gmax = b->max_max_proc > 0 ? b->max_max_proc : 0;
  a.out_gpus = d_out_gpus;
  a.out_status = d_out_status;

  if (ffdl_smem_bytes(a) > 200 * 1024)
    return report(errbuf, errbuf_len, VODA_E_UNSUPPORTED, "total_gpu / jobs too large for one CTA's shared memory");
  const size_t per = (ffdl_scratch_per_round(a) + 255) & ~(size_t)255;  // bytes one round needs
  if (per > VODA_STATIC_SCRATCH_BYTES) {
    char detail[160];
    std::snprintf(detail, sizeof detail, "one round needs %zu bytes of scratch, the fixed array holds %u", per,
                  (unsigned)VODA_STATIC_SCRATCH_BYTES);
    return report(errbuf, errbuf_len, VODA_E_TOO_LARGE, "batch shape exceeds the static scratch", detail);
  }
  // as many rounds per pass as the array holds; each pass is a DP launch + a backtrack launch
  const int fit = (int)(VODA_STATIC_SCRATCH_BYTES / per);
  for (int r0 = 0; r0 < a.rounds; r0 += fit) {
    const int count = a.rounds - r0 < fit ? a.rounds - r0 : fit;
    e = launch_ffdl(a, r0, count, sms, scratch, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return report(errbuf, errbuf_len, VODA_E_CUDA, "FfDL kernel launch", cudaGetErrorString(e));
  }
  return VODA_OK;
}
