// fill.cu — FIFO, SRJF, ElasticFIFO and ElasticSRJF for a batch of rounds (sm_100a).
//   FIFO         pkg/algorithm/fifo.go:25-52          ElasticFIFO  elastic_fifo.go:25-75
//   SRJF         pkg/algorithm/srjf.go:25-52          ElasticSRJF  elastic_srjf.go:25-72
//
// All four share: stable sort (SubmitTime, or EstimatedRemainningTimeSec with float32 '<'),
// then phase one "give MinNumProc while it fits" in sorted order.  The elastic pair adds the
// round-robin phase "+1 per eligible job per pass until no GPU is left or every job is
// satisfied" (eligible: r < max OR not satisfied — an OR, elastic_fifo.go:59).  Integer-only
// arithmetic; one CTA per round, per-job state in shared memory.
//
// Phase one is a serial dependence on `free` (a job that does not fit is skipped and a later,
// smaller one may fit), so it is evaluated by one warp, 32 sorted jobs at a time: prefix sums
// find the first job that fails given that all before it were accepted; it is rejected and
// the rest of the chunk is re-evaluated.  Exact for any sign of MinNumProc.
//
// Phase two is evaluated in closed form per "epoch": with E eligible jobs, U unsatisfied
// ones and d = min (max - r) over eligible jobs that can still reach their max, the next
// t = min(d, free / E) passes all hand +1 to the same E jobs and no flag changes before the
// end of pass t, so they collapse into r += t.  When free < E the last, partial pass gives +1
// to the first `free` eligible jobs in sorted order.  This reproduces the reference's result
// including its quirks, which simply run until free == 0.
#include "common.cuh"
#include "kernels.h"

namespace voda {
namespace {

constexpr int kThreads = 256;
constexpr int kBig = 0x7fffffff;

struct FillSmem {
  int64_t* khi;
  long long* red;   // [4]
  int32_t* klo;
  int* perm;
  int32_t* r;       // allocation, sorted order
  int32_t* mn;
  int32_t* mx;
  int* cnt;         // [kThreads/32 + 1] scan scratch
  unsigned char* sat;
};

__global__ void __launch_bounds__(kThreads) fill_kernel(AllocArgs a, int Jp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FillSmem sm;
  {
    unsigned char* p = smem_raw;
    sm.khi = reinterpret_cast<int64_t*>(p); p += sizeof(int64_t) * (size_t)Jp;
    sm.red = reinterpret_cast<long long*>(p); p += sizeof(long long) * 4;
    sm.klo = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.perm = reinterpret_cast<int*>(p); p += 4 * (size_t)Jp;
    sm.r = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.mn = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.mx = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.cnt = reinterpret_cast<int*>(p); p += 4 * (kThreads / 32 + 4);
    sm.sat = p;
  }
  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const bool byRemaining = a.algo == VODA_ALGO_SRJF || a.algo == VODA_ALGO_ELASTIC_SRJF;
  const bool elastic = a.algo == VODA_ALGO_ELASTIC_FIFO || a.algo == VODA_ALGO_ELASTIC_SRJF;
  const bool markEqMax = a.algo == VODA_ALGO_ELASTIC_FIFO;  // elastic_fifo.go:46-48 only

  for (int r = blockIdx.x; r < a.rounds; r += gridDim.x) {
    const int o = a.job_off[r] - a.job_base;
    const int n = a.job_off[r + 1] - a.job_off[r];
    const int total = a.total_gpu[r];
    if (n > a.max_jobs) {
      for (int i = tid; i < n; i += kThreads) a.out_gpus[o + i] = 0;
      if (tid == 0) a.out_status[r] = VODA_ST_BOUND;
      continue;
    }
    // 1. keys + stable sort.  The float32 remaining time always packs into a 64-bit key, a
    //    SubmitTime does when it lies in [0, 2^32) s (else the round takes the general argsort).
    const int P = next_pow2(n > 0 ? n : 1);
    unsigned long long* ukey = reinterpret_cast<unsigned long long*>(sm.khi);
    bool fits = true;
    for (int i = tid; i < P; i += kThreads) {
      if (i >= n) {
        ukey[i] = ~0ull;
      } else if (byRemaining) {
        ukey[i] = (unsigned long long)(f32_order_key(a.remaining_sec[o + i]) + (1ll << 31));
      } else {
        const long long sec = a.submit_sec[o + i];
        const int ns = a.submit_nsec ? a.submit_nsec[o + i] : 0;
        unsigned long long u;
        fits = pack_time_key(sec, ns, &u) && fits;
        ukey[i] = u;
      }
    }
    if (__syncthreads_and(fits)) {
      block_sort_kv(ukey, sm.perm, P);
    } else {  // (only SubmitTime keys get here)
      for (int i = tid; i < n; i += kThreads) {
        sm.khi[i] = a.submit_sec[o + i];
        sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
      }
      __syncthreads();
      block_argsort(sm.khi, sm.klo, sm.perm, n, P);
    }
    for (int i = tid; i < n; i += kThreads) {
      const int j = o + sm.perm[i];
      sm.mn[i] = a.min_proc[j];
      sm.mx[i] = a.max_proc[j];
    }
    __syncthreads();
    // 2. phase one (warp 0)
    if (warp == 0) {
      long long f1 = warp_skip_scan(sm.mn, n, (long long)total, sm.sat);  // sat[] holds "accepted" for now
      if (lane == 0) sm.red[2] = f1;
    }
    __syncthreads();
    long long free = sm.red[2];
    for (int i = tid; i < n; i += kThreads) {
      const bool accepted = sm.sat[i];
      sm.r[i] = accepted ? sm.mn[i] : 0;
      // not accepted => "unable to allocate minGPU": satisfied (elastic_fifo.go:49-51)
      sm.sat[i] = accepted ? (unsigned char)(markEqMax && sm.mn[i] == sm.mx[i]) : (unsigned char)1;
    }
    __syncthreads();
    // 3. phase two, epoch by epoch
    if (elastic) {
      while (true) {
        // E, U, dmin
        int E = 0, U = 0, dmin = kBig;
        for (int i = tid; i < n; i += kThreads) {
          const int ri = sm.r[i], mi = sm.mx[i];
          const bool st = sm.sat[i];
          const bool el = ri < mi || !st;
          E += el;
          U += !st;
          if (el && ri < mi) dmin = min(dmin, mi - ri);
        }
        for (int d = 16; d > 0; d >>= 1) {
          E += __shfl_xor_sync(0xffffffffu, E, d);
          U += __shfl_xor_sync(0xffffffffu, U, d);
          dmin = min(dmin, __shfl_xor_sync(0xffffffffu, dmin, d));
        }
        if (tid == 0) {
          sm.cnt[kThreads / 32] = 0;      // E
          sm.cnt[kThreads / 32 + 1] = 0;  // U
          sm.cnt[kThreads / 32 + 2] = kBig;
        }
        __syncthreads();
        if (lane == 0) {
          atomicAdd(&sm.cnt[kThreads / 32], E);
          atomicAdd(&sm.cnt[kThreads / 32 + 1], U);
          atomicMin(&sm.cnt[kThreads / 32 + 2], dmin);
        }
        __syncthreads();
        E = sm.cnt[kThreads / 32];
        U = sm.cnt[kThreads / 32 + 1];
        dmin = sm.cnt[kThreads / 32 + 2];
        __syncthreads();
        if (free <= 0 || U == 0) break;  // for freeGPU > 0 && !allTrue(sastified)
        long long t = free / E;          // E >= U > 0
        if (dmin != kBig && t > dmin) t = dmin;
        if (t >= 1) {
          for (int i = tid; i < n; i += kThreads) {
            const int ri = sm.r[i], mi = sm.mx[i];
            if (ri < mi || !sm.sat[i]) {
              const int nr = (int)(ri + t);
              sm.r[i] = nr;
              if (nr == mi) sm.sat[i] = 1;
            }
          }
          free -= t * (long long)E;
          __syncthreads();
          continue;
        }
        // partial pass: the first `free` eligible jobs (sorted order) get +1, then free == 0
        int baseCount = 0;
        for (int tile = 0; tile < n; tile += kThreads) {
          const int i = tile + tid;
          bool el = false;
          if (i < n) el = sm.r[i] < sm.mx[i] || !sm.sat[i];
          const unsigned bal = __ballot_sync(0xffffffffu, el);
          if (lane == 0) sm.cnt[warp] = __popc(bal);
          __syncthreads();
          int before = baseCount, tileTotal = 0;
          for (int w = 0; w < kThreads / 32; w++) {
            const int c = sm.cnt[w];
            if (w < warp) before += c;
            tileTotal += c;
          }
          const int rank = before + __popc(bal & ((1u << lane) - 1u));
          if (el && (long long)rank < free) {
            const int nr = sm.r[i] + 1;
            sm.r[i] = nr;
            if (nr == sm.mx[i]) sm.sat[i] = 1;
          }
          baseCount += tileTotal;
          __syncthreads();
        }
        free = 0;
        break;
      }
    }
    __syncthreads();
    // 4. scatter to input order, validate
    for (int i = tid; i < n; i += kThreads) a.out_gpus[o + sm.perm[i]] = sm.r[i];
    __syncthreads();
    const int v = block_validate(a, o, n, total, a.out_gpus + o, nullptr, sm.red);
    if (tid == 0) a.out_status[r] = v;
  }
}

size_t fill_smem(int Jp) {
  return (size_t)Jp * (8 + 4 * 5 + 1) + 8 * 4 + 4 * (kThreads / 32 + 4) + 16;
}

}  // namespace

size_t fill_smem_bytes(const AllocArgs& a) { return fill_smem(next_pow2(a.max_jobs > 0 ? a.max_jobs : 1)); }

cudaError_t launch_fill(const AllocArgs& a, int sms, cudaStream_t stream) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  const size_t smem = fill_smem(Jp);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int grid = sms * 8;
  if (a.rounds < grid) grid = a.rounds;
  fill_kernel<<<grid, kThreads, smem, stream>>>(a, Jp);
  return cudaGetLastError();
}

}  // namespace voda
