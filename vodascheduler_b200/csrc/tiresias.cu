// tiresias.cu — Tiresias (pkg/algorithm/tiresias.go:51-99) and ElasticTiresias
// (pkg/algorithm/elastic_tiresias.go:37-159) for a batch of rounds (sm_100a).
//
// One CTA per round, per-job state in shared memory; the speedup table is read in place (a
// round only looks up a few neighbouring entries per job).  Steps, with the reference lines they
// restate:
//   1. gain = S[min] / float32(min)                                   elastic_tiresias.go:58
//   2. queues by Priority, each stable-sorted by FirstStartTime       :54-69  -> ONE block
//      argsort on (queue, FirstStartTime, input index) (rank sort on 64-bit packed keys for
//      rounds up to 96 jobs whose times fit, bitonic otherwise); jobs whose priority is outside
//      [0, Q) sort last and are not visited by phases 3-4
//   3. give NumProc while it fits, queue by queue                     :75-85  -> warp skip-scan
//   4. compaction of queues >= 1 to MinNumProc when pendings > threshold   :89-102
//   5. prune satisfied / impossible jobs in input order               :107-115
//   6. greedy: +1 (or min for a job at 0) to the job with the largest marginal gain
//      S[r+1]-S[r] (float32 subtraction)                              :119-152
//      The reference re-sorts the live list twice per pick (stable by Priority, then stable by
//      gain descending) and takes the head.  Because only the head changes between sorts, the
//      head equals the maximum of (gain desc, Priority asc, stamp desc) with stamp = -(input
//      index) until a job is first picked and the pick number afterwards (SURVEY.md A.3/Q5;
//      checked on the CPU by oracle_dense.cpp against the literal double sort).  One warp runs
//      the picks: every lane caches the best of the jobs it owns, a shuffle arg-max finds the
//      winner, and only the owner lane rescans after the update.
//   7. validateResult                                                  :157
#include "common.cuh"
#include "kernels.h"

namespace voda {
namespace {

constexpr int kThreads = 128;     // upper bound; rounds of <= 64 jobs run with 64 threads
constexpr int kRankSortMax = 96;  // rounds up to this size use the rank sort
constexpr long long kMaxTimeSec = 0x7fffffffffffffffll - 62135596800ll;  // types.MaxTime (types.go:65)
constexpr int kMaxTimeNsec = 999999999;

struct TirSmem {
  int64_t* khi;
  long long* red;  // [4]
  int32_t* klo;
  int32_t* kq;     // queue index (Q for "no queue")
  int4* key4;      // [kRankSortMax] {queue, nsec, sec lo, sec hi}: one broadcast LDS.128 per compare
  unsigned long long* ukey;  // [Jp] the same key packed into 64 bits when every key of the round fits
  int* perm;
  int32_t* r;
  float* gain;
  int32_t* stamp;
  int32_t* prio;
  int32_t* mn;
  int32_t* mx;
  int32_t* need;   // NumProc in sorted order
  int* misc;       // [8]
  unsigned char* acc;
  unsigned char* live;
};

struct Cand {
  float gain;
  int prio, stamp, idx;
};

// (gain desc, Priority asc, stamp desc); idx < 0 = empty
__device__ __forceinline__ bool better(const Cand& a, const Cand& b) {
  if (a.idx < 0) return false;
  if (b.idx < 0) return true;
  if (a.gain > b.gain) return true;
  if (a.gain < b.gain) return false;
  if (a.prio != b.prio) return a.prio < b.prio;
  return a.stamp > b.stamp;
}

__global__ void __launch_bounds__(kThreads) tiresias_kernel(AllocArgs a, int Jp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TirSmem sm;
  {
    unsigned char* p = smem_raw;
    sm.key4 = reinterpret_cast<int4*>(p); p += 16 * kRankSortMax;
    sm.ukey = reinterpret_cast<unsigned long long*>(p); p += 8 * (size_t)Jp;
    sm.khi = reinterpret_cast<int64_t*>(p); p += 8 * (size_t)Jp;
    sm.red = reinterpret_cast<long long*>(p); p += 8 * 4;
    sm.klo = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.kq = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.perm = reinterpret_cast<int*>(p); p += 4 * (size_t)Jp;
    sm.r = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.gain = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.stamp = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.prio = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.mn = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.mx = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.need = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.misc = reinterpret_cast<int*>(p); p += 4 * 8;
    sm.acc = p; p += (size_t)Jp;
    sm.live = p; p += (size_t)Jp;
    p = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(p) + 15) & ~(uintptr_t)15);
  }
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nth = blockDim.x;
  const bool elastic = a.algo == VODA_ALGO_ELASTIC_TIRESIAS;
  const int Q = a.queue_num;
  const int W = a.stride;

  for (int rd = blockIdx.x; rd < a.rounds; rd += gridDim.x) {
    const int o = a.job_off[rd] - a.job_base;
    const int n = a.job_off[rd + 1] - a.job_off[rd];
    const int total = a.total_gpu[rd];
    if (n > a.max_jobs) {
      for (int i = tid; i < n; i += nth) a.out_gpus[o + i] = 0;
      if (tid == 0) a.out_status[rd] = VODA_ST_BOUND;
      continue;
    }
    __syncthreads();  // previous round's shared state is dead
    // per-job scalars.  The speedup table is NOT staged: a round looks up a handful of entries per
    // job (S[min], S[min+1], S[n], S[n+1], then two per GPU handed out), neighbours in the row, so
    // reading them in place touches one or two sectors per job where staging moved the whole
    // 34-wide row through shared memory (measured 1.3x to 3x slower on 32..1024 jobs per round).
    bool fits = true;  // (queue, FirstStartTime) of every job packs into 64 bits, order preserved
    for (int i = tid; i < n; i += nth) {
      const int pr = a.priority[o + i];
      sm.prio[i] = pr;
      sm.kq[i] = (pr >= 0 && pr < Q) ? pr : Q;
      sm.khi[i] = a.first_sec[o + i];
      sm.klo[i] = a.first_nsec ? a.first_nsec[o + i] : 0;
      if (n <= kRankSortMax)
        sm.key4[i] = make_int4(sm.kq[i], sm.klo[i], (int)(uint32_t)(uint64_t)sm.khi[i], (int)(sm.khi[i] >> 32));
      {
        // 2 bits of queue | 32 bits of seconds | 30 bits of nanoseconds.  types.MaxTime (the
        // never-started sentinel, far beyond 2^32 s) takes the seconds value 2^32-1, which no real
        // time may use; anything else out of range sends the round to the general compare.
        const long long sec = sm.khi[i];
        const int ns = sm.klo[i];
        const bool never = sec == kMaxTimeSec && ns == kMaxTimeNsec;
        const bool real = sec >= 0 && sec < 0xFFFFFFFFll && ns >= 0 && ns < (1 << 30);
        fits = fits && sm.kq[i] < 4 && (never || real);
        const unsigned long long sf = never ? 0xFFFFFFFFull : (unsigned long long)sec;
        sm.ukey[i] = ((unsigned long long)(unsigned)sm.kq[i] << 62) | (sf << 30) | (unsigned long long)(unsigned)ns;
      }
      sm.mn[i] = a.min_proc[o + i];
      sm.mx[i] = a.max_proc[o + i];
      sm.r[i] = 0;
    }
    if (tid == 0) {
      sm.misc[0] = 0;  // jobs inside a queue
      sm.misc[1] = 0;  // accepted in phase 3
      sm.misc[2] = 0;  // live jobs
      sm.misc[3] = 0;  // greedy loop cut: the reference would never leave it
    }
    const bool packedKeys = __syncthreads_and(fits) != 0;
    auto S = [&](int i, int g) -> float {  // Info.Speedup[fmt.Sprint(g)], map miss => 0
      if (g < 0 || g >= W) return 0.0f;
      // dense table, or a pool of curves the jobs point into (voda_alloc_batch.speedup_off)
      return a.speedup[(a.sp_off ? (size_t)a.sp_off[o + i] : (size_t)(o + i) * (size_t)W) + (size_t)g];
    };
    if (elastic)
      for (int i = tid; i < n; i += nth) sm.gain[i] = S(i, sm.mn[i]) / (float)sm.mn[i];  // :58
    // 2. queues
    {
      const int64_t* khi = sm.khi;
      const int32_t* klo = sm.klo;
      const int32_t* kq = sm.kq;
      auto less = [=](int x, int y) {
        if (y >= n) return x < n ? true : x < y;
        if (x >= n) return false;
        if (kq[x] != kq[y]) return kq[x] < kq[y];
        if (khi[x] != khi[y]) return khi[x] < khi[y];
        if (klo[x] != klo[y]) return klo[x] < klo[y];
        return x < y;
      };
      if (n <= kRankSortMax && packedKeys) {
        // rank sort on the 64-bit keys: rank(i) = #{j : u_j < u_i or (== and j < i)}
        //                                        = #{j : u_j < u_i + [j < i]}   (u_i + 1 cannot overflow)
        // The comparison (u_j, j) < (u_i, i) is the borrow of a three-word subtraction: four
        // carry-chain instructions per j (the 64-bit form with the [j < i] correction took nine).
        for (int i = tid; i < n; i += nth) {
          const uint2 ui = reinterpret_cast<const uint2*>(sm.ukey)[i];
          int negRank = 0;
#pragma unroll 8
          for (int j = 0; j < n; j++) {
            const uint2 uj = reinterpret_cast<const uint2*>(sm.ukey)[j];
            asm("{ .reg .u32 t;\n\t"
                "sub.cc.u32 t, %1, %2;\n\t"
                "subc.cc.u32 t, %3, %4;\n\t"
                "subc.cc.u32 t, %5, %6;\n\t"
                "subc.u32 %0, %0, 0; }"
                : "+r"(negRank)
                : "r"(j), "r"(i), "r"(uj.x), "r"(ui.x), "r"(uj.y), "r"(ui.y));
          }
          sm.perm[-negRank] = i;
        }
        __syncthreads();
      } else if (packedKeys) {
        // larger rounds: bitonic network over (64-bit key, slot) pairs
        const int P = next_pow2(n > 0 ? n : 1);
        for (int i = n + tid; i < P; i += nth) sm.ukey[i] = ~0ull;
        __syncthreads();
        block_sort_kv(sm.ukey, sm.perm, P);
      } else if (n <= kRankSortMax) {
        // rank sort on the packed records: rank(i) = #{j : key_j < key_i or (== and j < i)}
        for (int i = tid; i < n; i += nth) {
          const int4 ki = sm.key4[i];
          const long long si = ((long long)ki.w << 32) | (unsigned)ki.z;
          int rank = 0;
#pragma unroll 4
          for (int j = 0; j < n; j++) {
            const int4 kj = sm.key4[j];
            const long long sj = ((long long)kj.w << 32) | (unsigned)kj.z;
            bool lt;
            if (kj.x != ki.x) lt = kj.x < ki.x;
            else if (sj != si) lt = sj < si;
            else if (kj.y != ki.y) lt = kj.y < ki.y;
            else lt = j < i;
            rank += lt ? 1 : 0;
          }
          sm.perm[rank] = i;
        }
        __syncthreads();
      }
      else
        block_argsort_by(sm.perm, next_pow2(n > 0 ? n : 1), less);
    }
    int inq = 0;
    for (int t = tid; t < n; t += nth) {
      const int i = sm.perm[t];
      sm.need[t] = a.num_proc[o + i];
      inq += sm.kq[i] < Q;
    }
    for (int d = 16; d > 0; d >>= 1) inq += __shfl_xor_sync(0xffffffffu, inq, d);
    if (lane == 0 && inq) atomicAdd(&sm.misc[0], inq);
    __syncthreads();
    const int nq = sm.misc[0];
    // 3. basic portion
    if (warp == 0) {
      long long f = warp_skip_scan(sm.need, nq, (long long)total, sm.acc);
      if (lane == 0) sm.red[2] = f;
    }
    __syncthreads();
    long long free = sm.red[2];
    int accepted = 0;
    for (int t = tid; t < nq; t += nth) {
      if (sm.acc[t]) {
        const int i = sm.perm[t];
        const int np = sm.need[t];
        sm.r[i] = np;
        if (elastic) sm.gain[i] = S(i, np + 1) - S(i, np);
        accepted++;
      }
    }
    for (int d = 16; d > 0; d >>= 1) accepted += __shfl_xor_sync(0xffffffffu, accepted, d);
    if (lane == 0 && accepted) atomicAdd(&sm.misc[1], accepted);
    __syncthreads();
    if (!elastic) {
      // Tiresias: only jobs inside a queue are in the result map (tiresias.go:81-90)
      for (int i = tid; i < n; i += nth) {
        a.out_gpus[o + i] = sm.r[i];
        sm.live[i] = sm.kq[i] >= Q;  // "skip" flag for validateResult
      }
      __syncthreads();
      const int v = block_validate(a, o, n, total, a.out_gpus + o, sm.live, sm.red);
      if (tid == 0) a.out_status[rd] = v;
      continue;
    }
    const int pendings = n - sm.misc[1];
    // 4. compaction (:89-102)
    if (pendings > a.compaction) {
      long long freed = 0;
      for (int i = tid; i < n; i += nth) {
        const int q = sm.kq[i];
        if (q >= 1 && q < Q && sm.r[i] != 0) {
          freed += sm.r[i] - sm.mn[i];
          sm.r[i] = sm.mn[i];
          sm.gain[i] = S(i, sm.mn[i] + 1) - S(i, sm.mn[i]);
        }
      }
      for (int d = 16; d > 0; d >>= 1) freed += __shfl_xor_sync(0xffffffffu, freed, d);
      if (tid == 0) sm.red[3] = 0;
      __syncthreads();
      if (lane == 0 && freed) atomicAdd(reinterpret_cast<unsigned long long*>(&sm.red[3]), (unsigned long long)freed);
      __syncthreads();
      free += sm.red[3];
    }
    // 5. prune (:107-115); free is frozen here
    int nl = 0;
    for (int i = tid; i < n; i += nth) {
      const bool lv = !(sm.r[i] >= sm.mx[i] || free < (long long)sm.mn[i]);
      sm.live[i] = lv;
      sm.stamp[i] = -i;
      nl += lv;
    }
    for (int d = 16; d > 0; d >>= 1) nl += __shfl_xor_sync(0xffffffffu, nl, d);
    if (lane == 0 && nl) atomicAdd(&sm.misc[2], nl);
    __syncthreads();
    // 6. greedy picks by warp 0
    if (warp == 0) {
      int nlive = sm.misc[2];
      auto rescan = [&]() -> Cand {
        Cand b{0.0f, 0, 0, -1};
        for (int i = lane; i < n; i += 32) {
          if (!sm.live[i]) continue;
          Cand c{sm.gain[i], sm.prio[i], sm.stamp[i], i};
          if (better(c, b)) b = c;
        }
        return b;
      };
      Cand mine = rescan();
      int it = 0;
      while (free > 0 && nlive > 0) {
        Cand w = mine;
        for (int d = 16; d > 0; d >>= 1) {
          Cand c;
          c.gain = __shfl_xor_sync(0xffffffffu, w.gain, d);
          c.prio = __shfl_xor_sync(0xffffffffu, w.prio, d);
          c.stamp = __shfl_xor_sync(0xffffffffu, w.stamp, d);
          c.idx = __shfl_xor_sync(0xffffffffu, w.idx, d);
          if (better(c, w)) w = c;
        }
        if (w.idx < 0) break;
        it++;
        if (w.gain <= 0.0f) break;  // :128-134 no more gains
        const int j = w.idx;
        const int rj = sm.r[j], mnj = sm.mn[j], mxj = sm.mx[j];
        const bool owner = (j & 31) == lane;
        __syncwarp();  // every lane has read the winner's state before its owner rewrites it
        if (rj == 0) {
          if (free >= (long long)mnj) {
            if (mnj == 0) {
              // MinNumProc == 0: the pick leaves r, free and the live list as they were and sets
              // gain = S[1]-S[0] (:136-143).  If that is the gain the job already had, the next
              // iteration sees the identical state and picks it again: the reference never
              // returns.  Cut here (uniform across the warp: same shared values, same load).
              const float ng = S(j, 1) - S(j, 0);
              if (__float_as_int(ng) == __float_as_int(w.gain) && !(ng <= 0.0f)) {
                if (lane == 0) sm.misc[3] = 1;
                break;
              }
            }
            free -= mnj;
            if (owner) {
              sm.r[j] = mnj;
              sm.gain[j] = S(j, mnj + 1) - S(j, mnj);
              sm.stamp[j] = it;
            }
          } else {
            nlive--;
            if (owner) sm.live[j] = 0;
          }
        } else {
          free -= 1;
          const int nr = rj + 1;
          if (owner) {
            sm.r[j] = nr;
            sm.gain[j] = S(j, nr + 1) - S(j, nr);
            sm.stamp[j] = it;
          }
          if (nr >= mxj) {
            nlive--;
            if (owner) sm.live[j] = 0;
          }
        }
        __syncwarp();
        if (owner) mine = rescan();
      }
    }
    __syncthreads();
    for (int i = tid; i < n; i += nth) a.out_gpus[o + i] = sm.r[i];
    __syncthreads();
    if (sm.misc[3]) {  // validateResult (:157) is never reached
      if (tid == 0) a.out_status[rd] = VODA_ST_NONTERMINATING;
      continue;
    }
    const int v = block_validate(a, o, n, total, a.out_gpus + o, nullptr, sm.red);
    if (tid == 0) a.out_status[rd] = v;
  }
}

size_t tir_fixed_smem(int Jp) { return (size_t)Jp * (8 + 8 + 4 * 10 + 2) + 16 * kRankSortMax + 8 * 4 + 4 * 8 + 32; }

}  // namespace

size_t tiresias_smem_bytes(const AllocArgs& a, bool* table_in_smem) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  if (table_in_smem) *table_in_smem = false;  // the table is read in place
  return tir_fixed_smem(Jp);
}

cudaError_t launch_tiresias(const AllocArgs& a, int sms, cudaStream_t stream) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  const size_t smem = tiresias_smem_bytes(a, nullptr);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(tiresias_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int perSm = 0;
  const int threads = a.max_jobs <= 64 ? 64 : kThreads;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, tiresias_kernel, threads, smem);
  if (e != cudaSuccess) return e;
  if (perSm < 1) perSm = 1;
  int grid = sms * perSm;
  if (a.rounds < grid) grid = a.rounds;
  tiresias_kernel<<<grid, threads, smem, stream>>>(a, Jp);
  return cudaGetLastError();
}

}  // namespace voda
