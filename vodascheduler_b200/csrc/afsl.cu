// afsl.cu — AFSL.Schedule / topPriority / jobLength / evaluate (pkg/algorithm/afsl.go:31-106)
// for a batch of rounds (sm_100a).
//
// The reference hands out GPUs one at a time; each pick is a LEFT FOLD over the live jobs in
// SubmitTime order with a pairwise, state-dependent and non-transitive rule (afsl.go:72-92):
//     champion j, challenger jb
//     both at 0 workers:      j = jb  if rem_j >= rem_jb
//     otherwise:              lenA = len(j, r_j); lenB = len(jb, r_j)      <- champion's count for
//                             if lenA >= lenB: swap(j, jb)                    BOTH (bug kept, :82)
//                             if evaluate(j, jb): j = jb
//     len(x, w) = +Inf if w == 0 else float64(rem_x / S_x[w])               (float32 divide)
//     evaluate(short, long) = (S_l[r_l+1]-S_l[r_l]) / S_l[r_l+1]  >  (S_s[r_s+1]-S_s[r_s]) / S_s[r_s]
// A tree reduction would change results, so the fold order is kept exactly:
//   * one warp walks the live list 32 challengers at a time against the current champion; a
//     ballot finds the FIRST challenger that replaces it, that lane broadcasts its whole state
//     by shuffle and becomes the champion, and only the lanes after it re-evaluate
//     (speculative first-beat scan);
//   * the champion entering every position is cached; after the winner at position p gets its
//     GPU the next fold restarts at p with the cached champion (positions before p never
//     involve the winner, so their outcome cannot change);
//   * the divisions are hoisted: len(x, w) = rem_x / S_x[w] (the float32 quotient the reference
//     recomputes every step) is tabulated once per round for the HOT worker counts
//     w = 1 .. CH in shared memory, column-major over the jobs in sorted order so that 32
//     neighbouring challengers read 32 neighbouring banks; larger w (rare: the champion's count
//     is small) divide on the fly from the table in HBM; the two evaluate() ratios of a job are
//     recomputed only when its count changes.
// The dependent chain (picks x fold steps) is the cost; throughput comes from rounds in flight,
// so shared memory per round is kept small (32 bytes per job + CH * 4) to fit 4 CTAs per SM.
// Jobs are relabelled by their sorted position; removal at max keeps order (lazy deletion +
// periodic compaction of the live list).
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"
#include "afsl_warp.cuh"

namespace voda {
namespace {

constexpr int kThreads = 128;
constexpr int kWarps = kThreads / 32;
constexpr unsigned short kDead = 0xFFFF;

struct AfslSmem {
  // indexed by sorted position ("sr")
  float* rem;
  float* L;    // (S[r+1]-S[r]) / S[r+1]   "as the longer job"
  float* Rr;   // (S[r+1]-S[r]) / S[r]     "as the shorter job"
  float* self; // len(job, r[job]): its own length at its own count (the champion's lenA)
  int32_t* r;
  int32_t* mx;
  unsigned short* order;       // live list: sr at position p, kDead once removed
  unsigned short* champIn;     // champion entering each position ...
  unsigned short* champInPos;  // ... and the position it sits at
  unsigned short* perm;        // sr -> job index within the round
  float* lenT;                 // [CH][Jp]: len(sr, w), w = 1..CH
  long long* red;              // [4]
  int* sbeat;                  // [2][kWarps] first beating position per warp (double-buffered) + 1
  float* schamp;               // [6] state of a job that just became the champion
  // sort scratch, aliased over rem/L/Rr/r (16 bytes per job)
  int64_t* khi;
  int32_t* klo;
  int* perm32;
};

__global__ void __launch_bounds__(kThreads) afsl_kernel(AllocArgs a, int Jp, int CH) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  AfslSmem sm;
  {
    unsigned char* p = smem_raw;
    sm.red = reinterpret_cast<long long*>(p); p += 8 * 4;
    sm.sbeat = reinterpret_cast<int*>(p); p += 4 * (2 * kWarps + 2);
    sm.schamp = reinterpret_cast<float*>(p); p += 4 * 8;
    sm.khi = reinterpret_cast<int64_t*>(p);
    sm.klo = reinterpret_cast<int32_t*>(p + 8 * (size_t)Jp);
    sm.perm32 = reinterpret_cast<int*>(p + 12 * (size_t)Jp);
    sm.rem = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.L = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.Rr = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.r = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.mx = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.self = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.lenT = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp * (size_t)CH;
    sm.order = reinterpret_cast<unsigned short*>(p); p += 2 * (size_t)Jp;
    sm.champIn = reinterpret_cast<unsigned short*>(p); p += 2 * (size_t)Jp;
    sm.champInPos = reinterpret_cast<unsigned short*>(p); p += 2 * (size_t)Jp;
    sm.perm = reinterpret_cast<unsigned short*>(p);
  }
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int W = a.stride;
  const unsigned full = 0xffffffffu;
  const float INF = __int_as_float(0x7f800000);

  for (int rd = blockIdx.x; rd < a.rounds; rd += gridDim.x) {
    const int o = a.job_off[rd] - a.job_base;
    const int n = a.job_off[rd + 1] - a.job_off[rd];
    const int total = a.total_gpu[rd];
    if (n > a.max_jobs) {
      for (int i = tid; i < n; i += kThreads) a.out_gpus[o + i] = 0;
      if (tid == 0) a.out_status[rd] = VODA_ST_BOUND;
      continue;
    }
    __syncthreads();
    auto S = [&](int job, int g) -> float {  // Info.Speedup[fmt.Sprint(g)], map miss => 0
      if (g < 0 || g >= W) return 0.0f;
      return a.speedup[(a.sp_off ? (size_t)a.sp_off[o + job] : (size_t)(o + job) * (size_t)W) + (size_t)g];
    };
    // ---- stable sort by SubmitTime (:42-44), then relabel every job by its sorted position
    for (int i = tid; i < n; i += kThreads) {
      sm.khi[i] = a.submit_sec[o + i];
      sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
    }
    __syncthreads();
    block_argsort(sm.khi, sm.klo, sm.perm32, n, next_pow2(n > 0 ? n : 1));
    for (int t = tid; t < n; t += kThreads) sm.perm[t] = (unsigned short)sm.perm32[t];
    __syncthreads();  // the sort scratch is dead from here on
    for (int t = tid; t < n; t += kThreads) {
      const int j = sm.perm[t];
      sm.rem[t] = a.remaining_sec[o + j];
      sm.r[t] = 0;
      sm.mx[t] = a.max_proc[o + j];
      const float s0 = S(j, 0), s1 = S(j, 1);
      const float d = s1 - s0;
      sm.L[t] = d / s1;
      sm.Rr[t] = d / s0;
      sm.self[t] = INF;  // len(x, 0)
      sm.order[t] = (unsigned short)t;
    }
    for (int x = tid; x < n * CH; x += kThreads) {
      const int t = x / CH, w = x - t * CH + 1;
      const int j = sm.perm[t];
      sm.lenT[(size_t)(w - 1) * Jp + t] = a.remaining_sec[o + j] / S(j, w);  // afsl.go:94-100
    }
    __syncthreads();

    {
      // len(x, w): hot columns from shared memory, the rest divided on the fly (w >= W is a
      // map miss: rem / 0)
      auto LEN = [&](int x, int w) -> float {
        if (w == 0) return INF;
        if (w <= CH) return sm.lenT[(size_t)(w - 1) * Jp + x];
        return sm.rem[x] / S(sm.perm[x], w);
      };
      // Every thread of the CTA carries the same scalar state (computed from the same shared
      // memory values), so the control flow below is uniform across the CTA.
      int len = n;        // entries in order[] (live + dead)
      int dead = 0;
      int head = 0;       // first live position
      int startPos = 0;   // fold restart: position of the next challenger
      int champ = -1;     // champion entering startPos (-1: seed from head) ...
      int champPos = -1;  // ... and where it sits
      long long free = total;
      int nlive = n;
      int parity = 0;
      while (free > 0 && nlive > 0) {  // :47
        // ---- topPriority: fold from startPos ------------------------------------------
        if (champ < 0) {
          while (head < len && sm.order[head] == kDead) head++;
          champ = sm.order[head];
          champPos = head;
          startPos = head + 1;
        }
        // champion state, uniform across the CTA
        int c = champ, cp = champPos;
        int rc = sm.r[c];
        float remc = sm.rem[c], Lc = sm.L[c], Rc = sm.Rr[c];
        float lenA = sm.self[c];
        // SPECULATION ACROSS WARPS: the kThreads positions from `pos` on are all tested against
        // the champion entering `pos`.  On this workload the champion changes about once per
        // pick while a fold is ~10 chunks of 32 long, so most groups pass without a change and
        // the fold advances kThreads positions per step; when some position does beat the
        // champion, everything before it is final, that job becomes the champion and the
        // fold resumes right behind it — exactly the left fold, evaluated in wider steps.
        int pos = startPos;
        while (pos < len) {
          const int p = pos + tid;
          const int jb = p < len ? (int)sm.order[p] : (int)kDead;
          const bool pend = jb != kDead;
          int rj = 0;
          float remj = 0.f, Lj = 0.f, Rj = 0.f;
          bool beat = false;
          if (pend) {
            rj = sm.r[jb];
            remj = sm.rem[jb];
            Lj = sm.L[jb];
            Rj = sm.Rr[jb];
            if (rc == 0 && rj == 0) {
              beat = remc >= remj;  // :77-79 later job wins ties
            } else {
              const float lenB = LEN(jb, rc);  // champion's count for both (:82)
              if (lenA >= lenB) {
                // swap: short = challenger, long = champion; champion survives iff left > right
                beat = !(Lc > Rj);
              } else {
                beat = Lj > Rc;  // short = champion, long = challenger
              }
            }
          }
          const unsigned m = __ballot_sync(full, beat);
          if (lane == 0) sm.sbeat[parity * kWarps + warp] = m ? pos + 32 * warp + __ffs(m) - 1 : 0x7fffffff;
          __syncthreads();
          int first = 0x7fffffff;
#pragma unroll
          for (int w = 0; w < kWarps; w++) first = min(first, sm.sbeat[parity * kWarps + w]);
          parity ^= 1;
          if (pend && p <= first) {  // the champion entering these positions is final
            sm.champIn[p] = (unsigned short)c;
            sm.champInPos[p] = (unsigned short)cp;
          }
          if (first == 0x7fffffff) {
            pos += kThreads;
            continue;
          }
          if (p == first) {  // this job takes over: publish its state
            sm.schamp[0] = __int_as_float(jb);
            sm.schamp[1] = __int_as_float(rj);
            sm.schamp[2] = remj;
            sm.schamp[3] = Lj;
            sm.schamp[4] = Rj;
            sm.schamp[5] = sm.self[jb];
          }
          __syncthreads();
          c = __float_as_int(sm.schamp[0]);
          cp = first;
          rc = __float_as_int(sm.schamp[1]);
          remc = sm.schamp[2];
          Lc = sm.schamp[3];
          Rc = sm.schamp[4];
          lenA = sm.schamp[5];
          pos = first + 1;
        }
        // ---- give it one GPU (:49-57) ----------------------------------------------------
        const int w = c, wpos = cp;
        const int nr = rc + 1;
        free -= 1;
        const bool removed = nr >= sm.mx[w];
        __syncthreads();  // champIn of this fold is complete; everyone has read the old state of w
        if (tid == 0) {
          sm.r[w] = nr;
          const int job = sm.perm[w];
          const float s0 = S(job, nr), s1 = S(job, nr + 1);
          const float d = s1 - s0;
          sm.L[w] = d / s1;
          sm.Rr[w] = d / s0;
          sm.self[w] = sm.rem[w] / s0;  // len(w, nr) = rem / S[nr] (afsl.go:98)
          if (removed) sm.order[wpos] = kDead;
        }
        if (removed) {
          nlive--;
          dead++;
        }
        // ---- where the next fold restarts: positions before wpos never involved the winner,
        // so the champion that entered wpos is still right ---------------------------------
        if (wpos <= head) {
          champ = -1;  // the seed itself changed: start over from the first live job
        } else {
          champ = sm.champIn[wpos];
          champPos = sm.champInPos[wpos];
          startPos = removed ? wpos + 1 : wpos;
        }
        __syncthreads();
        // ---- compaction of the live list (warp 0; the condition is uniform) -----------------
        if (dead >= 64 || (dead > 0 && dead * 4 >= len)) {
          if (warp == 0) {
            int outp = 0;
            for (int base = 0; base < len; base += 32) {
              const int p = base + lane;
              const int v = p < len ? (int)sm.order[p] : (int)kDead;
              const unsigned keep = __ballot_sync(full, v != kDead);
              const int dst = outp + __popc(keep & ((1u << lane) - 1u));
              __syncwarp();
              if (v != kDead) sm.order[dst] = (unsigned short)v;
              outp += __popc(keep);
              __syncwarp();
            }
            if (lane == 0) sm.sbeat[2 * kWarps] = outp;
          }
          __syncthreads();
          len = sm.sbeat[2 * kWarps];
          dead = 0;
          head = 0;
          champ = -1;  // positions moved: cached champions are void
          __syncthreads();
        }
      }
    }
    __syncthreads();
    for (int t = tid; t < n; t += kThreads) a.out_gpus[o + sm.perm[t]] = sm.r[t];
    __syncthreads();
    const int v = block_validate(a, o, n, total, a.out_gpus + o, nullptr, sm.red);  // deferred :39
    if (tid == 0) a.out_status[rd] = v;
  }
}

// shared memory: 32 bytes per job + the hot len() columns
size_t afsl_fixed_smem(int Jp) { return (size_t)Jp * 32 + 8 * 4 + 4 * (2 * kWarps + 2) + 4 * 8 + 32; }

int afsl_hot_columns(const AllocArgs& a, int Jp) {
  // as many columns as still leave room for 4 CTAs per SM (about 55 KB each), at most W - 1
  const long budget = 55 * 1024;
  long ch = (budget - (long)afsl_fixed_smem(Jp)) / (4L * Jp);
  if (ch < 0) ch = 0;
  if (ch > a.stride - 1) ch = a.stride - 1;
  return (int)ch;
}

}  // namespace

// Rounds of up to 1024 jobs run with one warp per round (afsl_warp.cuh); larger rounds with the
// four-warp kernel above.
static int afsl_warp_columns(const AllocArgs& a, int Jp) {
  // columns of 16-bit ranks (column 0 = rem, w = len(., w)): as many as keep the CTA at about
  // 27 bytes of shared memory per job slot (eight 1024-job rounds per SM), more for small rounds
  // (the SM holds at most 32 CTAs anyway), never more than the table is wide
  long target = 27300L * Jp / 1024;
  if (target < 7000) target = 7000;
  long nc = (target - (long)Jp * 6 - 32) / (2L * (Jp + 32));
  if (nc < 1) nc = 1;
  if (const char* v = std::getenv("VODA_AFSL_NC")) nc = std::atoi(v);
  if (nc > a.stride) nc = a.stride;
  if (nc < 1) nc = 1;
  return (int)nc;
}

size_t afsl_smem_bytes(const AllocArgs& a) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  if (Jp <= 1024 && a.max_total <= 0xFFFF) {
    const int Jw = Jp < 32 ? 32 : Jp;
    return afslw::smem_bytes(Jw, afsl_warp_columns(a, Jw));
  }
  return afsl_fixed_smem(Jp) + (size_t)afsl_hot_columns(a, Jp) * 4 * (size_t)Jp;
}

// Resident CTAs per SM of the one-warp kernel for this shape (a round is one long dependent chain, so a
// launch takes as long as its fullest CTA: batches are best cut into multiples of sms * this).
template <int SLOTS>
static cudaError_t afsl_warp_residency(const AllocArgs& a, int* perSm, size_t* smemOut, int* ncOut) {
  const int Jp = 32 * SLOTS;
  const int NC = afsl_warp_columns(a, Jp);
  const size_t smem = afslw::smem_bytes(Jp, NC);
  const void* kern = a.sp_off ? (const void*)afslw::afsl_warp_kernel<SLOTS, true> : (const void*)afslw::afsl_warp_kernel<SLOTS, false>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(perSm, kern, 32, smem);
  if (e != cudaSuccess) return e;
  if (*perSm < 1) *perSm = 1;
  if (const char* v = std::getenv("VODA_AFSL_CTAS")) *perSm = std::max(1, std::min(*perSm, std::atoi(v)));
  if (smemOut) *smemOut = smem;
  if (ncOut) *ncOut = NC;
  return cudaSuccess;
}

template <int SLOTS>
static cudaError_t launch_afsl_warp(const AllocArgs& a, int sms, cudaStream_t stream) {
  int perSm = 0, NC = 0;
  size_t smem = 0;
  cudaError_t e = afsl_warp_residency<SLOTS>(a, &perSm, &smem, &NC);
  if (e != cudaSuccess) return e;
  int grid = sms * perSm;
  if (a.rounds < grid) grid = a.rounds;
  if (a.sp_off) afslw::afsl_warp_kernel<SLOTS, true><<<grid, 32, smem, stream>>>(a, NC);
  else afslw::afsl_warp_kernel<SLOTS, false><<<grid, 32, smem, stream>>>(a, NC);
  return cudaGetLastError();
}

int afsl_wave_rounds(const AllocArgs& a, int sms) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  if (a.max_total > 0xFFFF || Jp > 1024) return 0;
  int perSm = 0;
  cudaError_t e = Jp <= 32    ? afsl_warp_residency<1>(a, &perSm, nullptr, nullptr)
                  : Jp == 64  ? afsl_warp_residency<2>(a, &perSm, nullptr, nullptr)
                  : Jp == 128 ? afsl_warp_residency<4>(a, &perSm, nullptr, nullptr)
                  : Jp == 256 ? afsl_warp_residency<8>(a, &perSm, nullptr, nullptr)
                  : Jp == 512 ? afsl_warp_residency<16>(a, &perSm, nullptr, nullptr)
                              : afsl_warp_residency<32>(a, &perSm, nullptr, nullptr);
  return e == cudaSuccess ? sms * perSm : 0;
}

cudaError_t launch_afsl(const AllocArgs& a, int sms, cudaStream_t stream) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  const bool warp = a.max_total <= 0xFFFF;  // the one-warp kernel keeps a job's count in 16 bits
  if (warp && Jp <= 32) return launch_afsl_warp<1>(a, sms, stream);
  if (warp && Jp == 64) return launch_afsl_warp<2>(a, sms, stream);
  if (warp && Jp == 128) return launch_afsl_warp<4>(a, sms, stream);
  if (warp && Jp == 256) return launch_afsl_warp<8>(a, sms, stream);
  if (warp && Jp == 512) return launch_afsl_warp<16>(a, sms, stream);
  if (warp && Jp == 1024) return launch_afsl_warp<32>(a, sms, stream);
  const int CH = afsl_hot_columns(a, Jp);
  const size_t smem = afsl_fixed_smem(Jp) + (size_t)CH * 4 * (size_t)Jp;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(afsl_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  int perSm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, afsl_kernel, kThreads, smem);
  if (e != cudaSuccess) return e;
  if (perSm < 1) perSm = 1;
  int grid = sms * perSm;
  if (a.rounds < grid) grid = a.rounds;
  afsl_kernel<<<grid, kThreads, smem, stream>>>(a, Jp, CH);
  return cudaGetLastError();
}

}  // namespace voda
