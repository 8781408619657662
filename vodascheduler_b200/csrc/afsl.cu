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
//     ballot finds the FIRST challenger that replaces it, that one becomes the champion and
//     only the lanes after it re-evaluate (speculative first-beat scan);
//   * the champion entering every position is cached; after the winner at position p gets its
//     GPU the next fold restarts at p with the cached champion (positions before p never
//     involve the winner, so their outcome cannot change);
//   * the divisions are hoisted: len(x, w) = rem_x / S_x[w] is tabulated once per round for
//     every (job, w) in shared memory (same float32 quotient the reference computes each
//     time), and the two evaluate() ratios of a job are recomputed only when its count changes.
// Removal at max keeps order (lazy deletion + periodic compaction of the live list).
#include "common.cuh"
#include "kernels.h"

namespace voda {
namespace {

constexpr int kThreads = 128;

struct AfslSmem {
  int64_t* khi;
  long long* red;   // [4]
  int32_t* klo;
  int* order;       // live list (job ids in SubmitTime order), dead entries = -1
  int* champIn;     // champion entering each position of `order` ...
  int* champInPos;  // ... and the position that champion sits at
  int32_t* r;
  int32_t* mx;
  float* rem;
  float* L;         // (S[r+1]-S[r]) / S[r+1]   "as the longer job"
  float* Rr;        // (S[r+1]-S[r]) / S[r]     "as the shorter job"
  float* lenT;      // [n * W] rem / S[w]  (w = 0: +Inf), when staged
};

__global__ void __launch_bounds__(kThreads) afsl_kernel(AllocArgs a, int Jp, int tableInSmem) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  AfslSmem sm;
  {
    unsigned char* p = smem_raw;
    sm.khi = reinterpret_cast<int64_t*>(p); p += 8 * (size_t)Jp;
    sm.red = reinterpret_cast<long long*>(p); p += 8 * 4;
    sm.klo = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.order = reinterpret_cast<int*>(p); p += 4 * (size_t)Jp;
    sm.champIn = reinterpret_cast<int*>(p); p += 4 * (size_t)Jp;
    sm.champInPos = reinterpret_cast<int*>(p); p += 4 * (size_t)Jp;
    sm.r = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.mx = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Jp;
    sm.rem = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.L = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.Rr = reinterpret_cast<float*>(p); p += 4 * (size_t)Jp;
    sm.lenT = reinterpret_cast<float*>(p);
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
    auto S = [&](int i, int g) -> float {  // Info.Speedup[fmt.Sprint(g)], map miss => 0
      if (g < 0 || g >= W) return 0.0f;
      return a.speedup[(size_t)(o + i) * (size_t)W + (size_t)g];
    };
    for (int i = tid; i < n; i += kThreads) {
      sm.khi[i] = a.submit_sec[o + i];
      sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
      sm.r[i] = 0;
      sm.mx[i] = a.max_proc[o + i];
      sm.rem[i] = a.remaining_sec[o + i];
      const float s0 = S(i, 0), s1 = S(i, 1);
      const float d = s1 - s0;
      sm.L[i] = d / s1;
      sm.Rr[i] = d / s0;
    }
    if (tableInSmem) {
      const float* src = a.speedup + (size_t)o * (size_t)W;
      for (int x = tid; x < n * W; x += kThreads) {
        const int i = x / W, w = x - i * W;
        sm.lenT[x] = w == 0 ? INF : a.remaining_sec[o + i] / src[x];  // afsl.go:94-100
      }
    }
    __syncthreads();
    block_argsort(sm.khi, sm.klo, sm.order, n, next_pow2(n > 0 ? n : 1));  // :42-44
    __syncthreads();

    if (warp == 0) {
      // len(x, w) for w given; w >= W is a map miss: rem / 0
      auto LEN = [&](int x, int w) -> float {
        if (w == 0) return INF;
        if (w < 0 || w >= W) return sm.rem[x] / 0.0f;
        return tableInSmem ? sm.lenT[x * W + w] : sm.rem[x] / S(x, w);
      };
      int len = n;        // entries in order[] (live + dead)
      int dead = 0;
      int head = 0;       // first live position
      int startPos = 0;   // fold restart: position of the next challenger
      int champ = -1;     // champion entering startPos (-1: seed from head) ...
      int champPos = -1;  // ... and where it sits
      long long free = total;
      int nlive = n;
      while (free > 0 && nlive > 0) {  // :47
        // ---- topPriority: fold from startPos ------------------------------------------
        if (champ < 0) {
          while (head < len && sm.order[head] < 0) head++;
          champ = sm.order[head];
          champPos = head;
          startPos = head + 1;
        }
        int c = champ, cp = champPos;
        for (int base = startPos & ~31; base < len; base += 32) {
          const int p = base + lane;
          const int jb = (p < len && p >= startPos) ? sm.order[p] : -1;
          bool pend = jb >= 0;
          // challenger-side values (independent of the champion)
          int rj = 0;
          float remj = 0.f, Lj = 0.f, Rj = 0.f;
          if (pend) {
            rj = sm.r[jb];
            remj = sm.rem[jb];
            Lj = sm.L[jb];
            Rj = sm.Rr[jb];
          }
          while (true) {
            const int rc = sm.r[c];
            const float remc = sm.rem[c], Lc = sm.L[c], Rc = sm.Rr[c];
            const float lenA = LEN(c, rc);
            bool beat = false;
            if (pend) {
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
            if (!m) {
              if (pend) {
                sm.champIn[p] = c;
                sm.champInPos[p] = cp;
              }
              break;
            }
            const int f = __ffs(m) - 1;
            if (pend && lane <= f) {
              sm.champIn[p] = c;
              sm.champInPos[p] = cp;
              pend = false;
            }
            c = __shfl_sync(full, jb, f);
            cp = base + f;
          }
        }
        // ---- give it one GPU (:49-57) ----------------------------------------------------
        const int w = c, wpos = cp;
        __syncwarp();
        const int nr = sm.r[w] + 1;
        free -= 1;
        const bool removed = nr >= sm.mx[w];
        __syncwarp();
        if (lane == 0) {
          sm.r[w] = nr;
          const float s0 = S(w, nr), s1 = S(w, nr + 1);
          const float d = s1 - s0;
          sm.L[w] = d / s1;
          sm.Rr[w] = d / s0;
          if (removed) sm.order[wpos] = -1;
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
        __syncwarp();
        // ---- compaction of the live list ---------------------------------------------------
        if (dead >= 64 || (dead > 0 && dead * 4 >= len)) {
          int outp = 0;
          for (int base = 0; base < len; base += 32) {
            const int p = base + lane;
            const int v = p < len ? sm.order[p] : -1;
            const unsigned keep = __ballot_sync(full, v >= 0);
            const int dst = outp + __popc(keep & ((1u << lane) - 1u));
            __syncwarp();
            if (v >= 0) sm.order[dst] = v;
            outp += __popc(keep);
            __syncwarp();
          }
          len = outp;
          dead = 0;
          head = 0;
          champ = -1;  // positions moved: cached champions are void
        }
      }
    }
    __syncthreads();
    for (int i = tid; i < n; i += kThreads) a.out_gpus[o + i] = sm.r[i];
    __syncthreads();
    const int v = block_validate(a, o, n, total, a.out_gpus + o, nullptr, sm.red);  // deferred :39
    if (tid == 0) a.out_status[rd] = v;
  }
}

size_t afsl_fixed_smem(int Jp) { return (size_t)Jp * (8 + 4 * 10) + 8 * 4 + 32; }

}  // namespace

size_t afsl_smem_bytes(const AllocArgs& a) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  size_t fixed = afsl_fixed_smem(Jp);
  size_t table = (size_t)a.max_jobs * (size_t)a.stride * 4;
  return fixed + (fixed + table <= 220 * 1024 ? table : 0);
}

cudaError_t launch_afsl(const AllocArgs& a, int sms, cudaStream_t stream) {
  const int Jp = next_pow2(a.max_jobs > 0 ? a.max_jobs : 1);
  const size_t fixed = afsl_fixed_smem(Jp);
  const size_t table = (size_t)a.max_jobs * (size_t)a.stride * 4;
  const bool in = fixed + table <= 220 * 1024;
  const size_t smem = fixed + (in ? table : 0);
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
  afsl_kernel<<<grid, kThreads, smem, stream>>>(a, Jp, in ? 1 : 0);
  return cudaGetLastError();
}

}  // namespace voda
