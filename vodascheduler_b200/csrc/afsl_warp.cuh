// afsl_warp.cuh — AFS-L (pkg/algorithm/afsl.go:31-106) with ONE WARP per round, for rounds of up to
// 1024 jobs (sm_100a).  The rule being evaluated is stated at the top of afsl.cu; this file is
// about how one warp evaluates it with no CTA barrier and a quarter of the instructions.
//
// State.  Jobs are relabelled by their position after the stable SubmitTime sort ("id").  The
// live list is a sequence of positions p = 0 .. len-1; lane l of the warp owns the positions
// p = 32*k + l, k = 0 .. SLOTS-1, and keeps what a challenger contributes to one comparison in
// REGISTERS: id[k], L[k] = (S[r+1]-S[r])/S[r+1], Rr[k] = (S[r+1]-S[r])/S[r], a "count is zero" bit.
//
// One pass = every position from `pos` on against the champion entering `pos`: per row of 32
// positions one 16-bit shared-memory load, one integer and two float compares.  A lane remembers
// its own beating rows in a bit mask; one REDUX.MIN over (row, lane) gives the FIRST position
// whose job replaces the champion, everything before it is final, that job becomes the champion
// and the next pass starts right behind it: exactly the reference's left fold (a tree reduction
// would change results), about two passes per GPU handed out on the BASELINE workload.
//
// Lengths as ranks.  The rule compares len(champion, rc) with len(challenger, rc), both measured
// with the CHAMPION's worker count rc (reference bug at afsl.go:82, kept): two entries of the same
// column w = rc of the matrix len(j, w) = rem_j / S_j[w].  Only the order inside a column matters,
// so each column is stored as 16-bit ranks, rank(j) = #{i : len(i, w) < len(j, w)} — equal float32
// quotients get equal ranks, NaN (0/0) gets a flag that makes every compare false, as IEEE does.
// Column 0 holds the ranks of rem itself (the both-at-zero rule, :77-79).  Columns cost a quarter
// of a float32 table that also needs the quotient of every other column: 9 columns of 1024 jobs
// are 19 KB, so seven rounds are resident per SM.  A champion with more workers than the table has
// columns (rare) is compared by dividing on the fly.
//
// Speculation on the next state: when a job becomes the champion its next two table entries and
// its MaxNumProc are fetched from HBM at once; they arrive while the passes run, so the update of
// the winner (two IEEE divisions) does not wait for memory.
//
// The champion entering every take-over position is remembered; after the winner at position p
// got its GPU the next fold restarts at p with that champion (positions before p never involved
// the winner).  Jobs that reach MaxNumProc become dead slots and are squeezed out of the
// registers (through a 4 KB shared staging buffer) once 64 of them have accumulated.
#pragma once

#include "common.cuh"

namespace voda {
namespace afslw {

constexpr int kBig = 0x7fffffff;
constexpr int kNanRank = 0xFFFF;
constexpr int kCompactAt = 64;

#define AFSLW_FOR32(X)                                                                               \
  X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13) X(14) X(15) X(16) X(17) \
      X(18) X(19) X(20) X(21) X(22) X(23) X(24) X(25) X(26) X(27) X(28) X(29) X(30) X(31)

// order-preserving image of a float32 under IEEE '<' (-0 == +0); every NaN -> 0xFFFFFFFF
__device__ __forceinline__ unsigned rank_key(float x) {
  if (x != x) return 0xFFFFFFFFu;
  unsigned b = __float_as_uint(x);
  if (b == 0x80000000u) b = 0;
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// Warp-wide bitonic sort of P (a power of two >= 32) 32-bit keys in shared memory.
__device__ inline void warp_sort_u32(unsigned* key, int P, int lane) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (P >> 1); t += 32) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int ixj = i | j;
        const unsigned x = key[i], y = key[ixj];
        const unsigned lo = min(x, y), hi = max(x, y);
        const bool up = (i & k) == 0;
        key[i] = up ? lo : hi;
        key[ixj] = up ? hi : lo;
      }
      __syncwarp();
    }
  }
}

// Warp-wide bitonic sort of 32 * SLOTS keys held in registers, SLOTS per lane; element e of lane l is
// item l * SLOTS + e of the network.  Exchanges inside a lane are two min/max instructions on
// registers, exchanges between lanes one shuffle and one min-or-max; a lane whose block sorts
// descending in a phase works on the complemented keys instead (no per-exchange direction).
// About a quarter of the instructions of the same network on shared memory, and no barrier.
template <int SLOTS>
__device__ __forceinline__ void warp_sort_regs(unsigned (&v)[SLOTS], int lane) {
  constexpr int P = 32 * SLOTS;
#pragma unroll
  for (int k = 2; k <= P; k <<= 1) {
    const bool flip = k >= SLOTS && k < P && (lane & (k / SLOTS)) != 0;  // this lane's block is a descending one
    if (k >= SLOTS) {
#pragma unroll
      for (int e = 0; e < SLOTS; e++) v[e] = flip ? ~v[e] : v[e];
    }
#pragma unroll
    for (int j = k >> 1; j >= 1; j >>= 1) {
      if (j >= SLOTS) {
        const int partner = j / SLOTS;
        const bool lower = (lane & partner) == 0;
#pragma unroll
        for (int e = 0; e < SLOTS; e++) {
          const unsigned o = __shfl_xor_sync(0xffffffffu, v[e], partner);
          v[e] = lower ? min(v[e], o) : max(v[e], o);
        }
      } else {
#pragma unroll
        for (int e = 0; e < SLOTS; e++) {
          if ((e & j) == 0) {
            const bool asc = k >= SLOTS || (e & k) == 0;  // below SLOTS the direction is a bit of e
            const unsigned a = v[e], b = v[e | j];
            v[e] = asc ? min(a, b) : max(a, b);
            v[e | j] = asc ? max(a, b) : min(a, b);
          }
        }
      }
    }
    if (k >= SLOTS) {
#pragma unroll
      for (int e = 0; e < SLOTS; e++) v[e] = flip ? ~v[e] : v[e];
    }
  }
}

// One challenger against a champion that holds rc >= 1 workers (afsl.go:80-90), on predicates:
//   ge = rankA >= rankB   (lenA >= lenB: the challenger is the shorter job)
//   beat = ge ? !(L_c > Rr_b) : (L_b > Rr_c);   bits |= bit when beat
// (!(x > y) is "less, equal or unordered": a NaN ratio lets the challenger through, as in Go.)
__device__ __forceinline__ void row_rank(unsigned& bits, int rankA, int rankB, float Lc, float Rb, float Lb, float Rc,
                                         unsigned bit) {
  asm("{\n\t.reg .pred ge, t1, t2;\n\t"
      "setp.ge.s32 ge, %1, %2;\n\t"
      "setp.leu.and.f32 t1, %3, %4, ge;\n\t"
      "setp.gt.and.f32 t2, %5, %6, !ge;\n\t"
      "@t1 or.b32 %0, %0, %7;\n\t"
      "@t2 or.b32 %0, %0, %7;\n\t}"
      : "+r"(bits)
      : "r"(rankA), "r"(rankB), "f"(Lc), "f"(Rb), "f"(Lb), "f"(Rc), "r"(bit));
}

// One challenger against a champion at 0 workers: a challenger at 0 as well wins on
// rem_c >= rem_b (ranks of column 0, :77-79); any other meets lenA = lenB = +Inf, so the champion
// is the "long" job and is replaced iff !(L_c > Rr_b).
__device__ __forceinline__ void row_zero(unsigned& bits, int rankA, int rankB, unsigned zeroBits, float Lc, float Rb,
                                         unsigned bit) {
  asm("{\n\t.reg .pred z, t1, t2;\n\t.reg .b32 m;\n\t"
      "and.b32 m, %3, %6;\n\t"
      "setp.ne.b32 z, m, 0;\n\t"
      "setp.ge.and.s32 t1, %1, %2, z;\n\t"
      "setp.leu.and.f32 t2, %4, %5, !z;\n\t"
      "@t1 or.b32 %0, %0, %6;\n\t"
      "@t2 or.b32 %0, %0, %6;\n\t}"
      : "+r"(bits)
      : "r"(rankA), "r"(rankB), "r"(zeroBits), "f"(Lc), "f"(Rb), "r"(bit));
}

struct Smem {
  unsigned short* rankT;  // [NC][pitch], pitch = Jp + 32; entry Jp of a column is the dead slots' flag
  unsigned short* r;      // [Jp] workers per id (a job never holds more than min(total, MaxNumProc) <= 65535)
  unsigned* scr;          // [Jp] rank-sort keys during set-up; afterwards its first half is r[], its second
                          // half the champion cache, which also stages the compaction (half the positions at a time)
  unsigned short* perm;   // [Jp] id -> job index within the round
  long long* red;         // [2]
  // sort overlay on rankT
  int64_t* khi;
  int32_t* klo;
  int* perm32;
  unsigned long long* ukey;
};

__host__ __device__ inline size_t table_bytes(int Jp, int NC) {
  size_t t = (size_t)2 * (size_t)NC * (size_t)(Jp + 32);
  const size_t overlay = (size_t)16 * (size_t)Jp;
  t = t < overlay ? overlay : t;
  return (t + 15) & ~(size_t)15;
}
__host__ __device__ inline size_t smem_bytes(int Jp, int NC) {
  return table_bytes(Jp, NC) + (size_t)Jp * (4 + 2) + 16 + 16;
}

template <int SLOTS, bool POOL>
__global__ void __launch_bounds__(32) afsl_warp_kernel(AllocArgs a, int NC) {
  constexpr int Jp = 32 * SLOTS;
  constexpr int pitch = Jp + 32;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem sm;
  {
    unsigned char* p = smem_raw;
    sm.red = reinterpret_cast<long long*>(p); p += 16;
    sm.rankT = reinterpret_cast<unsigned short*>(p);
    sm.khi = reinterpret_cast<int64_t*>(p);
    sm.ukey = reinterpret_cast<unsigned long long*>(p);
    sm.klo = reinterpret_cast<int32_t*>(p + 8 * (size_t)Jp);
    sm.perm32 = reinterpret_cast<int*>(p + 12 * (size_t)Jp);
    p += table_bytes(Jp, NC);
    sm.scr = reinterpret_cast<unsigned*>(p);
    sm.r = reinterpret_cast<unsigned short*>(p); p += 4 * (size_t)Jp;
    sm.perm = reinterpret_cast<unsigned short*>(p);
  }
  const int lane = threadIdx.x;
  const int W = a.stride;
  const int CH = NC - 1;  // len() columns 1..CH are tabulated
  const unsigned full = 0xffffffffu;
  const unsigned ltmask = (1u << lane) - 1u;
  const float NEG_INF = VODA_NEG_INF;

  for (int rd = blockIdx.x; rd < a.rounds; rd += gridDim.x) {
    const int o = a.job_off[rd] - a.job_base;
    const int n = a.job_off[rd + 1] - a.job_off[rd];
    const int total = a.total_gpu[rd];
    if (n > a.max_jobs || n > Jp) {
      for (int i = lane; i < n; i += 32) a.out_gpus[o + i] = 0;
      if (lane == 0) a.out_status[rd] = VODA_ST_BOUND;
      continue;
    }
    __syncwarp();
    // Info.Speedup[fmt.Sprint(g)], map miss => 0.  Offsets inside a round fit 32 bits (at most
    // 1024 jobs x W entries): one IMAD per lookup instead of 64-bit address arithmetic.
    const float* __restrict__ Sb = a.speedup + (size_t)o * (size_t)W;
    const int32_t* __restrict__ mxb = a.max_proc + o;
    const float* __restrict__ remb = a.remaining_sec + o;
    // POOL: the table is a pool of curves the jobs point into (voda_alloc_batch.speedup_off); a separate
    // instantiation, because even a uniform branch in this lookup costs the dense path 10 %
    auto S = [&](int job, int g) -> float {
      if ((unsigned)g >= (unsigned)W) return 0.0f;
      if constexpr (POOL) return a.speedup[(size_t)a.sp_off[o + job] + (size_t)g];
      else return Sb[(unsigned)(job * W + g)];
    };
    // ---- stable sort by SubmitTime (:42-44); ids are sorted positions from here on ---------
    {
      bool fits = true;
      for (int i = lane; i < Jp; i += 32) {
        unsigned long long u = ~0ull;
        if (i < n) fits = pack_time_key(a.submit_sec[o + i], a.submit_nsec ? a.submit_nsec[o + i] : 0, &u) && fits;
        sm.ukey[i] = u;
      }
      if (__all_sync(full, fits)) {
        __syncthreads();
        block_sort_kv(sm.ukey, sm.perm32, Jp);
      } else {
        __syncthreads();
        for (int i = lane; i < n; i += 32) {
          sm.khi[i] = a.submit_sec[o + i];
          sm.klo[i] = a.submit_nsec ? a.submit_nsec[o + i] : 0;
        }
        __syncthreads();
        block_argsort(sm.khi, sm.klo, sm.perm32, n, next_pow2(n > 0 ? n : 1));
      }
      unsigned short tmp[SLOTS];
#pragma unroll
      for (int k = 0; k < SLOTS; k++) {
        const int t = 32 * k + lane;
        tmp[k] = t < n ? (unsigned short)sm.perm32[t] : (unsigned short)0;
      }
      __syncwarp();  // the overlay is dead from here on
#pragma unroll
      for (int k = 0; k < SLOTS; k++) sm.perm[32 * k + lane] = tmp[k];
    }
    // ---- rank columns: 0 = rem, w = rem / S[w] (afsl.go:94-100), w up to the largest count a
    //      champion can hold in this round (below the largest MaxNumProc).  Built before the per-slot
    //      state so that the sort has the registers to itself.
    int maxmax = 0;
    {
      float remv[SLOTS];
#pragma unroll
      for (int k = 0; k < SLOTS; k++) {
        const int t = 32 * k + lane;
        remv[k] = 0.0f;
        if (t < n) {
          const int job = sm.perm[t];
          remv[k] = remb[job];
          maxmax = max(maxmax, mxb[job]);
        }
      }
      maxmax = __reduce_max_sync(full, maxmax);
      const int cols = min(NC, max(1, maxmax));
      for (int c = 0; c < cols; c++) {
        unsigned key[SLOTS], srt[SLOTS];
#pragma unroll
        for (int k = 0; k < SLOTS; k++) {
          const int t = 32 * k + lane;
          float v = remv[k];
          if (c > 0 && t < n) v = v / S(sm.perm[t], c);
          key[k] = t < n ? rank_key(v) : 0xFFFFFFFFu;
          srt[k] = key[k];
        }
        warp_sort_regs<SLOTS>(srt, lane);
#pragma unroll
        for (int k = 0; k < SLOTS; k++) sm.scr[lane * SLOTS + k] = srt[k];
        __syncwarp();
        // rank = number of keys below mine (lower bound in the sorted keys)
        int lb[SLOTS];
#pragma unroll
        for (int k = 0; k < SLOTS; k++) lb[k] = 0;
#pragma unroll 1
        for (int step = Jp >> 1; step >= 1; step >>= 1) {
#pragma unroll
          for (int k = 0; k < SLOTS; k++)
            if (sm.scr[lb[k] + step - 1] < key[k]) lb[k] += step;
        }
        unsigned short* col = sm.rankT + (size_t)c * pitch;
#pragma unroll
        for (int k = 0; k < SLOTS; k++)
          col[32 * k + lane] = key[k] == 0xFFFFFFFFu ? (unsigned short)kNanRank : (unsigned short)lb[k];
        if (lane == 0) col[Jp] = (unsigned short)kNanRank;
        __syncwarp();
      }
    }
    // ---- per-slot registers ---------------------------------------------------------------
    int id[SLOTS];
    float L[SLOTS], Rr[SLOTS];
    unsigned zeroBits = 0xffffffffu;  // bit k: slot k holds a job with 0 workers (dead slots keep the bit)
    unsigned aliveBits = 0;           // bit k: slot k holds a live job
#pragma unroll
    for (int k = 0; k < SLOTS; k++) {
      const int t = 32 * k + lane;
      if (t < n) {
        const int job = sm.perm[t];
        id[k] = t;
        const float s0 = S(job, 0), s1 = S(job, 1);
        const float d = s1 - s0;
        L[k] = d / s1;
        Rr[k] = d / s0;
        aliveBits |= 1u << k;
      } else {
        id[k] = Jp;
        L[k] = NEG_INF;
        Rr[k] = 0.0f;
      }
    }
    // r[] and the champion cache take over the sort scratch
    unsigned short* champIn = sm.r + Jp;  // position of the champion entering a take-over position
    unsigned* stage = reinterpret_cast<unsigned*>(champIn);  // Jp/2 words: compaction staging
#pragma unroll
    for (int k = 0; k < SLOTS; k++) sm.r[32 * k + lane] = 0;
    if (min(total, max(maxmax, 1)) > 0xFFFF) {  // a count could outgrow 16 bits: not this kernel's shape
      for (int i = lane; i < n; i += 32) a.out_gpus[o + i] = 0;
      if (lane == 0) a.out_status[rd] = VODA_ST_BOUND;
      continue;
    }
    __syncwarp();

    // ---- uniform state (same value in every lane) ------------------------------------------------
    int len = n;                 // positions in use (live + dead)
    int sEnd = (len + 31) >> 5;  // rows in use
    int dead = 0, nlive = n, head = 0;
    long long free = total;
    int np = -1;   // position whose job becomes the champion next (-1: seed from the first live job)
    int pos = 0;   // first position the next pass tests

    while (free > 0 && nlive > 0) {  // :47; one iteration = one pass
      if (np < 0) {  // seed: the first live job (:73)
        const int k0 = aliveBits ? (((__ffs(aliveBits) - 1) << 5) | lane) : kBig;
        head = __reduce_min_sync(full, k0);
        np = head;
        pos = head + 1;
      }
      // ---- the job at position np becomes the champion (the only place its state is fetched) ------
      const int cp = np;
      int c;
      float Lc, Rc;
      {
        const int sx = cp >> 5, lx = cp & 31;
        int xi = 0;
        float xl = 0.f, xr = 0.f;
        switch (sx) {
#define X(k) case k: if constexpr (k < SLOTS) { xi = id[k]; xl = L[k]; xr = Rr[k]; } break;
          AFSLW_FOR32(X)
#undef X
        }
        c = __shfl_sync(full, xi, lx);
        Lc = __shfl_sync(full, xl, lx);
        Rc = __shfl_sync(full, xr, lx);
      }
      const int rc = sm.r[c];
      const int jobc = sm.perm[c];
      // what the update needs if this champion wins: in flight while the pass runs
      const int mxc = mxb[jobc];
      const float sA = S(jobc, rc + 1), sB = S(jobc, rc + 2);

      // ---- one pass: every position >= pos against this champion -----------------------------------
      unsigned bits = 0;
      if (pos < len) {
        const int s0 = pos >> 5;
        if (__builtin_expect(rc <= CH, 1)) {
          const unsigned short* col = sm.rankT + rc * pitch;
          const int rk = col[c];
          const int rankA = rk == kNanRank ? -1 : rk;  // NaN: every '>=' is false
          if (rc == 0) {
            // champion at 0 workers: a challenger at 0 wins on rem_c >= rem_b (:77-79); against any
            // other lenA = lenB = +Inf, the champion is the "long" one and survives iff L_c > Rr_b
#define ROW(k) row_zero(bits, rankA, (int)col[id[k]], zeroBits, Lc, Rr[k], 1u << (k));
#pragma unroll
            for (int g = 0; g < SLOTS; g += 4) {
              if (g + 3 >= s0 && g < sEnd) {
                ROW(g) if constexpr (SLOTS > 1) { ROW(g + 1) } if constexpr (SLOTS > 2) { ROW(g + 2) ROW(g + 3) }
              }
            }
#undef ROW
          } else {
            // both lengths at the champion's count rc (:81-82): ranks of column rc
#define ROW(k) row_rank(bits, rankA, (int)col[id[k]], Lc, Rr[k], L[k], Rc, 1u << (k));
#pragma unroll
            for (int g = 0; g < SLOTS; g += 4) {
              if (g + 3 >= s0 && g < sEnd) {
                ROW(g) if constexpr (SLOTS > 1) { ROW(g + 1) } if constexpr (SLOTS > 2) { ROW(g + 2) ROW(g + 3) }
              }
            }
#undef ROW
          }
        } else {
          // more workers than the table has columns: divide on the fly (w >= W is a map miss: rem / 0)
          const float lenAf = remb[jobc] / S(jobc, rc);
#pragma unroll
          for (int k = 0; k < SLOTS; k++) {
            if (k >= s0 && k < sEnd && id[k] != Jp) {
              const int jb = sm.perm[id[k]];
              const float lenB = remb[jb] / S(jb, rc);
              const bool b = (lenAf >= lenB) ? !(Lc > Rr[k]) : (L[k] > Rc);
              if (b) bits |= 1u << k;
            }
          }
        }
        // rows are evaluated whole; drop the positions before pos (rows past the end hold dead slots)
        bits &= (lane < (pos & 31)) ? ((0xffffffffu << s0) & ~(1u << s0)) : (0xffffffffu << s0);
      }
      const int key = bits ? (((__ffs(bits) - 1) << 5) | lane) : kBig;
      const int first = __reduce_min_sync(full, key);
      if (first != kBig) {  // that job takes over; everything before it is final
        if (lane == 0) champIn[first] = (unsigned short)cp;
        np = first;
        pos = first + 1;
        continue;
      }
      // ---- the fold is over: give the champion one GPU (:49-57) ----------------------------------
      const int nr = rc + 1;
      free -= 1;
      const bool removed = nr >= mxc;
      const float d = sB - sA;
      const float nL = d / sB, nR = d / sA;
      const int sx = cp >> 5, lx = cp & 31;
      __syncwarp();
      const int prevPos = cp > head ? (int)champIn[cp] : -1;
      if (lane == 0) sm.r[c] = (unsigned short)nr;
      if (lane == lx) {
        if (removed) {
          a.out_gpus[o + jobc] = nr;
          aliveBits &= ~(1u << sx);
          zeroBits |= 1u << sx;
        } else {
          zeroBits &= ~(1u << sx);
        }
        switch (sx) {
#define X(k) case k: if constexpr (k < SLOTS) { if (removed) { id[k] = Jp; L[k] = NEG_INF; Rr[k] = 0.0f; } else { L[k] = nL; Rr[k] = nR; } } break;
          AFSLW_FOR32(X)
#undef X
        }
      }
      if (removed) {
        nlive--;
        dead++;
      }
      __syncwarp();
      // the next fold restarts at the winner's position with the champion that entered it: positions
      // before it never involved the winner.  If the winner was the seed, start over.
      np = prevPos;
      pos = removed ? cp + 1 : cp;
      // ---- squeeze the dead slots out (uniform condition) ---------------------------------------
      if (__builtin_expect(dead >= kCompactAt || (dead > 0 && dead * 4 >= len), 0)) {
        unsigned nzero = 0;
        int base = 0;
        constexpr int kHalf = Jp / 2;  // the staging buffer holds half of the positions
#pragma unroll
        for (int pass = 0; pass < 3; pass++) {
          unsigned nz = 0;
#pragma unroll
          for (int h = 0; h < 2; h++) {
            base = 0;
#pragma unroll
            for (int k = 0; k < SLOTS; k++) {
              const bool alive = (aliveBits >> k) & 1u;
              const unsigned m = __ballot_sync(full, alive);
              const int dst = base + __popc(m & ltmask);
              if (alive && dst / kHalf == h) {
                unsigned v;
                if (pass == 0) v = (unsigned)id[k] | (((zeroBits >> k) & 1u) << 16);
                else if (pass == 1) v = __float_as_uint(L[k]);
                else v = __float_as_uint(Rr[k]);
                stage[dst - h * kHalf] = v;
              }
              base += __popc(m);
            }
            __syncwarp();
            // in place: a live element never moves to a later position, so the second half's sources
            // (positions >= Jp/2) are untouched by the first half's writes; zeroBits is rebuilt aside
#pragma unroll
            for (int k = 0; k < SLOTS; k++) {
              const int p = 32 * k + lane;
              if (p / kHalf != h) continue;
              const bool in = p < base;
              const unsigned v = in ? stage[p - h * kHalf] : 0u;
              if (pass == 0) {
                id[k] = in ? (int)(v & 0xFFFFu) : Jp;
                nz |= (in ? ((v >> 16) & 1u) : 1u) << k;
              } else if (pass == 1) {
                L[k] = in ? __uint_as_float(v) : NEG_INF;
              } else {
                Rr[k] = in ? __uint_as_float(v) : 0.0f;
              }
            }
            __syncwarp();
          }
          if (pass == 0) nzero = nz;
        }
        zeroBits = nzero;
        aliveBits = 0;
#pragma unroll
        for (int k = 0; k < SLOTS; k++) aliveBits |= (32 * k + lane < base ? 1u : 0u) << k;
        len = base;
        sEnd = (len + 31) >> 5;
        dead = 0;
        np = -1;  // positions moved: the champion cache is void
      }
    }
    // ---- results (removed jobs were written when they left) --------------------------------------
#pragma unroll
    for (int k = 0; k < SLOTS; k++)
      if (id[k] != Jp) a.out_gpus[o + sm.perm[id[k]]] = sm.r[id[k]];
    __syncthreads();
    const int v = block_validate(a, o, n, total, a.out_gpus + o, nullptr, sm.red);  // deferred :39
    if (lane == 0) a.out_status[rd] = v;
  }
}

#undef AFSLW_FOR32

}  // namespace afslw
}  // namespace voda
