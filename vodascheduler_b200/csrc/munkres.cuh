// munkres.cuh — warp-level six-step Munkres on a cost matrix held in shared memory (sm_100a).
//
// Restates munkres.ComputeMunkresMax of github.com/heyfey/munkres@d4f16a25250d (not vendored
// in the reference; call site pkg/placement/placement_manager.go:505-512) in the step order of
// SURVEY.md Appendix C, preserving every "first in row-major order" choice, so the starred
// cells — the tie-breaking — match the CPU restatement bit for bit.
//
// ONE WARP solves one n x n matrix (n <= 128):
//   * the matrix after "maximise" + step 1 is C[i][j] = rowmax_i - A[i][j] >= 0 (the constant
//     of the max->min transform cancels in the row-minimum subtraction); it stays immutable
//     in shared memory in the narrowest cell type that holds it;
//   * step 6 (add m to covered rows, subtract m from uncovered columns) only moves row / column
//     potentials: value(i,j) = C[i][j] + ra[i] - cs[j];
//   * each lane owns 4 consecutive rows and keeps their 128-bit ZERO MASKS in registers, so
//     "first uncovered zero in row-major order" (step 4, ~n^2/16 times per matrix) is one
//     ballot + two find-first-set instead of an n^2 scan; step 6 patches the masks (covered
//     rows lose their zeros in covered columns; uncovered x uncovered cells equal to m gain one);
//   * stars / primes are starCol[row], starRow[col], primeCol[row] in shared memory, and so
//     are the column covers and a column-major copy of the zero masks (which rows have a zero
//     in column j), so that a step-4 iteration is ~40 instructions.
#pragma once

#include "common.cuh"

namespace voda {

template <typename CellT>
struct MunkresTypes {
  using Pot = int32_t;  // potentials are bounded by n * max(C) < 2^31 for 8/16-bit cells
};
template <>
struct MunkresTypes<int32_t> {
  using Pot = int64_t;
};
template <>
struct MunkresTypes<int64_t> {
  using Pot = int64_t;
};

struct MunkresSmem {
  short* starCol;   // [128] column of the star in row i, -1 if none
  short* starRow;   // [128] row of the star in column j, -1 if none
  short* primeCol;  // [128] column of the prime in row i (valid for rows covered in this phase)
  void* ra;         // [128] row potentials (Pot)
  uint32_t* zt;     // [128][4] column-major zero masks: bit i of zt[j] = (value(i, j) == 0);
                    //          the caller fills it together with the row-major masks Z
  uint32_t* ccs;    // [4] column covers
};

__device__ __forceinline__ void mask_set(uint32_t (&m)[4], int idx) {
  const uint32_t bit = 1u << (idx & 31);
  const int w = idx >> 5;
  m[0] |= w == 0 ? bit : 0u;
  m[1] |= w == 1 ? bit : 0u;
  m[2] |= w == 2 ? bit : 0u;
  m[3] |= w == 3 ? bit : 0u;
}
__device__ __forceinline__ void mask_clear(uint32_t (&m)[4], int idx) {
  const uint32_t bit = 1u << (idx & 31);
  const int w = idx >> 5;
  m[0] &= w == 0 ? ~bit : ~0u;
  m[1] &= w == 1 ? ~bit : ~0u;
  m[2] &= w == 2 ? ~bit : ~0u;
  m[3] &= w == 3 ? ~bit : ~0u;
}
__device__ __forceinline__ bool mask_test(const uint32_t (&m)[4], int idx) {
  const int w = idx >> 5;
  const uint32_t word = w == 0 ? m[0] : w == 1 ? m[1] : w == 2 ? m[2] : m[3];
  return (word >> (idx & 31)) & 1u;
}
// the 4 cover bits of the rows a lane owns (rows 4*lane .. 4*lane+3)
__device__ __forceinline__ uint32_t own_rows(const uint32_t (&rowCov)[4], int lane) {
  const int w = lane >> 3;
  const uint32_t word = w == 0 ? rowCov[0] : w == 1 ? rowCov[1] : w == 2 ? rowCov[2] : rowCov[3];
  return (word >> (4 * (lane & 7))) & 0xFu;
}

template <typename T>
__device__ __forceinline__ T warp_min(T v) {
  for (int d = 16; d > 0; d >>= 1) {
    T o = __shfl_xor_sync(0xffffffffu, v, d);
    v = o < v ? o : v;
  }
  return v;
}

// Solve.  C: n*n cells (row stride n) already holding rowmax_i - A[i][j]; Z: the zero masks of
// the calling lane's rows (bit j of Z[q] = C[4*lane+q][j] == 0).  On return sm.starCol[i] is
// the column assigned to row i.  Must be called by one full warp.
template <typename CellT>
__device__ void munkres_solve(const CellT* __restrict__ C, int n, uint32_t (&Z)[4][4], const MunkresSmem& sm) {
  using Pot = typename MunkresTypes<CellT>::Pot;
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  Pot* ra = static_cast<Pot*>(sm.ra);
  Pot cs[4] = {0, 0, 0, 0};  // potentials of columns lane + 32 t
  for (int i = lane; i < 128; i += 32) {
    ra[i] = 0;
    sm.starCol[i] = -1;
    sm.starRow[i] = -1;
    sm.primeCol[i] = -1;
  }
  __syncwarp();

  // ---- step 2: star the first zero of each row whose column holds no star yet, rows in order
  uint32_t starred[4] = {0, 0, 0, 0};  // columns holding a star (uniform)
  int nstar = 0;
  for (int l = 0; l < 32 && 4 * l < n; l++) {
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int i = 4 * l + q;
      if (i >= n) break;
      int col = -1;
#pragma unroll
      for (int t = 3; t >= 0; t--) {
        const uint32_t cand = Z[q][t] & ~starred[t];
        if (cand) col = 32 * t + __ffs(cand) - 1;
      }
      col = __shfl_sync(full, col, l);
      if (col >= 0) {
        mask_set(starred, col);
        nstar++;
        if (lane == 0) {
          sm.starCol[i] = (short)col;
          sm.starRow[col] = (short)i;
        }
      }
    }
  }
  __syncwarp();

  // Covers.  Columns: 128 bits in shared memory (sm.ccs), copied to registers when a whole
  // mask is needed.  Rows: each lane keeps the state of the 4 rows it owns.
  //
  // k[q] = (q << 8) | first uncovered zero column of owned row q, or kBig when the row is
  // covered or has none.  Inside one step-4 phase columns only get UNcovered and rows only get
  // covered, so uncovering column sc turns k[q] into min(k[q], (q << 8) | sc) for the live rows
  // that have a zero there — read from the column-major zero masks sm.zt — and the search
  // "first uncovered zero in row-major order" never rescans anything: it is the minimum key
  // of the lowest lane that has one.
  constexpr int kBig = 0x7fffffff;
  uint32_t colCov[4] = {starred[0], starred[1], starred[2], starred[3]};  // step 3
  bool ucolValid[4];
#pragma unroll
  for (int t = 0; t < 4; t++) ucolValid[t] = lane + 32 * t < n;
  int k[4];
  uint32_t live = 0xFu;  // owned rows not covered
  auto reset_phase = [&]() {  // covers = {starred columns}, no row; keys from the zero masks
    if (lane < 4) sm.ccs[lane] = lane == 0 ? colCov[0] : lane == 1 ? colCov[1] : lane == 2 ? colCov[2] : colCov[3];
    live = 0xFu;
#pragma unroll
    for (int q = 0; q < 4; q++) {
      k[q] = kBig;
#pragma unroll
      for (int t = 3; t >= 0; t--) {
        const uint32_t cand = Z[q][t] & ~colCov[t];
        if (cand) k[q] = (q << 8) | (32 * t + __ffs(cand) - 1);
      }
    }
    __syncwarp();
  };
  reset_phase();
  const int ztWord = lane >> 3, ztShift = 4 * (lane & 7);

  while (nstar < n) {
    // ---- step 4: first uncovered zero in row-major order
    const int key = min(min(k[0], k[1]), min(k[2], k[3]));
    const unsigned b = __ballot_sync(full, key != kBig);
    if (b) {
      const int ls = __ffs(b) - 1;
      const int packed = __shfl_sync(full, key, ls);
      const int row = 4 * ls + (packed >> 8), col = packed & 255;
      const int sc = sm.starCol[row];
      if (sc >= 0) {  // prime it, cover its row, uncover the star's column
        if (lane == 0) {
          sm.primeCol[row] = (short)col;
          sm.ccs[sc >> 5] &= ~(1u << (sc & 31));
        }
        if (lane == ls) {  // the chosen key is this lane's k[q] of the covered row
          live &= ~(1u << (packed >> 8));
#pragma unroll
          for (int q = 0; q < 4; q++) k[q] = k[q] == packed ? kBig : k[q];
        }
        const uint32_t hits = (sm.zt[sc * 4 + ztWord] >> ztShift) & live;
#pragma unroll
        for (int q = 0; q < 4; q++)
          if ((hits >> q) & 1u) k[q] = min(k[q], (q << 8) | sc);
        continue;
      }
      // ---- step 5: augment along the alternating path that starts at the prime (row, col)
      int last = col;
      if (lane == 0) {
        int r = row, c = col;
        while (true) {
          const int rn = sm.starRow[c];
          sm.starRow[c] = (short)r;
          sm.starCol[r] = (short)c;
          if (rn < 0) break;
          r = rn;
          c = sm.primeCol[r];
        }
        last = c;
      }
      last = __shfl_sync(full, last, 0);
      __syncwarp();
      mask_set(starred, last);  // the path ends in the one column that gains a star
      nstar++;
#pragma unroll
      for (int t = 0; t < 4; t++) colCov[t] = starred[t];  // step 3; covers cleared, primes stale
      reset_phase();
      continue;
    }
    // ---- step 6 needs whole masks: columns from shared memory, rows from the per-lane bits
    __syncwarp();
#pragma unroll
    for (int t = 0; t < 4; t++) colCov[t] = sm.ccs[t];
    const uint32_t ownCov = ~live & 0xFu;
    // rebuild the uniform row-cover mask from the per-lane bits (rows 4*lane .. 4*lane+3)
    uint32_t rowCov[4];
    {
      const uint32_t nib = ownCov << (4 * (lane & 7));
#pragma unroll
      for (int w = 0; w < 4; w++) rowCov[w] = __reduce_or_sync(full, (lane >> 3) == w ? nib : 0u);
    }
    const uint32_t own = ownCov;
    // ---- step 6: m = smallest uncovered value; covered rows += m, uncovered columns -= m
    bool ucol[4];
#pragma unroll
    for (int t = 0; t < 4; t++) ucol[t] = ucolValid[t] && !((colCov[t] >> lane) & 1u);
    Pot mn = sizeof(Pot) == 8 ? (Pot)0x7fffffffffffffffLL : (Pot)0x7fffffff;
    for (int i = 0; i < n; i++) {
      if (mask_test(rowCov, i)) continue;
      const Pot rai = ra[i];
      const CellT* row = C + (size_t)i * n;
#pragma unroll
      for (int t = 0; t < 4; t++) {
        if (ucol[t]) {
          const Pot v = (Pot)row[lane + 32 * t] + rai - cs[t];
          mn = v < mn ? v : mn;
        }
      }
    }
    mn = warp_min(mn);
    for (int i = lane; i < n; i += 32)
      if (mask_test(rowCov, i)) ra[i] += mn;
#pragma unroll
    for (int t = 0; t < 4; t++)
      if (ucol[t]) cs[t] += mn;
    // zeros of covered rows in covered columns become positive (row-major and column-major masks)
#pragma unroll
    for (int q = 0; q < 4; q++)
      if ((own >> q) & 1u) {
#pragma unroll
        for (int t = 0; t < 4; t++) Z[q][t] &= ~colCov[t];
      }
#pragma unroll
    for (int t = 0; t < 4; t++)
      if (ucolValid[t] && ((colCov[t] >> lane) & 1u)) {
        uint32_t* z = sm.zt + (lane + 32 * t) * 4;
#pragma unroll
        for (int w = 0; w < 4; w++) z[w] &= ~rowCov[w];
      }
    __syncwarp();
    // uncovered x uncovered cells that were equal to m become zeros
    for (int l = 0; l < 32 && 4 * l < n; l++) {
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int i = 4 * l + q;
        if (i >= n) break;
        if (mask_test(rowCov, i)) continue;
        const Pot rai = ra[i];
        const CellT* row = C + (size_t)i * n;
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const bool z = ucol[t] && ((Pot)row[lane + 32 * t] + rai - cs[t] == 0);
          const unsigned bz = __ballot_sync(full, z);
          if (lane == l) Z[q][t] |= bz;
          if (z) sm.zt[(lane + 32 * t) * 4 + (i >> 5)] |= 1u << (i & 31);
        }
      }
    }
    __syncwarp();
    // the masks changed; the covers did not: rebuild the keys of the live rows
#pragma unroll
    for (int q = 0; q < 4; q++) {
      k[q] = kBig;
      if ((live >> q) & 1u) {
#pragma unroll
        for (int t = 3; t >= 0; t--) {
          const uint32_t cand = Z[q][t] & ~colCov[t];
          if (cand) k[q] = (q << 8) | (32 * t + __ffs(cand) - 1);
        }
      }
    }
  }
  __syncwarp();
}

}  // namespace voda
