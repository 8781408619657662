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
//   * every row has a 128-bit ZERO MASK (shared memory, row-major and column-major copies), so
//     "first uncovered zero in row-major order" (step 4, ~n^2/16 times per matrix) never scans
//     the matrix; step 6 patches the masks (covered rows lose their zeros in covered columns;
//     uncovered x uncovered cells equal to m gain one); each lane owns 4 consecutive rows;
//   * stars / primes are starCol[row], starRow[col], primeCol[row] in shared memory, and so is
//     a column-major copy of the zero masks (which rows have a zero in column j); every owned
//     row carries a key (row, first uncovered zero column, column of its star), so a step-4
//     iteration is one redux over the keys, one shared-memory load and a dozen ALU operations;
//   * with 32-bit potentials step 6 takes the minimum row by row (a redux each) and then looks
//     for new zeros only in the rows whose minimum is the global one.
#pragma once

#include "common.cuh"

namespace voda {

constexpr int kMunkresSlackRows = 7;  // the vector pass of step 6 loads whole groups of 8 rows

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

// Four adjacent cells of a row in one load (the row must start on a multiple of four cells).
template <typename CellT>
struct Vec4;
template <>
struct Vec4<uint8_t> {
  using type = uint32_t;
  static __device__ __forceinline__ void unpack(uint32_t w, int (&b)[4]) {
    b[0] = (int)__byte_perm(w, 0, 0x4440);
    b[1] = (int)__byte_perm(w, 0, 0x4441);
    b[2] = (int)__byte_perm(w, 0, 0x4442);
    b[3] = (int)__byte_perm(w, 0, 0x4443);
  }
};
template <>
struct Vec4<uint16_t> {
  using type = uint2;
  static __device__ __forceinline__ void unpack(uint2 w, int (&b)[4]) {
    b[0] = (int)(w.x & 0xFFFFu);
    b[1] = (int)(w.x >> 16);
    b[2] = (int)(w.y & 0xFFFFu);
    b[3] = (int)(w.y >> 16);
  }
};
template <>
struct Vec4<int32_t> {
  using type = int4;
  static __device__ __forceinline__ void unpack(int4 w, int (&b)[4]) { b[0] = w.x; b[1] = w.y; b[2] = w.z; b[3] = w.w; }
};
template <>
struct Vec4<int64_t> {  // (64-bit potentials: not used by the vector path)
  using type = int4;
  static __device__ __forceinline__ void unpack(int4 w, int (&b)[4]) { b[0] = w.x; b[1] = w.y; b[2] = w.z; b[3] = w.w; }
};

struct MunkresSmem {
  short* starCol;   // [128] column of the star in row i, -1 if none (8-byte aligned)
  short* starRow;   // [128] row of the star in column j, -1 if none
  unsigned char* primeCol;  // [128] column of the prime in row i (valid for rows covered in this phase)
  void* ra;         // [128] row potentials (Pot)
  void* cs;         // [128] column potentials (Pot), 16-byte aligned
  uint32_t* zr;     // [128][4] row-major zero masks: bit j of zr[i] = (value(i, j) == 0) (16-byte aligned)
  uint32_t* zt;     // [128][4] column-major zero masks: bit i of zt[j] = (value(i, j) == 0);
                    //          the caller fills both (rows and columns beyond n: no bits)
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

// Solve.  C: n*n cells (row stride n) already holding rowmax_i - A[i][j], followed by at least
// kMunkresSlackRows * n readable cells when the potentials are 32-bit; sm.zr / sm.zt: its zero masks.  On return sm.starCol[i] is the column assigned to row i.  Must be called by one full warp.
// Pot: the type of the potentials; they are bounded by n * max(C), and a caller that knows its
// cells stay below 2^31 / n can ask for 32 bits with a wider CellT.
template <typename CellT, typename Pot = typename MunkresTypes<CellT>::Pot>
__device__ void munkres_solve(const CellT* __restrict__ C, int n, const MunkresSmem& sm) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  Pot* ra = static_cast<Pot*>(sm.ra);
  for (int i = lane; i < 128; i += 32) {
    ra[i] = 0;
    static_cast<Pot*>(sm.cs)[i] = 0;
    sm.starCol[i] = -1;
    sm.starRow[i] = -1;
    sm.primeCol[i] = 0;
  }
  __syncwarp();

  // ---- step 2: star the first zero of each row whose column holds no star yet, rows in order
  uint4* zr4 = reinterpret_cast<uint4*>(sm.zr);
  uint32_t starred[4] = {0, 0, 0, 0};  // columns holding a star (uniform)
  int nstar = 0;
  for (int i = 0; i < n; i++) {
    const uint4 z = zr4[i];  // the same address in every lane
    const uint32_t c0 = z.x & ~starred[0], c1 = z.y & ~starred[1], c2 = z.z & ~starred[2], c3 = z.w & ~starred[3];
    int col = -1;
    if (c3) col = 96 + __ffs(c3) - 1;
    if (c2) col = 64 + __ffs(c2) - 1;
    if (c1) col = 32 + __ffs(c1) - 1;
    if (c0) col = __ffs(c0) - 1;
    if (col >= 0) {
      mask_set(starred, col);
      nstar++;
      if (lane == 0) {
        sm.starCol[i] = (short)col;
        sm.starRow[col] = (short)i;
      }
    }
  }
  __syncwarp();

  // Covers.  Rows: each lane keeps the state of the 4 rows it owns (in RB, below).  Columns: inside one
  // step-4 phase the covered columns are the starred ones minus the star columns of the rows
  // covered so far, so nothing is stored per iteration; step 6 rebuilds the mask when it needs it.
  //
  // K[q] = (row << 16) | (first uncovered zero column of owned row q << 8) | (column of the row's
  // star, 0xFF if it has none), or kNone when the row is covered or has no uncovered zero.  RB[q] is
  // the key of owned row q without its column; `live` has a bit per owned row that is not covered.  Inside
  // a phase columns only get UNcovered and rows only get covered, so uncovering column sc turns
  // K[q] into min(K[q], key with column sc) for the live rows that have a zero there — read from
  // the column-major zero masks sm.zt — and the search "first uncovered zero in row-major order"
  // never rescans anything: it is the minimum key of the warp (one redux), which carries the row,
  // the column and the row's star.
  constexpr int kNone = 0x7fffffff;
  uint32_t colCov[4] = {starred[0], starred[1], starred[2], starred[3]};  // step 3
  bool ucolValid[4];
#pragma unroll
  for (int t = 0; t < 4; t++) ucolValid[t] = lane + 32 * t < n;
  int K[4], RB[4];
  uint32_t live = 0xFu;
  auto keys_from_masks = [&]() {
#pragma unroll
    for (int q = 0; q < 4; q++) {
      K[q] = kNone;
      if ((live >> q) & 1u) {
        const uint4 z = zr4[4 * lane + q];
        const uint32_t c0 = z.x & ~colCov[0], c1 = z.y & ~colCov[1], c2 = z.z & ~colCov[2], c3 = z.w & ~colCov[3];
        if (c3) K[q] = RB[q] | ((96 + __ffs(c3) - 1) << 8);
        if (c2) K[q] = RB[q] | ((64 + __ffs(c2) - 1) << 8);
        if (c1) K[q] = RB[q] | ((32 + __ffs(c1) - 1) << 8);
        if (c0) K[q] = RB[q] | ((__ffs(c0) - 1) << 8);
      }
    }
  };
  auto reset_phase = [&]() {  // covers = {starred columns}, no row; keys from the zero masks
    live = 0xFu;
    const short4 sc4 = *reinterpret_cast<const short4*>(sm.starCol + 4 * lane);
    RB[0] = ((4 * lane + 0) << 16) | (sc4.x & 0xFF);
    RB[1] = ((4 * lane + 1) << 16) | (sc4.y & 0xFF);
    RB[2] = ((4 * lane + 2) << 16) | (sc4.z & 0xFF);
    RB[3] = ((4 * lane + 3) << 16) | (sc4.w & 0xFF);
    keys_from_masks();
  };
  reset_phase();
  // shared-memory addresses and shifts of the step-4 loop, made opaque so that they stay in
  // registers (recomputing them costs a fifth of an iteration)
  uint32_t ztAddr = (uint32_t)__cvta_generic_to_shared(sm.zt) + 4u * (uint32_t)(lane >> 3);
  uint32_t primeAddr = (uint32_t)__cvta_generic_to_shared(sm.primeCol);
  uint32_t ztShift = 4u * (uint32_t)(lane & 7);
  asm volatile("" : "+r"(ztAddr), "+r"(primeAddr), "+r"(ztShift));

  while (nstar < n) {
    // ---- step 4: first uncovered zero in row-major order
    const int red = __reduce_min_sync(full, min(min(K[0], K[1]), min(K[2], K[3])));
    if ((red & 0x80) == 0) {  // it lies in a row that holds a star: prime it, cover its row, uncover the star's column
      const int sc = red & 0x7f;
      const int row = red >> 16;
      uint32_t word;
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(word) : "r"(ztAddr + 16u * (uint32_t)sc));
      // every lane stores the same value: no predicate to keep
      asm volatile("st.shared.u8 [%0], %1;" ::"r"(primeAddr + (uint32_t)row), "r"(red >> 8) : "memory");
      // the chosen key is some lane's K[q] of the row that gets covered (two predicated operations per
      // q; left to itself the compiler builds the new `live` from four selects)
#define VODA_MK_COVER(q)                                                                               \
  asm("{.reg .pred p; setp.eq.s32 p, %0, %2; @p mov.s32 %0, 0x7fffffff; @p and.b32 %1, %1, %3;}" \
      : "+r"(K[q]), "+r"(live)                                                                        \
      : "r"(red), "n"(~(1u << q)))
      VODA_MK_COVER(0);
      VODA_MK_COVER(1);
      VODA_MK_COVER(2);
      VODA_MK_COVER(3);
#undef VODA_MK_COVER
      // (the column's star sits in the row just covered: without `live` that bit would always hit)
      const uint32_t hits = (word >> ztShift) & live;
      if (hits) {  // zeros are sparse: most lanes, most of the time, have none in that column
        const int cb = sc << 8;
#pragma unroll
        for (int q = 0; q < 4; q++)
          if ((hits >> q) & 1u) K[q] = min(K[q], RB[q] + cb);
      }
      continue;
    }
    if (red != kNone) {
      // ---- step 5: augment along the alternating path that starts at the prime (row, col)
      const int row = red >> 16, col = (red >> 8) & 0xFF;
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
    // ---- step 6 needs whole masks.  Columns: the starred ones minus the star columns of the covered rows
    __syncwarp();
    {
      uint32_t unc[4] = {0, 0, 0, 0};
#pragma unroll
      for (int q = 0; q < 4; q++)
        if (!((live >> q) & 1u)) mask_set(unc, RB[q] & 0x7f);  // a covered row holds a star
#pragma unroll
      for (int t = 0; t < 4; t++) colCov[t] = starred[t] & ~__reduce_or_sync(full, unc[t]);
    }
    const uint32_t ownCov = ~live & 0xFu;
    // rebuild the uniform row-cover mask from the per-lane bits (rows 4*lane .. 4*lane+3)
    uint32_t rowCov[4];
    {
      const uint32_t nib = ownCov << (4 * (lane & 7));
#pragma unroll
      for (int w = 0; w < 4; w++) rowCov[w] = __reduce_or_sync(full, (lane >> 3) == w ? nib : 0u);
    }
    const uint32_t own = ownCov;
    // ---- step 6: m = smallest uncovered value; covered rows += m, uncovered columns -= m.
    // Column potentials live in shared memory (sm.cs): nothing of step 6 stays in registers
    // during step 4.
    Pot* csm = static_cast<Pot*>(sm.cs);
    bool ucol[4];  // columns lane + 32 t that are valid and not covered
#pragma unroll
    for (int t = 0; t < 4; t++) ucol[t] = ucolValid[t] && !((colCov[t] >> lane) & 1u);
    const bool vec = sizeof(Pot) == 4 && (n & 3) == 0;
    Pot mn;
    // 32-bit potentials: the minimum of every uncovered row is one redux, kept by lane l for rows
    // l + 32 w (rmin[w]), so that afterwards only the rows whose minimum IS m are searched for new zeros
    int rmin[4] = {0x7fffffff, 0x7fffffff, 0x7fffffff, 0x7fffffff};
    int d[4] = {0, 0, 0, 0};  // vector path: -potential of column 4*lane + t, or a large value when covered
    const bool colsValid = 4 * lane < n;
    if constexpr (sizeof(Pot) == 4) {
      if (vec) {
        // a lane reads 4 adjacent cells of a row in one load (columns 4*lane .. 4*lane+3), kRows rows in flight
        constexpr int kRows = kMunkresSlackRows + 1;
        using VecT = typename Vec4<CellT>::type;
        const int n4 = n >> 2, laneC = min(lane, n4 - 1);
        {
          const int4 c4 = colsValid ? reinterpret_cast<const int4*>(csm)[lane] : make_int4(0, 0, 0, 0);
          const uint32_t cw = (lane >> 3) == 0 ? colCov[0] : (lane >> 3) == 1 ? colCov[1] : (lane >> 3) == 2 ? colCov[2] : colCov[3];
          const uint32_t cov = (cw >> (4 * (lane & 7))) & 0xFu;
          d[0] = (colsValid && !(cov & 1u)) ? -c4.x : 0x3fffffff;
          d[1] = (colsValid && !(cov & 2u)) ? -c4.y : 0x3fffffff;
          d[2] = (colsValid && !(cov & 4u)) ? -c4.z : 0x3fffffff;
          d[3] = (colsValid && !(cov & 8u)) ? -c4.w : 0x3fffffff;
        }
#pragma unroll
        for (int w = 0; w < 4; w++) {
#pragma unroll 1
          for (int c8 = 0; c8 < 32 && 32 * w + c8 < n; c8 += kRows) {
            const int i0 = 32 * w + c8;
            // rows beyond n count as covered; every load is unconditional (covered columns and lanes
            // beyond the matrix carry a large d[t]; the caller's C has kRows - 1 rows of readable slack)
            const uint32_t rw = (rowCov[w] >> c8) | (n - i0 >= kRows ? 0u : ~0u << (n - i0));
            const VecT* p = reinterpret_cast<const VecT*>(C) + i0 * n4 + laneC;
            VecT raw[kRows];
            int rai[kRows];
#pragma unroll
            for (int u = 0; u < kRows; u++) {
              raw[u] = p[u * n4];
              rai[u] = (int)ra[i0 + u];  // ra has 128 entries
            }
#pragma unroll
            for (int u = 0; u < kRows; u++) {
              if ((rw >> u) & 1u) continue;
              int b[4];
              Vec4<CellT>::unpack(raw[u], b);
              const int v = min(min(b[0] + rai[u] + d[0], b[1] + rai[u] + d[1]), min(b[2] + rai[u] + d[2], b[3] + rai[u] + d[3]));
              const int rm = __reduce_min_sync(full, v);
              if (lane == c8 + u) rmin[w] = rm;
            }
          }
        }
      } else {
        int cs[4];
#pragma unroll
        for (int t = 0; t < 4; t++) cs[t] = ucolValid[t] ? (int)csm[lane + 32 * t] : 0;
#pragma unroll
        for (int w = 0; w < 4; w++) {
#pragma unroll 1
          for (int l = 0; l < 32 && 32 * w + l < n; l++) {
            if ((rowCov[w] >> l) & 1u) continue;
            const int i = 32 * w + l;
            const int rai = (int)ra[i];
            const CellT* row = C + (size_t)i * n;
            int v = 0x7fffffff;
#pragma unroll
            for (int t = 0; t < 4; t++)
              if (ucol[t]) v = min(v, (int)row[lane + 32 * t] + rai - cs[t]);
            const int rm = __reduce_min_sync(full, v);
            if (lane == l) rmin[w] = rm;
          }
        }
      }
      mn = (Pot)__reduce_min_sync(full, min(min(rmin[0], rmin[1]), min(rmin[2], rmin[3])));
    } else {
      Pot cs[4];
#pragma unroll
      for (int t = 0; t < 4; t++) cs[t] = ucolValid[t] ? csm[lane + 32 * t] : 0;
      mn = (Pot)0x7fffffffffffffffLL;
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
    }
    for (int i = lane; i < n; i += 32)
      if (mask_test(rowCov, i)) ra[i] += mn;
    Pot cs[4];  // the updated potentials of columns lane + 32 t
#pragma unroll
    for (int t = 0; t < 4; t++) {
      cs[t] = ucolValid[t] ? csm[lane + 32 * t] : 0;
      if (ucol[t]) {
        cs[t] += mn;
        csm[lane + 32 * t] = cs[t];
      }
    }
    // zeros of covered rows in covered columns become positive (row-major and column-major masks)
#pragma unroll
    for (int q = 0; q < 4; q++)
      if ((own >> q) & 1u) {
        uint4 z = zr4[4 * lane + q];
        z.x &= ~colCov[0];
        z.y &= ~colCov[1];
        z.z &= ~colCov[2];
        z.w &= ~colCov[3];
        zr4[4 * lane + q] = z;
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
    if constexpr (sizeof(Pot) == 4) {
      if (vec) {
#pragma unroll
        for (int t = 0; t < 4; t++) d[t] = d[t] == 0x3fffffff ? d[t] : d[t] - (int)mn;  // the new potentials
      }
#pragma unroll
      for (int w = 0; w < 4; w++) {
        unsigned rows = __ballot_sync(full, rmin[w] == (int)mn);  // only these rows hold such a cell
        while (rows) {
          const int l = __ffs(rows) - 1;
          rows &= rows - 1;
          const int i = 32 * w + l;
          const int rai = (int)ra[i];
          if (vec) {
            int b[4];
            Vec4<CellT>::unpack(colsValid ? reinterpret_cast<const typename Vec4<CellT>::type*>(C + (size_t)i * n)[lane]
                                          : typename Vec4<CellT>::type{}, b);
            const uint32_t nib = (b[0] + rai + d[0] == 0 ? 1u : 0u) | (b[1] + rai + d[1] == 0 ? 2u : 0u) |
                                 (b[2] + rai + d[2] == 0 ? 4u : 0u) | (b[3] + rai + d[3] == 0 ? 8u : 0u);
            if (nib) {  // columns 4*lane + t: this lane's nibble of the row mask, and its own column masks
              atomicOr(&sm.zr[i * 4 + (lane >> 3)], nib << (4 * (lane & 7)));
#pragma unroll
              for (int t = 0; t < 4; t++)
                if ((nib >> t) & 1u) sm.zt[(4 * lane + t) * 4 + w] |= 1u << l;
            }
          } else {
            const CellT* row = C + (size_t)i * n;
#pragma unroll
            for (int t = 0; t < 4; t++) {
              const bool z = ucol[t] && ((int)row[lane + 32 * t] + rai - (int)cs[t] == 0);
              const unsigned bz = __ballot_sync(full, z);
              if (lane == 0 && bz) sm.zr[i * 4 + t] |= bz;
              if (z) sm.zt[(lane + 32 * t) * 4 + w] |= 1u << l;
            }
          }
        }
      }
    } else {
      // 64-bit potentials serve the widest cells, and those may hold the reference's WRAPPED values
      // (MaxInt64 - score with mixed signs), negative ones included: then m can be negative and a
      // covered x covered cell equal to -m becomes a zero as well.  No case analysis here: both masks
      // are rebuilt from the values, which is what the reference's `== 0` tests see.
      for (int x = lane; x < 128 * 4; x += 32) sm.zt[x] = 0;
      __syncwarp();
      for (int i = 0; i < n; i++) {
        const Pot rai = ra[i];
        const CellT* row = C + (size_t)i * n;
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const bool z = ucolValid[t] && ((Pot)row[lane + 32 * t] + rai - cs[t] == 0);
          const unsigned bz = __ballot_sync(full, z);
          if (lane == 0) sm.zr[i * 4 + t] = bz;
          if (z) sm.zt[(lane + 32 * t) * 4 + (i >> 5)] |= 1u << (i & 31);
        }
      }
    }
    __syncwarp();
    // the masks changed; the covers did not: rebuild the keys of the live rows
    keys_from_masks();
  }
  __syncwarp();
}

}  // namespace voda
