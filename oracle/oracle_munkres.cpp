// oracle_munkres.cpp — CPU restatement of munkres.ComputeMunkresMax.
//
// TEST INFRASTRUCTURE, NOT PRODUCT (see voda_oracle.h).
//
// PARITY UNPINNED.  The arithmetic lives in a third-party Go module that is NOT in the
// reference tree: github.com/heyfey/munkres, pinned at v0.0.0-20210705070356-d4f16a25250d
// (reference go.mod:7, go.sum:337-338; no vendor/ directory, no module cache on this image).
// Its only call site is pkg/placement/placement_manager.go:505-512:
//     m := munkres.NewMatrix(size); m.A = scoringMatrix   // []int64, row-major size*row+col
//     result := munkres.ComputeMunkresMax(m)               // []RowCol{Row, Col}
// and no reference test pins its output.  What follows restates the PUBLISHED algorithm that
// module's lineage implements (a Go port of the classic Clapper/Pilgrim six-step Munkres:
// work on a copy of A, row/column covers, an n*n mark array {unset, starred, primed}, every
// search a row-major scan from the top-left) — the step order of SURVEY.md Appendix C.  What
// can be checked is checked in tests/: the result is a permutation, its score equals
// scipy.optimize.linear_sum_assignment(maximize=True), and the Appendix C tie-break KATs.
// Identity with the real module's tie-breaking cannot be verified offline.

#include <cstdint>
#include <limits>
#include <thread>
#include <vector>

#include "voda_oracle.h"

namespace {

enum Mark : uint8_t { Unset = 0, Starred = 1, Primed = 2 };

thread_local int64_t g_step4 = 0, g_step5 = 0, g_step6 = 0;

struct Context {
  int n;
  std::vector<int64_t> A;  // working copy
  std::vector<char> rowCovered, colCovered;
  std::vector<uint8_t> marked;
  int z0row = 0, z0col = 0;
  std::vector<int> rowPath, colPath;
};

void clearCovers(Context& c) {
  c.rowCovered.assign((size_t)c.n, 0);
  c.colCovered.assign((size_t)c.n, 0);
}

// Step 1: subtract each row's minimum from the row.
void step1(Context& c) {
  const int n = c.n;
  for (int i = 0; i < n; i++) {
    int64_t mn = std::numeric_limits<int64_t>::max();
    for (int j = 0; j < n; j++) mn = std::min(mn, c.A[(size_t)i * n + j]);
    for (int j = 0; j < n; j++) c.A[(size_t)i * n + j] -= mn;
  }
}

// Step 2: star a zero when its row and column hold no star yet (row-major scan).
void step2(Context& c) {
  const int n = c.n;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++)
      if (c.A[(size_t)i * n + j] == 0 && !c.colCovered[(size_t)j] && !c.rowCovered[(size_t)i]) {
        c.marked[(size_t)i * n + j] = Starred;
        c.colCovered[(size_t)j] = 1;
        c.rowCovered[(size_t)i] = 1;
      }
  clearCovers(c);
}

// Step 3: cover every column that holds a star; all n covered => done.
bool step3(Context& c) {
  const int n = c.n;
  int count = 0;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++)
      if (c.marked[(size_t)i * n + j] == Starred) {
        c.colCovered[(size_t)j] = 1;
        count++;
      }
  return count >= n;
}

bool findAZero(const Context& c, int& row, int& col) {
  const int n = c.n;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++)
      if (c.A[(size_t)i * n + j] == 0 && !c.rowCovered[(size_t)i] && !c.colCovered[(size_t)j]) {
        row = i;
        col = j;
        return true;
      }
  return false;
}
int findInRow(const Context& c, int row, Mark m) {
  for (int j = 0; j < c.n; j++)
    if (c.marked[(size_t)row * c.n + j] == m) return j;
  return -1;
}
int findStarInCol(const Context& c, int col) {
  for (int i = 0; i < c.n; i++)
    if (c.marked[(size_t)i * c.n + col] == Starred) return i;
  return -1;
}

// Step 4: prime uncovered zeros until one has no star in its row (-> step 5, returns true)
// or none is left (-> step 6, returns false).
bool step4(Context& c) {
  for (;;) {
    int row, col;
    g_step4++;
    if (!findAZero(c, row, col)) return false;
    c.marked[(size_t)row * c.n + col] = Primed;
    int starCol = findInRow(c, row, Starred);
    if (starCol >= 0) {
      c.rowCovered[(size_t)row] = 1;
      c.colCovered[(size_t)starCol] = 0;
    } else {
      c.z0row = row;
      c.z0col = col;
      return true;
    }
  }
}

// Step 5: alternating path from z0 (star in column, then prime in that row), flip, reset.
void step5(Context& c) {
  g_step5++;
  int count = 0;
  c.rowPath[0] = c.z0row;
  c.colPath[0] = c.z0col;
  for (;;) {
    int row = findStarInCol(c, c.colPath[(size_t)count]);
    if (row < 0) break;
    count++;
    c.rowPath[(size_t)count] = row;
    c.colPath[(size_t)count] = c.colPath[(size_t)count - 1];
    int col = findInRow(c, c.rowPath[(size_t)count], Primed);
    count++;
    c.rowPath[(size_t)count] = c.rowPath[(size_t)count - 1];
    c.colPath[(size_t)count] = col;
  }
  for (int i = 0; i <= count; i++) {
    size_t off = (size_t)c.rowPath[(size_t)i] * c.n + (size_t)c.colPath[(size_t)i];
    c.marked[off] = (c.marked[off] == Starred) ? Unset : Starred;
  }
  clearCovers(c);
  for (auto& m : c.marked)
    if (m == Primed) m = Unset;
}

// Step 6: smallest uncovered value; add to covered rows, subtract from uncovered columns.
void step6(Context& c) {
  g_step6++;
  const int n = c.n;
  int64_t mn = std::numeric_limits<int64_t>::max();
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++)
      if (!c.rowCovered[(size_t)i] && !c.colCovered[(size_t)j]) mn = std::min(mn, c.A[(size_t)i * n + j]);
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      if (c.rowCovered[(size_t)i]) c.A[(size_t)i * n + j] += mn;
      if (!c.colCovered[(size_t)j]) c.A[(size_t)i * n + j] -= mn;
    }
}

void solve(int n, const int64_t* A, int32_t* colOfRow) {
  Context c;
  c.n = n;
  c.A.resize((size_t)n * n);
  // maximise: A' = C - A with C = MaxInt64 (any constant gives the same matrix after step 1)
  for (size_t i = 0; i < (size_t)n * n; i++) c.A[i] = std::numeric_limits<int64_t>::max() - A[i];
  c.marked.assign((size_t)n * n, Unset);
  c.rowPath.assign((size_t)2 * n + 2, 0);
  c.colPath.assign((size_t)2 * n + 2, 0);
  clearCovers(c);
  step1(c);
  step2(c);
  for (;;) {
    if (step3(c)) break;
    while (!step4(c)) step6(c);
    step5(c);
  }
  for (int i = 0; i < n; i++) colOfRow[i] = -1;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++)
      if (c.marked[(size_t)i * n + j] == Starred && colOfRow[i] < 0) colOfRow[i] = j;
}

void run(int32_t n, const int64_t* A, int32_t* out, int32_t* status, int32_t b0, int32_t b1) {
  for (int32_t b = b0; b < b1; b++) {
    solve(n, A + (size_t)b * n * n, out + (size_t)b * n);
    if (status) status[b] = VODA_ST_OK;
  }
}

}  // namespace

// exposed to oracle_placement.cpp
void voda_oracle_munkres_solve(int n, const int64_t* A, int32_t* colOfRow) { solve(n, A, colOfRow); }

extern "C" int voda_oracle_munkres_max(int32_t n, int32_t batch, const int64_t* A,
                                       int32_t* out_col_of_row, int32_t* out_status,
                                       int32_t threads) {
  if (n < 0 || batch < 0 || (!A && n > 0 && batch > 0)) return VODA_E_INVALID;
  if (n == 0) {
    for (int32_t b = 0; b < batch; b++) if (out_status) out_status[b] = VODA_ST_OK;
    return VODA_OK;
  }
  if (threads <= 1 || batch < 2) {
    g_step4 = g_step5 = g_step6 = 0;
    run(n, A, out_col_of_row, out_status, 0, batch);
    return VODA_OK;
  }
  std::vector<std::thread> pool;
  for (int32_t t = 0; t < threads; t++) {
    int32_t b0 = (int32_t)((int64_t)batch * t / threads), b1 = (int32_t)((int64_t)batch * (t + 1) / threads);
    pool.emplace_back(run, n, A, out_col_of_row, out_status, b0, b1);
  }
  for (auto& th : pool) th.join();
  return VODA_OK;
}

extern "C" void voda_oracle_munkres_counters(int64_t* step4, int64_t* step6, int64_t* step5) {
  if (step4) *step4 = g_step4;
  if (step6) *step6 = g_step6;
  if (step5) *step5 = g_step5;
}
