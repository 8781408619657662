// oracle_dense.cpp — array-based CPU restatement of pkg/algorithm (same results, bit for bit,
// as oracle_algorithm.cpp; no strings, no hash maps).
//
// TEST INFRASTRUCTURE, NOT PRODUCT (see voda_oracle.h).  It has two jobs: (1) a second,
// independently written restatement that the structure-faithful one is cross-checked against
// in tests/ (it uses the derived forms the CUDA kernels rely on — SURVEY.md A.3's
// "(gain desc, priority asc, stamp desc)" argmax for ElasticTiresias — so those derivations
// are tested on the CPU too); (2) the "what a tuned CPU implementation costs" figure that
// bench.py prints beside the structure-faithful timing.  Reference file:line citations are
// on each function; paths are relative to the reference tree.

#include <cstring>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <numeric>
#include <thread>
#include <vector>

#include "voda_oracle.h"

namespace {

struct Round {
  const voda_alloc_batch* b;
  int32_t o, n;
  long total;
  long W;
  long mn(int i) const { return b->min_proc ? b->min_proc[o + i] : 0; }
  long mx(int i) const { return b->max_proc ? b->max_proc[o + i] : 0; }
  long np(int i) const { return b->num_proc ? b->num_proc[o + i] : 0; }
  long prio(int i) const { return b->priority ? b->priority[o + i] : 0; }
  float rem(int i) const { return b->remaining_sec ? b->remaining_sec[o + i] : 0.0f; }
  // Info.Speedup[fmt.Sprint(g)]: a key outside the dense table is a map miss => 0
  float S(int i, long g) const {
    if (!b->speedup || g < 0 || g >= W) return 0.0f;
    return b->speedup[(b->speedup_off ? (size_t)b->speedup_off[o + i] : (size_t)(o + i) * (size_t)W) + (size_t)g];
  }
  bool submitBefore(int x, int y) const {
    int64_t sx = b->submit_sec ? b->submit_sec[o + x] : 0, sy = b->submit_sec ? b->submit_sec[o + y] : 0;
    int32_t nx = b->submit_nsec ? b->submit_nsec[o + x] : 0, ny = b->submit_nsec ? b->submit_nsec[o + y] : 0;
    return sx < sy || (sx == sy && nx < ny);
  }
  bool firstBefore(int x, int y) const {
    int64_t sx = b->first_sec ? b->first_sec[o + x] : 0, sy = b->first_sec ? b->first_sec[o + y] : 0;
    int32_t nx = b->first_nsec ? b->first_nsec[o + x] : 0, ny = b->first_nsec ? b->first_nsec[o + y] : 0;
    return sx < sy || (sx == sy && nx < ny);
  }
};

// validateResult, pkg/algorithm/utils.go:18-42 with the tie rule of include/voda_accel.h
int validate(const Round& R, const std::vector<long>& r, const std::vector<char>* present = nullptr) {
  long sum = 0;
  int worst = 0;
  for (int i = 0; i < R.n; i++) {
    if (present && !(*present)[(size_t)i]) continue;
    long n = r[(size_t)i];
    int st = 0;
    if (n < 0) st = VODA_ST_NEGATIVE;
    else if (n > 0 && n < R.mn(i)) st = VODA_ST_LESS_THAN_MIN;
    else if (n > R.mx(i)) st = VODA_ST_EXCEEDED_MAX;
    if (st && (!worst || st < worst)) worst = st;
    sum += n;
  }
  if (worst) return worst;
  return sum > R.total ? VODA_ST_EXCEEDED_TOTAL : VODA_ST_OK;
}

std::vector<int> sortedBySubmit(const Round& R) {
  std::vector<int> ord((size_t)R.n);
  std::iota(ord.begin(), ord.end(), 0);
  std::stable_sort(ord.begin(), ord.end(), [&](int x, int y) { return R.submitBefore(x, y); });
  return ord;
}
std::vector<int> sortedByRemaining(const Round& R) {
  std::vector<int> ord((size_t)R.n);
  std::iota(ord.begin(), ord.end(), 0);
  std::stable_sort(ord.begin(), ord.end(), [&](int x, int y) { return R.rem(x) < R.rem(y); });
  return ord;
}

// FIFO fifo.go:25-52 / SRJF srjf.go:25-52
int simple(const Round& R, std::vector<long>& r, bool srjf) {
  auto ord = srjf ? sortedByRemaining(R) : sortedBySubmit(R);
  long free = R.total;
  for (int i : ord)
    if (free >= R.mn(i)) { r[(size_t)i] = R.mn(i); free -= R.mn(i); }
  return validate(R, r);
}

// ElasticFIFO elastic_fifo.go:25-75 / ElasticSRJF elastic_srjf.go:25-72
int elasticFill(const Round& R, std::vector<long>& r, bool srjf) {
  auto ord = srjf ? sortedByRemaining(R) : sortedBySubmit(R);
  std::vector<char> sat((size_t)R.n, 0);
  long free = R.total, unsat = 0;
  for (int i : ord) {
    if (free >= R.mn(i)) {
      r[(size_t)i] = R.mn(i);
      free -= R.mn(i);
      if (!srjf && r[(size_t)i] == R.mx(i)) sat[(size_t)i] = 1;
    } else {
      sat[(size_t)i] = 1;
    }
    if (!sat[(size_t)i]) unsat++;
  }
  while (free > 0 && unsat > 0) {
    for (int i : ord) {
      if (r[(size_t)i] < R.mx(i) || !sat[(size_t)i]) {
        r[(size_t)i] += 1;
        free -= 1;
        if (r[(size_t)i] == R.mx(i) && !sat[(size_t)i]) { sat[(size_t)i] = 1; unsat--; }
        if (free == 0) break;
      }
    }
  }
  return validate(R, r);
}

// Tiresias tiresias.go:51-99
int tiresias(const Round& R, std::vector<long>& r, long Q) {
  std::vector<char> present((size_t)R.n, 0);
  long free = R.total;
  for (long p = 0; p < Q; p++) {
    std::vector<int> q;
    for (int i = 0; i < R.n; i++) if (R.prio(i) == p) q.push_back(i);
    std::stable_sort(q.begin(), q.end(), [&](int x, int y) { return R.firstBefore(x, y); });
    for (int i : q) {
      present[(size_t)i] = 1;
      if (free >= R.np(i)) { r[(size_t)i] = R.np(i); free -= R.np(i); }
    }
  }
  return validate(R, r, &present);
}

// ElasticTiresias elastic_tiresias.go:37-159, greedy phase in the argmax form of SURVEY.md A.3
// (Q5): the head of the doubly stable-sorted list equals the maximum of
// (gain desc, Priority asc, stamp desc), stamp = -(position in the pruned list) at first and
// the pick number once a job has been picked.
bool sameBits(float a, float b) {
  uint32_t x, y;
  std::memcpy(&x, &a, 4);
  std::memcpy(&y, &b, 4);
  return x == y;
}

int elasticTiresias(const Round& R, std::vector<long>& r, long Q, long compaction) {
  const int n = R.n;
  std::vector<float> gain((size_t)n);
  long free = R.total, pend = n;
  for (int i = 0; i < n; i++) gain[(size_t)i] = R.S(i, R.mn(i)) / (float)R.mn(i);
  std::vector<std::vector<int>> queues((size_t)Q);
  for (long p = 0; p < Q; p++) {
    auto& q = queues[(size_t)p];
    for (int i = 0; i < n; i++) if (R.prio(i) == p) q.push_back(i);
    std::stable_sort(q.begin(), q.end(), [&](int x, int y) { return R.firstBefore(x, y); });
  }
  for (long p = 0; p < Q; p++)
    for (int i : queues[(size_t)p])
      if (free >= R.np(i)) {
        r[(size_t)i] = R.np(i); free -= R.np(i); pend -= 1;
        gain[(size_t)i] = R.S(i, r[(size_t)i] + 1) - R.S(i, r[(size_t)i]);
      }
  if (pend > compaction)
    for (long p = 1; p < Q; p++)
      for (int i : queues[(size_t)p])
        if (r[(size_t)i] != 0) {
          free += r[(size_t)i] - R.mn(i);
          r[(size_t)i] = R.mn(i);
          gain[(size_t)i] = R.S(i, r[(size_t)i] + 1) - R.S(i, r[(size_t)i]);
        }
  std::vector<char> live((size_t)n, 0);
  std::vector<long> stamp((size_t)n, 0);
  long nlive = 0;
  for (int i = 0; i < n; i++)
    if (!(r[(size_t)i] >= R.mx(i) || free < R.mn(i))) { live[(size_t)i] = 1; stamp[(size_t)i] = -(nlive++); }
  long it = 0;
  while (free > 0 && nlive > 0) {
    int best = -1;
    for (int i = 0; i < n; i++) {
      if (!live[(size_t)i]) continue;
      if (best < 0) { best = i; continue; }
      float gi = gain[(size_t)i], gb = gain[(size_t)best];
      if (gi > gb || (gi == gb && (R.prio(i) < R.prio(best) ||
                                   (R.prio(i) == R.prio(best) && stamp[(size_t)i] > stamp[(size_t)best]))))
        best = i;
    }
    it++;
    if (gain[(size_t)best] <= 0) break;
    if (r[(size_t)best] == 0) {
      if (free >= R.mn(best)) {
        const float before = gain[(size_t)best];
        r[(size_t)best] = R.mn(best); free -= R.mn(best);
        gain[(size_t)best] = R.S(best, r[(size_t)best] + 1) - R.S(best, r[(size_t)best]);
        stamp[(size_t)best] = it;
        // MinNumProc == 0: nothing changed but (possibly) the gain; if that is unchanged too the
        // reference repeats this pick forever (guard, same rule as the kernel's)
        if (R.mn(best) == 0 && sameBits(before, gain[(size_t)best]) && !(gain[(size_t)best] <= 0))
          return VODA_ST_NONTERMINATING;
      } else { live[(size_t)best] = 0; nlive--; }
    } else {
      r[(size_t)best] += 1; free -= 1;
      gain[(size_t)best] = R.S(best, r[(size_t)best] + 1) - R.S(best, r[(size_t)best]);
      stamp[(size_t)best] = it;
      if (r[(size_t)best] >= R.mx(best)) { live[(size_t)best] = 0; nlive--; }
    }
  }
  return validate(R, r);
}

// FfDLOptimizer ffdl_optimizer.go:32-130 with two rolling rows of P and a compact SOL table
int ffdl(const Round& R, std::vector<long>& r) {
  if (R.n == 0) return VODA_ST_OK;
  int st = VODA_ST_OK;
  const long K = R.total;
  if (K < 0) {
    st = VODA_ST_RUNTIME_PANIC;
  } else {
    auto ord = sortedBySubmit(R);
    const long J = std::min<long>(R.n, K);
    std::vector<float> prev((size_t)K + 1, 0.0f), cur((size_t)K + 1);
    std::vector<int32_t> SOL((size_t)J * ((size_t)K + 1), 0);
    std::vector<float> Srow;
    for (long j = 1; j <= J; j++) {
      const int job = ord[(size_t)j - 1];
      const long mx = R.mx(job);
      const long gmax = std::max<long>(0, std::min<long>(mx, K));
      Srow.assign((size_t)gmax + 1, 0.0f);
      for (long g = 1; g <= gmax; g++) Srow[(size_t)g] = R.S(job, g);
      cur[0] = -10000.0f;
      int32_t* sol = &SOL[(size_t)(j - 1) * ((size_t)K + 1)];
      for (long k = 1; k <= K; k++) {
        float best = -10000.0f;
        int32_t bg = 0;
        const long lim = std::min<long>(gmax, k);
        for (long g = 1; g <= lim; g++) {
          float p = Srow[(size_t)g] + prev[(size_t)(k - g)];
          if (p > best) { best = p; bg = (int32_t)g; }
        }
        cur[(size_t)k] = best;
        sol[k] = bg;
      }
      std::swap(prev, cur);
    }
    if (prev[(size_t)K] <= 0) {
      st = VODA_ST_INFEASIBLE;
    } else {
      long k = K;
      for (long j = J; j > 0; j--) {
        long g = SOL[(size_t)(j - 1) * ((size_t)K + 1) + (size_t)k];
        r[(size_t)ord[(size_t)j - 1]] = g;
        k -= g;
      }
    }
  }
  int v = validate(R, r);  // deferred validateResult replaces a panic in flight
  return v ? v : st;
}

// AFS-L afsl.go:31-106 (straight left fold; no shortcuts)
int afsl(const Round& R, std::vector<long>& r) {
  auto live = sortedBySubmit(R);
  long free = R.total;
  auto len = [&](int job, long w) -> double {
    if (w == 0) return std::numeric_limits<double>::infinity();
    return (double)(R.rem(job) / R.S(job, w));
  };
  while (free > 0 && !live.empty()) {
    int j = live[0];
    size_t pos = 0;
    for (size_t i = 1; i < live.size(); i++) {
      int jb = live[i];
      int win;
      if (r[(size_t)j] == 0 && r[(size_t)jb] == 0) {
        win = (R.rem(j) >= R.rem(jb)) ? jb : j;
      } else {
        double lenA = len(j, r[(size_t)j]), lenB = len(jb, r[(size_t)j]);
        int s = j, l = jb;
        if (lenA >= lenB) { s = jb; l = j; }
        float left = (R.S(l, r[(size_t)l] + 1) - R.S(l, r[(size_t)l])) / R.S(l, r[(size_t)l] + 1);
        float right = (R.S(s, r[(size_t)s] + 1) - R.S(s, r[(size_t)s])) / R.S(s, r[(size_t)s]);
        win = (left > right) ? l : s;
      }
      if (win != j) { j = win; pos = i; }
    }
    r[(size_t)j] += 1;
    free -= 1;
    if (r[(size_t)j] >= R.mx(j)) live.erase(live.begin() + (long)pos);
  }
  return validate(R, r);
}

void run_rounds(const voda_alloc_batch* b, int32_t* out_gpus, int32_t* out_status, int32_t r0,
                int32_t r1) {
  const long Q = b->queue_num > 0 ? b->queue_num : 2;
  const long compaction = b->compaction_threshold >= 0 ? b->compaction_threshold : 10;
  std::vector<long> r;
  for (int32_t rd = r0; rd < r1; rd++) {
    Round R{b, b->job_off[rd], b->job_off[rd + 1] - b->job_off[rd], b->total_gpu[rd], b->speedup_stride};
    r.assign((size_t)R.n, 0);
    int st;
    switch (b->algo) {
      case VODA_ALGO_FIFO: st = simple(R, r, false); break;
      case VODA_ALGO_SRJF: st = simple(R, r, true); break;
      case VODA_ALGO_ELASTIC_FIFO: st = elasticFill(R, r, false); break;
      case VODA_ALGO_ELASTIC_SRJF: st = elasticFill(R, r, true); break;
      case VODA_ALGO_TIRESIAS: st = tiresias(R, r, Q); break;
      case VODA_ALGO_ELASTIC_TIRESIAS: st = elasticTiresias(R, r, Q, compaction); break;
      case VODA_ALGO_FFDL_OPTIMIZER: st = ffdl(R, r); break;
      case VODA_ALGO_AFSL: st = afsl(R, r); break;
      default: st = VODA_ST_NOT_FOUND;
    }
    out_status[rd] = st;
    for (int i = 0; i < R.n; i++) {
      long v = r[(size_t)i];
      if (v > INT32_MAX) v = INT32_MAX;
      if (v < INT32_MIN) v = INT32_MIN;
      out_gpus[R.o + i] = (int32_t)v;
    }
  }
}

}  // namespace

extern "C" int voda_oracle_allocate_dense(const voda_alloc_batch* b, int32_t* out_gpus,
                                          int32_t* out_status, int32_t threads) {
  if (!b || !b->job_off || !b->total_gpu || b->rounds < 0) return VODA_E_INVALID;
  if (b->algo < 0 || b->algo >= VODA_ALGO_COUNT) return VODA_E_INVALID;
  if (threads <= 1 || b->rounds < 2) {
    run_rounds(b, out_gpus, out_status, 0, b->rounds);
    return VODA_OK;
  }
  std::vector<std::thread> pool;
  for (int32_t t = 0; t < threads; t++) {
    int32_t r0 = (int32_t)((int64_t)b->rounds * t / threads);
    int32_t r1 = (int32_t)((int64_t)b->rounds * (t + 1) / threads);
    pool.emplace_back(run_rounds, b, out_gpus, out_status, r0, r1);
  }
  for (auto& th : pool) th.join();
  return VODA_OK;
}
