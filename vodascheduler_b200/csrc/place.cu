// place.cu — the arithmetic core of PlacementManager.Place for a batch of rounds (sm_100a):
//   bestFit            pkg/placement/placement_manager.go:415-487
//   scoreCandidates    :525-544   M[logical r][physical c] = sum over jobs of r of
//                                 (cand_c[job] <= pos_r[job] ? cand_c[job] : pos_r[job])
//   ComputeMunkresMax  :505-507   (munkres.cuh)
//   rename + score     :509-512
// One CTA per round.  Warp 0 runs bestFit (a serial walk over the sorted requests; the two
// node searches per step are warp reductions over (freeSlots, index) keys that keep the
// reference's first-index tie-breaks), all threads build the n x n score matrix with atomics
// (a join of the new layout and the previous one on the job id), then warp 0 solves the
// assignment in place.  The matrix lives in an L2-resident per-CTA slab of global scratch, not
// in shared memory: bestFit and the solve are single-warp dependent chains, so throughput comes
// from resident rounds per SM, and 64 KB of shared memory per round would allow only two.
// Reference quirks kept: the last chunk of a split job records the WHOLE request (Q11, :479)
// so freeSlots can go negative, and the loop that never ends once a request meets a cluster
// without usable slots (Q12) is cut with VODA_ST_NONTERMINATING.
// Large batches without a pod table run in a split form instead: place_kernel<true> stops after the
// score matrix (a slab per round) and place_solve_kernel solves with one warp per round, which doubles
// the rounds an SM works on during the assignment (see launch_place).
#include "kernels.h"
#include "munkres.cuh"

namespace voda {
namespace {

constexpr int kThreads = 64;
constexpr int kMaxNodes = 128;
constexpr int kLimit = 1 << 22;  // |freeSlots| * 128 must fit an int32 key

struct PlaceSmem {
  int32_t* M;        // [n*n] scores, then rowmax - score (global scratch slab of this CTA)
  int64_t* khi;      // [Mp] sort key (-workers)
  int32_t* klo;      // [Mp] zeros
  int* perm;         // [Mp]
  int32_t* workers;  // [Mp]
  int* jobHead;      // [Mp] first layout entry of a job, -1 if none
  int32_t* eNode;    // [Mp + 128]
  int32_t* eJob;
  int32_t* eW;
  int* eNext;
  int32_t* rowmax;   // [128]
  long long* red;    // [4]
  int* range;        // [2 per warp] min and max of the score matrix
  MunkresSmem mk;
};

// Warp 0 of the CTA: rowmax - M as CellT cells written over M's own slab (row i of the narrow matrix
// lies inside the rows of M already consumed), the zero masks, the solve, and the total score of
// the assignment.  Returns the score (identical in every lane).
template <typename CellT>
__device__ long long place_solve(const PlaceSmem& sm, int n, int lane) {
  const unsigned full = 0xffffffffu;
  CellT* C = reinterpret_cast<CellT*>(sm.M);
  for (int x = lane; x < 128 * 4; x += 32) {
    sm.mk.zt[x] = 0;
    sm.mk.zr[x] = 0;
  }
  __syncwarp();
  bool wide = false;  // some cell of rowmax - M is too large for 32-bit potentials
  uint32_t acc[4] = {0, 0, 0, 0};  // zeros of the current block of 32 rows in columns lane + 32 t
  // Rows mixing negative and non-negative scores (a previous layout with negative counts): the
  // reference's MaxInt64 - score wraps there (munkres.cu).  With k = the smallest |score| among the
  // row's negative ones, what its step 1 leaves is |score| - k for the negative scores and
  // -(score + k) for the others: in both cases (-k) - score, so rowmax[i] keeps -k for such a row and
  // the total score is recovered from the cells the same way as for the others.
  for (int i = 0; i < n; i++) {
    int v[4];
    int mx = (int)0x80000000, negMax = (int)0x80000000;
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const int col = lane + 32 * t;
      v[t] = col < n ? sm.M[i * n + col] : (int)0x80000000;
      mx = max(mx, v[t]);
      if (sizeof(CellT) == 4 && col < n && v[t] < 0) negMax = max(negMax, v[t]);
    }
    mx = __reduce_max_sync(full, mx);
    bool mixed = false;
    int k = 0;
    if (sizeof(CellT) == 4 && mx >= 0) {  // (a row with a non-negative maximum may also hold negative scores)
      negMax = __reduce_max_sync(full, negMax);
      mixed = negMax != (int)0x80000000;
      k = -negMax;
    }
    if (lane == 0) sm.rowmax[i] = mixed ? -k : mx;
    const uint32_t rowBit = 1u << (i & 31);
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const int col = lane + 32 * t;
      const bool ok = col < n;
      int c = ok ? mx - v[t] : 1;
      if (mixed && ok) c = -k - v[t];  // = |score| - k for a negative score, -(score + k) otherwise
      wide |= ok && (unsigned)c >= (1u << 23);
      if (ok) C[i * n + col] = (CellT)c;
      const bool z = ok && c == 0;
      const unsigned bz = __ballot_sync(full, z);
      if (lane == 0) sm.mk.zr[i * 4 + t] = bz;
      acc[t] |= z ? rowBit : 0u;  // this lane owns column col
    }
    if ((i & 31) == 31 || i == n - 1) {  // the block of 32 rows is complete
#pragma unroll
      for (int t = 0; t < 4; t++) {
        if (lane + 32 * t < n) sm.mk.zt[(lane + 32 * t) * 4 + (i >> 5)] = acc[t];
        acc[t] = 0;
      }
    }
  }
  __syncwarp();
  // the potentials are bounded by n * max cell: 32 bits unless a cell reaches 2^23 (scores are sums
  // of workers on one node: only a degenerate input gets there)
  // (negative cells count as wide: the 64-bit form of step 6 rebuilds its zero masks from the values)
  if (sizeof(CellT) == 4 && __any_sync(full, wide)) munkres_solve<CellT, int64_t>(C, n, sm.mk);
  else munkres_solve<CellT, int32_t>(C, n, sm.mk);
  long long sc = 0;
  for (int i = lane; i < n; i += 32) sc += (long long)sm.rowmax[i] - (long long)C[i * n + sm.mk.starCol[i]];
  for (int d = 16; d > 0; d >>= 1) sc += __shfl_xor_sync(full, sc, d);
  return sc;
}

#ifndef VODA_PLACE_MINBLOCKS
#define VODA_PLACE_MINBLOCKS 10  // 96 registers: 10 CTAs per SM (118 registers unbounded hold 8: 12.7 vs 9.65 ms per 20k rounds; 12 and 16 spill and are slower)
#endif
constexpr size_t kMunkresSmemBytes = 16 + 2 * 4 * 4 * 128 + 3 * 2 * 128 + 2 * 8 * 128;
// The solver's shared-memory state, from the next 16-byte boundary at or after p.
__device__ __forceinline__ void carve_munkres(MunkresSmem& mk, unsigned char* p) {
  p = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(p) + 15) & ~(uintptr_t)15);
  mk.zr = reinterpret_cast<uint32_t*>(p); p += 4 * 4 * 128;  // read as uint4
  mk.zt = reinterpret_cast<uint32_t*>(p); p += 4 * 4 * 128;
  mk.cs = p; p += 8 * 128;                                   // read as int4
  mk.ra = p; p += 8 * 128;
  mk.starCol = reinterpret_cast<short*>(p); p += 2 * 128;    // read as short4
  mk.starRow = reinterpret_cast<short*>(p); p += 2 * 128;
  mk.primeCol = p; p += 2 * 128;
}

// SPLIT: the kernel stops after the score matrix (one slab per ROUND of [r0, r1)) and leaves the
// assignment to place_solve_kernel, which runs a warp per round at twice the residency; the fused form
// (one slab per CTA) serves small batches, where a second launch would cost more than it gains, and
// rounds with a pod table, whose migration counts need the layout this kernel still holds.
constexpr int kPending = -2;  // status of a round whose assignment place_solve_kernel has yet to compute
template <bool SPLIT>
__global__ void __launch_bounds__(kThreads, SPLIT ? 1 : VODA_PLACE_MINBLOCKS)
    place_kernel(PlaceArgs a, int Mp, int32_t* slabs, size_t slabInts, int r0, int r1) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PlaceSmem sm;
  {
    unsigned char* p = smem_raw;
    sm.M = slabs + (size_t)blockIdx.x * slabInts;
    sm.khi = reinterpret_cast<int64_t*>(p); p += 8 * (size_t)Mp;
    sm.red = reinterpret_cast<long long*>(p); p += 8 * 4;
    sm.klo = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Mp;
    sm.perm = reinterpret_cast<int*>(p); p += 4 * (size_t)Mp;
    sm.workers = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)Mp;
    sm.jobHead = reinterpret_cast<int*>(p); p += 4 * (size_t)Mp;
    sm.eNode = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)(Mp + kMaxNodes);
    sm.eJob = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)(Mp + kMaxNodes);
    sm.eW = reinterpret_cast<int32_t*>(p); p += 4 * (size_t)(Mp + kMaxNodes);
    sm.eNext = reinterpret_cast<int*>(p); p += 4 * (size_t)(Mp + kMaxNodes);
    sm.rowmax = reinterpret_cast<int32_t*>(p); p += 4 * 128;
    sm.range = reinterpret_cast<int*>(p); p += 4 * 2 * (kThreads / 32);
    if (!SPLIT) carve_munkres(sm.mk, p);  // last: the split form does not allocate it
  }
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned full = 0xffffffffu;

  for (int rd = r0 + blockIdx.x; rd < r1; rd += gridDim.x) {
    if (SPLIT) sm.M = slabs + (size_t)(rd - r0) * slabInts;
    const int n0 = a.node_off[rd], n = a.node_off[rd + 1] - n0;
    const int q0 = a.req_off[rd], m = a.req_off[rd + 1] - q0;
    const int e0 = q0 + n0;
    __syncthreads();
    if (tid == 0) {
      sm.red[0] = VODA_ST_OK;  // status
      sm.red[1] = 0;           // entries
      sm.red[2] = 0;           // cross
      sm.red[3] = 0;           // two 32-bit counters: migrated workers, migrated launchers
      if (a.migrated_workers) {
        a.migrated_workers[rd] = 0;
        a.migrated_launchers[rd] = 0;
      }
    }
    if (n <= 0) {  // no nodes: bestFit returns at once, size-0 matrix
      if (tid == 0) {
        a.lay_cnt[rd] = 0; a.score[rd] = 0; a.cross[rd] = 0; a.status[rd] = VODA_ST_OK;
      }
      continue;
    }
    if (n > a.max_nodes || n > kMaxNodes || m > a.max_requests) {
      if (tid == 0) {
        a.lay_cnt[rd] = 0; a.score[rd] = 0; a.cross[rd] = 0; a.status[rd] = VODA_ST_BOUND;
      }
      continue;
    }
    // requests sorted by numWorkers descending, stable (:422-424); the len(jobRequests)
    // zero-valued leading entries of the reference (:416, Q10) never enter the loop body
    int tooBig = 0;
    for (int i = tid; i < m; i += kThreads) {
      const int w = a.req_workers[q0 + i];
      sm.workers[i] = w;
      sm.khi[i] = -(int64_t)w;
      sm.klo[i] = 0;
      sm.jobHead[i] = -1;
      tooBig |= (w >= kLimit);
    }
    for (int i = tid; i < n; i += kThreads) {
      const int s = a.node_slots[n0 + i];
      tooBig |= (s >= kLimit || s <= -kLimit);
    }
    for (int x = tid; x < n * n; x += kThreads) sm.M[x] = 0;
    if (__syncthreads_or(tooBig)) {
      if (tid == 0) {
        a.lay_cnt[rd] = 0; a.score[rd] = 0; a.cross[rd] = 0; a.status[rd] = VODA_ST_BOUND;
      }
      continue;
    }
    block_argsort(sm.khi, sm.klo, sm.perm, m, next_pow2(m > 0 ? m : 1));

    // ---- bestFit (warp 0) -------------------------------------------------------------
    if (warp == 0) {
      int fr[4];
      bool valid[4];
      long long total = 0;
#pragma unroll
      for (int q = 0; q < 4; q++) {
        const int i = lane + 32 * q;
        valid[q] = i < n;
        fr[q] = valid[q] ? a.node_slots[n0 + i] : 0;  // NewNodeState("TBD", totalSlots) (:317-320)
        total += fr[q];
      }
      for (int d = 16; d > 0; d >>= 1) total += __shfl_xor_sync(full, total, d);
      int cnt = 0, cross = 0, status = VODA_ST_OK;
      bool stop = false;
      for (int t = 0; t < m && !stop; t++) {
        const int job = sm.perm[t];
        const int w = sm.workers[job];
        int requested = w;
        int tail = -1;
        while (requested > 0) {
          if (total == 0) {  // :455-457 return
            stop = true;
            break;
          }
          int keyb = 0x7fffffff, keym = (int)0x80000000;
#pragma unroll
          for (int q = 0; q < 4; q++) {
            if (valid[q]) {
              const int i = lane + 32 * q;
              if (fr[q] >= requested) keyb = min(keyb, fr[q] * 128 + i);
              keym = max(keym, fr[q] * 128 + (127 - i));
            }
          }
          const int kb = __reduce_min_sync(full, keyb);
          int node, give;
          if (kb == 0x7fffffff) {  // no node fits: take everything from the fullest one (:472-477)
            const int km = __reduce_max_sync(full, keym);
            node = 127 - (km & 127);
            give = km >> 7;
            if (give == 0) {  // the state can never change again (Q12)
              status = VODA_ST_NONTERMINATING;
              stop = true;
              break;
            }
            requested -= give;
            total -= give;
#pragma unroll
            for (int q = 0; q < 4; q++)
              if (lane + 32 * q == node) fr[q] = 0;
            cross++;
          } else {  // best fit; records and subtracts the WHOLE request (:479-482)
            node = kb & 127;
            give = w;
            requested -= w;
            total -= w;
#pragma unroll
            for (int q = 0; q < 4; q++)
              if (lane + 32 * q == node) fr[q] -= w;
          }
          if (lane == 0) {
            sm.eNode[cnt] = node;
            sm.eJob[cnt] = job;
            sm.eW[cnt] = give;
            sm.eNext[cnt] = -1;
            if (tail < 0) sm.jobHead[job] = cnt; else sm.eNext[tail] = cnt;
          }
          tail = cnt;
          cnt++;
        }
      }
#pragma unroll
      for (int q = 0; q < 4; q++)
        if (valid[q]) a.free_slots[n0 + lane + 32 * q] = fr[q];
      if (lane == 0) {
        sm.red[0] = status;
        sm.red[1] = cnt;
        sm.red[2] = cross;
      }
    }
    __syncthreads();
    const int status = (int)sm.red[0];
    const int cnt = (int)sm.red[1];
    if (status != VODA_ST_OK) {
      if (tid == 0) {
        a.lay_cnt[rd] = 0; a.score[rd] = 0; a.cross[rd] = 0; a.status[rd] = status;
      }
      continue;
    }
    for (int e = tid; e < cnt; e += kThreads) {
      a.lay_node[e0 + e] = sm.eNode[e];
      a.lay_job[e0 + e] = sm.eJob[e];
      a.lay_workers[e0 + e] = sm.eW[e];
    }
    // ---- score matrix: join previous (physical) entries with the new layout on the job ----
    const int p0 = a.prev_off[n0], p1 = a.prev_off[n0 + n];
    for (int pe = p0 + tid; pe < p1; pe += kThreads) {
      const int job = a.prev_job[pe];
      if (job < 0 || job >= m) continue;  // a job that is not requested cannot be in the new layout
      const int cw = a.prev_workers[pe];
      // physical node of this entry: the node whose [prev_off) range holds pe
      int lo = 0, hi = n - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (a.prev_off[n0 + mid] <= pe) lo = mid; else hi = mid - 1;
      }
      const int c = lo;
      for (int e = sm.jobHead[job]; e >= 0; e = sm.eNext[e]) {
        const int w = sm.eW[e];
        if (w < 0) continue;  // handled below for every column
        atomicAdd(&sm.M[sm.eNode[e] * n + c], cw <= w ? cw : w);
      }
    }
    // rare: a layout entry with a negative count (reference quirk) contributes to EVERY
    // column, because an absent candidate reads 0 and 0 <= negative is false (:537-541)
    for (int e = tid; e < cnt; e += kThreads) {
      const int w = sm.eW[e];
      if (w >= 0) continue;
      const int job = sm.eJob[e];
      for (int c = 0; c < n; c++) {
        int cand = 0;
        for (int pe = a.prev_off[n0 + c]; pe < a.prev_off[n0 + c + 1]; pe++)
          if (a.prev_job[pe] == job) cand = a.prev_workers[pe];
        atomicAdd(&sm.M[sm.eNode[e] * n + c], cand <= w ? cand : w);
      }
    }
    __syncthreads();
    if (SPLIT) {  // the matrix is complete (barrier above): the rest of the round is place_solve_kernel's
      if (tid == 0) {
        a.lay_cnt[rd] = cnt;
        a.cross[rd] = (int)sm.red[2];
        a.status[rd] = kPending;
      }
      continue;
    }
    // ---- ComputeMunkresMax on M (warp 0), then rename + total score ------------------------
    // The cells of rowmax - M are bounded by max(M) - min(M): found by all threads, it picks the
    // narrowest cell type, and the solve then works on a matrix a quarter (or half) the size of M,
    // written over M's own slab.
    {
      int lo = 0x7fffffff, hi = (int)0x80000000;
      for (int x = tid; x < n * n; x += kThreads) {
        const int v = sm.M[x];
        lo = min(lo, v);
        hi = max(hi, v);
      }
      lo = __reduce_min_sync(full, lo);
      hi = __reduce_max_sync(full, hi);
      if (lane == 0) {
        sm.range[2 * warp] = lo;
        sm.range[2 * warp + 1] = hi;
      }
      __syncthreads();
    }
    if (warp == 0) {
      int lo = sm.range[0], hi = sm.range[1];
#pragma unroll
      for (int w = 1; w < kThreads / 32; w++) {
        lo = min(lo, sm.range[2 * w]);
        hi = max(hi, sm.range[2 * w + 1]);
      }
      const long long span = (long long)hi - (long long)lo;
      long long sc;
      if (lo >= 0 && span < 256) sc = place_solve<uint8_t>(sm, n, lane);
      else if (lo >= 0 && span < 65536) sc = place_solve<uint16_t>(sm, n, lane);
      else sc = place_solve<int32_t>(sm, n, lane);  // also every matrix with a negative score (signed cells)
      for (int i = lane; i < n; i += 32) a.phys_of_logical[n0 + i] = sm.mk.starCol[i];
      if (lane == 0) {
        a.score[rd] = sc;
        a.lay_cnt[rd] = cnt;
        a.cross[rd] = (int)sm.red[2];
        a.status[rd] = VODA_ST_OK;
      }
    }
    // ---- updateJobStates + updatePodNodeName (:548-617) as counts: which worker pods change node ----
    // A job's new nodeNumSlotsList = its layout entries renamed to their physical nodes, taken by
    // physical index; its old one is the caller's run-length pod table.  Pod i of the job sits on
    // the i-th slot of either list; a pod present in both whose node differs is deleted (:589-596),
    // and a job all of whose workers are deleted loses its launcher too (:606-609).
    if (a.pods_off) {
      __syncthreads();  // the binding (starCol) is final
      int* mig = reinterpret_cast<int*>(&sm.red[3]);
      for (int job = tid; job < m; job += kThreads) {
        if (sm.jobHead[job] < 0) continue;  // no entry: the job is not in the new jobStates
        int run = a.pods_off[q0 + job];
        const int runEnd = a.pods_off[q0 + job + 1];
        int left = run < runEnd ? a.pods_slots[run] : 0;
        int last = -1, deleted = 0, numWorkers = 0;
        for (;;) {
          int best = 0x7fffffff, bw = 0;
          for (int e = sm.jobHead[job]; e >= 0; e = sm.eNext[e]) {
            const int p = sm.mk.starCol[sm.eNode[e]];
            if (p > last && p <= best) {  // '<=': a later entry on the same node overwrites (map assignment)
              best = p;
              bw = sm.eW[e];
            }
          }
          if (best == 0x7fffffff) break;
          last = best;
          numWorkers += bw;
          int w = bw;
          while (w > 0) {
            while (run < runEnd && left <= 0) {
              run++;
              left = run < runEnd ? a.pods_slots[run] : 0;
            }
            if (run >= runEnd) break;  // the remaining pods are new: nothing to delete
            const int take = w < left ? w : left;
            if (a.pods_node[run] != best) deleted += take;
            w -= take;
            left -= take;
          }
        }
        if (deleted) atomicAdd(&mig[0], deleted);
        if (deleted == numWorkers) atomicAdd(&mig[1], 1);
      }
      __syncthreads();
      if (tid == 0) {
        a.migrated_workers[rd] = mig[0];
        a.migrated_launchers[rd] = mig[1];
      }
    }
  }
}

// The assignment half of the split form: ONE WARP per round (a one-warp CTA, so that an SM holds twice as
// many rounds as the fused kernel does).  Reads the score matrix place_kernel<true> left in the round's
// slab and writes the binding, the total score and the final status.
constexpr size_t kSolveSmemBytes = kMunkresSmemBytes + 4 * 128;
#ifndef VODA_SOLVE_MINBLOCKS
#define VODA_SOLVE_MINBLOCKS 20  // 96 registers: 20 one-warp CTAs per SM (16 / 20 / 24 / 32: 6.89 / 6.47 / 6.46 / 6.50 ms per 20k rounds)
#endif
__global__ void __launch_bounds__(32, VODA_SOLVE_MINBLOCKS) place_solve_kernel(PlaceArgs a, int32_t* slabs, size_t slabInts, int r0, int r1) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PlaceSmem sm{};
  sm.rowmax = reinterpret_cast<int32_t*>(smem_raw);
  carve_munkres(sm.mk, smem_raw + 4 * 128);
  const int lane = threadIdx.x;
  const unsigned full = 0xffffffffu;
  // one CTA per round, not a persistent grid: rounds differ in length by a factor of several, and the
  // hardware hands the next round to whichever SM has room
  for (int rd = r0 + blockIdx.x; rd < r1; rd += gridDim.x) {
    if (a.status[rd] != kPending) continue;  // refused or empty: place_kernel has written the outcome
    const int n0 = a.node_off[rd], n = a.node_off[rd + 1] - n0;
    sm.M = slabs + (size_t)(rd - r0) * slabInts;
    int lo = 0x7fffffff, hi = (int)0x80000000;
    for (int x = lane; x < n * n; x += 32) {
      const int v = sm.M[x];
      lo = min(lo, v);
      hi = max(hi, v);
    }
    lo = __reduce_min_sync(full, lo);
    hi = __reduce_max_sync(full, hi);
    const long long span = (long long)hi - (long long)lo;
    long long sc;
    if (lo >= 0 && span < 256) sc = place_solve<uint8_t>(sm, n, lane);
    else if (lo >= 0 && span < 65536) sc = place_solve<uint16_t>(sm, n, lane);
    else sc = place_solve<int32_t>(sm, n, lane);  // also every matrix with a negative score (signed cells)
    for (int i = lane; i < n; i += 32) a.phys_of_logical[n0 + i] = sm.mk.starCol[i];
    if (lane == 0) {
      a.score[rd] = sc;
      a.status[rd] = VODA_ST_OK;
    }
    __syncwarp();
  }
}

size_t place_smem(int Mp, bool split = false) {
  return 16 + (size_t)Mp * (8 + 4 * 4) + 8 * 4 + 4 * 4 * (size_t)(Mp + kMaxNodes) + 4 * 128 + 4 * 2 * (kThreads / 32) + 64 +
         (split ? 0 : kMunkresSmemBytes);
}

}  // namespace

size_t place_smem_bytes(const PlaceArgs& a) { return place_smem(next_pow2(a.max_requests > 0 ? a.max_requests : 1)); }

static size_t place_slab_ints(const PlaceArgs& a) { return (((size_t)a.max_nodes * a.max_nodes) + 63) & ~(size_t)63; }

template <bool SPLIT>
static int place_grid(const PlaceArgs& a, int rounds, int sms, size_t smem, cudaError_t* err) {
  int perSm = 0;
  *err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, place_kernel<SPLIT>, kThreads, smem);
  if (perSm < 1) perSm = 1;
  if (const char* v = std::getenv("VODA_PLACE_CTAS")) perSm = std::max(1, std::min(perSm, std::atoi(v)));
  int grid = sms * perSm;
  if (rounds < grid) grid = rounds;
  return grid;
}

// The split form pays a second launch and a slab per round: it is used for batches of at least a few
// waves of the fused kernel, without a pod table (the migration counts need the fused kernel's state).
static bool place_split(const PlaceArgs& a, int rounds, int sms) {
  if (const char* v = std::getenv("VODA_PLACE_SPLIT")) return std::atoi(v) != 0 && !a.pods_off;
  return !a.pods_off && rounds >= 2 * 10 * sms;
}
// rounds of the split form processed per pair of launches: their slabs may take 2 GiB; equal sub-batches
static int place_split_rounds(const PlaceArgs& a, int rounds) {
  const size_t per = place_slab_ints(a) * 4;
  const size_t cap = std::max<size_t>(1, ((size_t)2 << 30) / per);
  const size_t parts = ((size_t)rounds + cap - 1) / cap;
  return (int)(((size_t)rounds + parts - 1) / parts);
}

// bytes of global scratch the launch needs (score-matrix slabs: one per CTA, or one per round of a sub-batch)
size_t place_scratch_bytes(const PlaceArgs& a, int sms, int rounds) {
  if (rounds < 0) rounds = a.rounds;
  const int Mp = next_pow2(a.max_requests > 0 ? a.max_requests : 1);
  // + the rows the last slab's vector loads may touch beyond its matrix (munkres.cuh)
  const size_t slack = (size_t)kMunkresSlackRows * kMaxNodes * 4 + 256;
  if (place_split(a, rounds, sms)) return (size_t)place_split_rounds(a, rounds) * place_slab_ints(a) * 4 + slack;
  const size_t smem = place_smem(Mp);
  if (smem > 48 * 1024) cudaFuncSetAttribute(place_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaError_t e;
  const int grid = place_grid<false>(a, rounds, sms, smem, &e);
  return (size_t)grid * place_slab_ints(a) * 4 + slack;
}

// Enqueues the placement of rounds [first, last) (last < 0: all of them); *launched = kernels launched.
cudaError_t launch_place(const PlaceArgs& a, int sms, void* scratch, cudaStream_t stream, int* launched, int first,
                         int last) {
  if (last < 0) last = a.rounds;
  const int rounds = last - first;
  const int Mp = next_pow2(a.max_requests > 0 ? a.max_requests : 1);
  int32_t* slabs = static_cast<int32_t*>(scratch);
  cudaError_t e;
  *launched = 0;
  if (!place_split(a, rounds, sms)) {
    const size_t smem = place_smem(Mp);
    if (smem > 48 * 1024) {
      e = cudaFuncSetAttribute(place_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    const int grid = place_grid<false>(a, rounds, sms, smem, &e);
    if (e != cudaSuccess) return e;
    place_kernel<false><<<grid, kThreads, smem, stream>>>(a, Mp, slabs, place_slab_ints(a), first, last);
    *launched = 1;
    return cudaGetLastError();
  }
  const size_t smem = place_smem(Mp, true);
  if (smem > 48 * 1024) {
    e = cudaFuncSetAttribute(place_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const int sub = place_split_rounds(a, rounds);
  for (int r0 = first; r0 < last; r0 += sub) {
    const int r1 = std::min(last, r0 + sub);
    const int grid = place_grid<true>(a, r1 - r0, sms, smem, &e);
    if (e != cudaSuccess) return e;
    place_kernel<true><<<grid, kThreads, smem, stream>>>(a, Mp, slabs, place_slab_ints(a), r0, r1);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    place_solve_kernel<<<r1 - r0, 32, kSolveSmemBytes, stream>>>(a, slabs, place_slab_ints(a), r0, r1);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    *launched += 2;
  }
  return cudaSuccess;
}

}  // namespace voda
