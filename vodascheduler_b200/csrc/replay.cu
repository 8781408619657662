// replay.cu — voda_replay: many what-if replicas of the scheduler's round loop (BASELINE configs[4]).
//
// Per tick, the rounds of all replicas that need one are evaluated on the GPU as ONE ElasticSRJF
// batch, one compareResults batch and ONE placement batch (the device-pointer entry points, on the
// context's stream, fed by two packed copies each way per tick).  What runs on the host here is the reference's state bookkeeping around those
// kernels — releaseSlots (pkg/placement/placement_manager.go:337-411), updateJobStates (:548-566),
// updatePodNodeName (:571-617) — map manipulation in the reference, vectors here, replicas spread
// over host threads.  No allocation or placement decision is computed on the CPU.
//
// Iteration orders the reference leaves to Go's map randomisation are fixed as documented in
// include/voda_accel.h (nodes by physical index, requests = live jobs in arrival order, a node's
// jobs in bestFit's creation order, a job's nodes by physical index).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/voda_accel.h"
#include "hostpack.h"

namespace {

using voda::HostPool;

struct NodeSt {
  int total = 0, free = 0;
  std::vector<std::pair<int, int>> jw;  // node.jobNumWorkers: (job, workers), insertion order
  int* find(int job) {
    for (auto& e : jw)
      if (e.first == job) return &e.second;
    return nullptr;
  }
  void erase(int job) {
    for (size_t i = 0; i < jw.size(); i++)
      if (jw[i].first == job) {
        jw.erase(jw.begin() + (long)i);
        return;
      }
  }
};

struct JobSt {
  int numWorkers = 0;
  std::vector<std::pair<int, int>> ns;  // nodeNumSlotsList: (node, numSlots)
};

inline void fnv(uint64_t& h, int64_t v) { h = (h ^ (uint64_t)v) * 0x100000001B3ull; }

struct Replica {
  int j0 = 0, J = 0, nodes = 0, slots = 0;
  double rate = 30.0, now = 0.0;
  int next = 0;
  bool dirty = false;
  std::vector<int> live;          // job ids (replica-local), arrival order
  std::vector<double> remaining;  // [J]
  std::vector<int> alloc;         // [J] GPUs held
  std::vector<NodeSt> nodeSt;     // by physical index
  std::vector<JobSt> jobSt;       // [J]; jobs listed in `placed` are the keys of pm.jobStates
  std::vector<int> placed;
  std::vector<std::vector<int>> podNode;  // [J] node of worker pod idx (pm.podNodeName)
  std::vector<int> hasPods;               // jobs with a non-empty podNode entry
  int64_t done = 0, rounds = 0, migW = 0, migL = 0, skipped = 0, act[4] = {0, 0, 0, 0};
  double jct = 0.0;
  uint64_t hist = 0xCBF29CE484222325ull;
  // per-tick scratch of the placement round
  std::vector<int> prevOff, prevJob, prevW, liveIndex;

  bool finished() const { return done == J; }
};

// One rate-limit period of simulated time: progress, completions, arrivals (header: "Model").
void advance(Replica& r, const voda_replay_batch* b) {
  const double* arrival = b->arrival_sec + r.j0;
  const double* work = b->work_sec + r.j0;
  const int W = b->speedup_stride;
  double t1 = r.now + r.rate;
  if (r.live.empty() && r.next < r.J && arrival[r.next] > t1) t1 = arrival[r.next];  // idle: jump
  const double dt = t1 - r.now;
  size_t keep = 0;
  for (size_t i = 0; i < r.live.size(); i++) {
    const int j = r.live[i];
    const int w = r.alloc[j];
    if (w > 0) {
      const int g = w < W - 1 ? w : W - 1;
      r.remaining[j] -= dt * (double)b->speedup[(size_t)(r.j0 + j) * (size_t)W + (size_t)g];
    }
    if (r.remaining[j] <= 0.0) {
      r.done++;
      r.jct += t1 - arrival[j];
      r.alloc[j] = 0;
      r.dirty = true;  // job finished (scheduler.go:648)
    } else {
      r.live[keep++] = j;
    }
  }
  r.live.resize(keep);
  while (r.next < r.J && arrival[r.next] <= t1) {
    const int j = r.next++;
    r.live.push_back(j);
    r.remaining[j] = work[j];
    r.dirty = true;  // job created (scheduler.go:887)
  }
  r.now = t1;
}

// releaseSlots, placement_manager.go:337-411.  jobRequests = the live jobs with their new counts
// (liveIndex[j] >= 0 marks membership).
void releaseSlots(Replica& r) {
  for (int job : r.placed) {
    JobSt& js = r.jobSt[job];
    if (r.liveIndex[job] < 0) {
      // not scheduled any more: terminated, release everything (:344-360)
      for (auto& ns : js.ns) {
        NodeSt& node = r.nodeSt[ns.first];
        node.free += ns.second;
        node.erase(job);
      }
      js.ns.clear();
    } else if (r.alloc[job] < js.numWorkers) {
      // scaled down: release from the last allocated node backwards (:362-409)
      int toRelease = js.numWorkers - r.alloc[job];
      while (toRelease > 0 && !js.ns.empty()) {
        auto& last = js.ns.back();
        NodeSt& node = r.nodeSt[last.first];
        int* nw = node.find(job);
        if (last.second >= toRelease) {
          last.second -= toRelease;
          node.free += toRelease;
          if (nw) *nw -= toRelease;
          toRelease = 0;
        } else {
          toRelease -= last.second;
          last.second = 0;
          if (nw) {
            node.free += *nw;
            *nw = 0;
          }
        }
        const bool emptyHere = last.second == 0;
        if (nw && *nw == 0) node.erase(job);
        if (emptyHere) js.ns.pop_back();
      }
    }
  }
}

// The round's "current (post-release) contents of the physical nodes" as CSR over the node index.
void packPrev(Replica& r) {
  r.prevOff.assign(1, 0);
  r.prevJob.clear();
  r.prevW.clear();
  for (int c = 0; c < r.nodes; c++) {
    for (auto& e : r.nodeSt[c].jw) {
      const int q = r.liveIndex[e.first];
      if (q >= 0) {
        r.prevJob.push_back(q);
        r.prevW.push_back(e.second);
      }
    }
    r.prevOff.push_back((int)r.prevJob.size());
  }
}

// updateJobStates, placement_manager.go:548-566 (nodes by physical index, a node's jobs in
// insertion order).
void updateJobStates(Replica& r) {
  for (int job : r.placed) {
    r.jobSt[job].ns.clear();
    r.jobSt[job].numWorkers = 0;
  }
  r.placed.clear();
  for (int c = 0; c < r.nodes; c++)
    for (auto& e : r.nodeSt[c].jw) {
      JobSt& js = r.jobSt[e.first];
      if (js.ns.empty()) r.placed.push_back(e.first);
      js.ns.push_back({c, e.second});
      js.numWorkers += e.second;
    }
}

// updatePodNodeName, placement_manager.go:571-617: the new pod -> node table and the pods whose
// node changed (the migration cost of the round).
void updatePodNodeName(Replica& r, int64_t* workers, int64_t* launchers) {
  int64_t dw = 0, dl = 0;
  std::vector<std::vector<int>> fresh;
  fresh.reserve(r.placed.size());
  for (int job : r.placed) {
    const JobSt& js = r.jobSt[job];
    const std::vector<int>& old = r.podNode[job];
    std::vector<int> now;
    int deleted = 0;
    for (auto& ns : js.ns)
      for (int i = 0; i < ns.second; i++) {
        const size_t idx = now.size();
        if (idx < old.size() && old[idx] != ns.first) {
          deleted++;
          dw++;
        }
        now.push_back(ns.first);
      }
    if (deleted == js.numWorkers) dl++;  // every worker moved: the launcher goes too (:606-609)
    fresh.push_back(std::move(now));
  }
  for (int job : r.hasPods) r.podNode[job].clear();
  r.hasPods.clear();
  for (size_t i = 0; i < r.placed.size(); i++) {
    r.podNode[r.placed[i]] = std::move(fresh[i]);
    r.hasPods.push_back(r.placed[i]);
  }
  *workers = dw;
  *launchers = dl;
}

struct Pinned {  // page-locked staging arrays, grown on demand
  void* p = nullptr;
  size_t cap = 0;
  ~Pinned() { voda_host_free(p); }
  template <typename T>
  T* get(size_t n) {
    const size_t bytes = (n ? n : 1) * sizeof(T);
    if (bytes > cap) {
      voda_host_free(p);
      cap = bytes + bytes / 2 + 256;
      p = voda_host_alloc(cap);
      if (!p) cap = 0;
    }
    return static_cast<T*>(p);
  }
};

struct Device {  // device staging, grown on demand
  char* p = nullptr;
  size_t cap = 0;
  ~Device() { cudaFree(p); }
  char* get(size_t bytes, cudaStream_t s) {
    if (bytes > cap) {
      if (p) {
        cudaStreamSynchronize(s);
        cudaFree(p);
      }
      cap = bytes + bytes / 2 + 4096;
      if (cudaMalloc((void**)&p, cap) != cudaSuccess) {
        p = nullptr;
        cap = 0;
      }
    }
    return p;
  }
};

}  // namespace

extern "C" int voda_replay(voda_ctx* ctx, const voda_replay_batch* b, const voda_replay_out* out) {
  if (!ctx || !b || !out) return VODA_E_INVALID;
  if (b->replicas < 0 || !b->job_off || b->speedup_stride < 2) return VODA_E_INVALID;
  const int R = b->replicas;
  if (R == 0) return VODA_OK;
  if (!b->arrival_sec || !b->work_sec || !b->min_proc || !b->max_proc || !b->speedup || !b->nodes || !b->slots ||
      !b->rate_limit_sec || !out->rounds || !out->done || !out->sim_time_sec || !out->jct_sum_sec ||
      !out->migrated_workers || !out->migrated_launchers || !out->placements_skipped || !out->actions ||
      !out->history || !out->ticks)
    return VODA_E_INVALID;
  {  // makes the context's device current for the staging buffers allocated below
    const int rc0 = voda_ctx_sync(ctx);
    if (rc0) return rc0;
  }
  std::vector<Replica> reps((size_t)R);
  for (int k = 0; k < R; k++) {
    Replica& r = reps[(size_t)k];
    r.j0 = b->job_off[k];
    r.J = b->job_off[k + 1] - b->job_off[k];
    r.nodes = b->nodes[k];
    r.slots = b->slots[k];
    r.rate = b->rate_limit_sec[k];
    if (r.J < 0 || r.nodes < 1 || r.nodes > 128 || r.slots < 1 || !(r.rate > 0.0)) return VODA_E_INVALID;
    r.remaining.assign((size_t)r.J, 0.0);
    r.alloc.assign((size_t)r.J, 0);
    r.jobSt.resize((size_t)r.J);
    r.podNode.resize((size_t)r.J);
    r.liveIndex.assign((size_t)r.J, -1);
    r.nodeSt.resize((size_t)r.nodes);
    for (auto& n : r.nodeSt) n.total = n.free = r.slots;
  }
  int threads = b->host_threads > 0 ? b->host_threads : std::min(16, std::max(1, R / 4));
  threads = std::min(threads, R);
  HostPool pool(threads);
  auto parallel = [&](const std::vector<int>& todo, const std::function<void(int)>& f) {
    if (threads == 1 || todo.size() < 2) {
      for (size_t i = 0; i < todo.size(); i++) f((int)i);
      return;
    }
    pool.run([&](int w) {
      for (size_t i = (size_t)w; i < todo.size(); i += (size_t)threads) f((int)i);
    });
  };

  // Per tick two packed transfers each way: (A) the allocation rounds in, counts + statuses + actions
  // out; (B) the placement rounds in (their request list is the allocation's output, already on the
  // device), bindings + layouts out.  Kernels run through the device-pointer entry points on the
  // context's stream.
  Pinned hA, hAo, hB, hBo;
  Device dA, dAo, dB, dBo;
  cudaStream_t stream = static_cast<cudaStream_t>(voda_ctx_stream(ctx));
  auto a16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
  std::vector<int> todo;
  int64_t ticks = 0;
  int rc = VODA_OK;
  for (;;) {
    bool all = true;
    for (auto& r : reps) all = all && r.finished();
    if (all || (b->max_ticks > 0 && ticks >= b->max_ticks)) break;
    ticks++;
    todo.clear();
    for (int k = 0; k < R; k++) {
      Replica& r = reps[(size_t)k];
      if (r.finished()) continue;
      advance(r, b);
      if (r.dirty && !r.live.empty()) todo.push_back(k);
      r.dirty = false;
    }
    if (todo.empty()) continue;
    const int T = (int)todo.size();
    // ---- (A) one allocation batch: a round per replica (ElasticSRJF over its live jobs) ----------
    std::vector<int> joff((size_t)T + 1, 0), noff((size_t)T + 1, 0);
    int maxJobs = 1, maxTotal = 1, maxNodes = 1;
    for (int i = 0; i < T; i++) {
      const Replica& r = reps[(size_t)todo[(size_t)i]];
      joff[(size_t)i + 1] = joff[(size_t)i] + (int)r.live.size();
      noff[(size_t)i + 1] = noff[(size_t)i] + r.nodes;
      maxJobs = std::max(maxJobs, (int)r.live.size());
      maxTotal = std::max(maxTotal, r.nodes * r.slots);
      maxNodes = std::max(maxNodes, r.nodes);
    }
    const size_t NJ = (size_t)joff[(size_t)T], NN = (size_t)noff[(size_t)T];
    const size_t oOff = 0, oTot = oOff + a16(4 * ((size_t)T + 1)), oMin = oTot + a16(4 * (size_t)T), oMax = oMin + a16(4 * NJ),
                 oRem = oMax + a16(4 * NJ), oOld = oRem + a16(4 * NJ), oJst = oOld + a16(4 * NJ), szA = oJst + a16(4 * NJ);
    const size_t rG = 0, rSt = rG + a16(4 * NJ), rAct = rSt + a16(4 * (size_t)T), szAo = rAct + a16(4 * NJ);
    char* pa = hA.get<char>(szA);
    char* pao = hAo.get<char>(szAo);
    char* da = dA.get(szA, stream);
    char* dao = dAo.get(szAo, stream);
    if (!pa || !pao || !da || !dao) return VODA_E_NOMEM;
    int32_t* hOff = (int32_t*)(pa + oOff);
    int32_t* hTot = (int32_t*)(pa + oTot);
    int32_t* hMin = (int32_t*)(pa + oMin);
    int32_t* hMax = (int32_t*)(pa + oMax);
    float* hRem = (float*)(pa + oRem);
    int32_t* hOld = (int32_t*)(pa + oOld);
    int32_t* hJst = (int32_t*)(pa + oJst);
    const int32_t* hG = (const int32_t*)(pao + rG);
    const int32_t* hSt = (const int32_t*)(pao + rSt);
    const int32_t* hAct = (const int32_t*)(pao + rAct);
    std::memcpy(hOff, joff.data(), sizeof(int32_t) * ((size_t)T + 1));
    parallel(todo, [&](int i) {
      Replica& r = reps[(size_t)todo[(size_t)i]];
      hTot[i] = r.nodes * r.slots;
      const int base = joff[(size_t)i];
      for (size_t q = 0; q < r.live.size(); q++) {
        const int j = r.live[q];
        hMin[base + (int)q] = b->min_proc[r.j0 + j];
        hMax[base + (int)q] = b->max_proc[r.j0 + j];
        hRem[base + (int)q] = (float)r.remaining[(size_t)j];  // EstimatedRemainningTimeSec is float32
        hOld[base + (int)q] = r.alloc[(size_t)j];
        hJst[base + (int)q] = r.alloc[(size_t)j] > 0 ? VODA_JOB_RUNNING : VODA_JOB_WAITING;
      }
    });
    if (cudaMemcpyAsync(da, pa, szA, cudaMemcpyHostToDevice, stream) != cudaSuccess) return VODA_E_CUDA;
    voda_alloc_batch ab;
    std::memset(&ab, 0, sizeof(ab));
    ab.algo = VODA_ALGO_ELASTIC_SRJF;
    ab.rounds = T;
    ab.job_off = (const int32_t*)(da + oOff);
    ab.total_gpu = (const int32_t*)(da + oTot);
    ab.min_proc = (const int32_t*)(da + oMin);
    ab.max_proc = (const int32_t*)(da + oMax);
    ab.remaining_sec = (const float*)(da + oRem);
    ab.compaction_threshold = -1;
    ab.max_jobs = maxJobs;
    ab.max_total_gpu = maxTotal;
    int32_t* dG = (int32_t*)(dao + rG);
    rc = voda_allocate_dev(ctx, &ab, dG, (int32_t*)(dao + rSt));
    if (rc) return rc;
    // compareResults (scheduler.go:448-480): what the scheduler would do to every job
    rc = voda_compare_results_dev(ctx, (int32_t)NJ, (const int32_t*)(da + oOld), dG, (const int32_t*)(da + oJst),
                                  (int32_t*)(dao + rAct));
    if (rc) return rc;
    if (cudaMemcpyAsync(pao, dao, szAo, cudaMemcpyDeviceToHost, stream) != cudaSuccess) return VODA_E_CUDA;
    if (cudaStreamSynchronize(stream) != cudaSuccess) return VODA_E_CUDA;
    for (int i = 0; i < T; i++)
      if (hSt[i] != VODA_ST_OK) return VODA_E_INVALID;  // the reference would panic on this trace
    // ---- releaseSlots + the placement round of every replica -----------------------------------
    parallel(todo, [&](int i) {
      Replica& r = reps[(size_t)todo[(size_t)i]];
      const int base = joff[(size_t)i];
      for (size_t q = 0; q < r.live.size(); q++) {
        const int j = r.live[q];
        r.alloc[(size_t)j] = hG[base + (int)q];
        r.liveIndex[(size_t)j] = (int)q;
        const int a = hAct[base + (int)q];
        if (a >= 1 && a <= 4) r.act[a - 1]++;
      }
      releaseSlots(r);
      packPrev(r);
    });
    std::vector<int> poff((size_t)T + 1, 0);
    for (int i = 0; i < T; i++) poff[(size_t)i + 1] = poff[(size_t)i] + (int)reps[(size_t)todo[(size_t)i]].prevJob.size();
    const size_t NP = (size_t)poff[(size_t)T];
    // ---- (B) one placement batch ------------------------------------------------------------------
    const size_t iNo = 0, iSl = iNo + a16(4 * ((size_t)T + 1)), iPo = iSl + a16(4 * NN), iPj = iPo + a16(4 * (NN + 1)),
                 iPw = iPj + a16(4 * NP), szB = iPw + a16(4 * NP);
    const size_t cap = NN + NJ;
    const size_t uSc = 0, uPh = uSc + a16(8 * (size_t)T), uFr = uPh + a16(4 * NN), uCn = uFr + a16(4 * NN),
                 uLn = uCn + a16(4 * (size_t)T), uLj = uLn + a16(4 * cap), uLw = uLj + a16(4 * cap),
                 uCr = uLw + a16(4 * cap), uSt = uCr + a16(4 * (size_t)T), szBo = uSt + a16(4 * (size_t)T);
    char* pb_ = hB.get<char>(szB);
    char* pbo = hBo.get<char>(szBo);
    char* db = dB.get(szB, stream);
    char* dbo = dBo.get(szBo, stream);
    if (!pb_ || !pbo || !db || !dbo) return VODA_E_NOMEM;
    int32_t* hNodeOff = (int32_t*)(pb_ + iNo);
    int32_t* hSlots = (int32_t*)(pb_ + iSl);
    int32_t* hPrevOff = (int32_t*)(pb_ + iPo);
    int32_t* hPrevJob = (int32_t*)(pb_ + iPj);
    int32_t* hPrevW = (int32_t*)(pb_ + iPw);
    std::memcpy(hNodeOff, noff.data(), sizeof(int32_t) * ((size_t)T + 1));
    hPrevOff[0] = 0;
    parallel(todo, [&](int i) {
      Replica& r = reps[(size_t)todo[(size_t)i]];
      const int nb = noff[(size_t)i], pb0 = poff[(size_t)i];
      for (int c = 0; c < r.nodes; c++) {
        hSlots[nb + c] = r.nodeSt[(size_t)c].total;
        hPrevOff[nb + c + 1] = pb0 + r.prevOff[(size_t)c + 1];
      }
      if (!r.prevJob.empty()) {
        std::memcpy(hPrevJob + pb0, r.prevJob.data(), sizeof(int32_t) * r.prevJob.size());
        std::memcpy(hPrevW + pb0, r.prevW.data(), sizeof(int32_t) * r.prevW.size());
      }
    });
    if (cudaMemcpyAsync(db, pb_, szB, cudaMemcpyHostToDevice, stream) != cudaSuccess) return VODA_E_CUDA;
    voda_place_batch pb;
    std::memset(&pb, 0, sizeof(pb));
    pb.rounds = T;
    pb.node_off = (const int32_t*)(db + iNo);
    pb.node_slots = (const int32_t*)(db + iSl);
    pb.req_off = (const int32_t*)(da + oOff);   // the requests of a round = its jobs, same offsets
    pb.req_workers = dG;                         // ... with the counts the allocation just produced
    pb.prev_off = (const int32_t*)(db + iPo);
    pb.prev_job = (const int32_t*)(db + iPj);
    pb.prev_workers = (const int32_t*)(db + iPw);
    pb.max_nodes = maxNodes;
    pb.max_requests = maxJobs;
    voda_place_out dpo;
    dpo.score = (int64_t*)(dbo + uSc);
    dpo.phys_of_logical = (int32_t*)(dbo + uPh);
    dpo.free_slots = (int32_t*)(dbo + uFr);
    dpo.lay_cnt = (int32_t*)(dbo + uCn);
    dpo.lay_node = (int32_t*)(dbo + uLn);
    dpo.lay_job = (int32_t*)(dbo + uLj);
    dpo.lay_workers = (int32_t*)(dbo + uLw);
    dpo.cross = (int32_t*)(dbo + uCr);
    dpo.status = (int32_t*)(dbo + uSt);
    rc = voda_place_dev(ctx, &pb, &dpo);
    if (rc) return rc;
    if (cudaMemcpyAsync(pbo, dbo, szBo, cudaMemcpyDeviceToHost, stream) != cudaSuccess) return VODA_E_CUDA;
    if (cudaStreamSynchronize(stream) != cudaSuccess) return VODA_E_CUDA;
    voda_place_out po;
    po.score = (int64_t*)(pbo + uSc);
    po.phys_of_logical = (int32_t*)(pbo + uPh);
    po.free_slots = (int32_t*)(pbo + uFr);
    po.lay_cnt = (int32_t*)(pbo + uCn);
    po.lay_node = (int32_t*)(pbo + uLn);
    po.lay_job = (int32_t*)(pbo + uLj);
    po.lay_workers = (int32_t*)(pbo + uLw);
    po.cross = (int32_t*)(pbo + uCr);
    po.status = (int32_t*)(pbo + uSt);
    for (int i = 0; i < T; i++)
      if (po.status[i] != VODA_ST_OK && po.status[i] != VODA_ST_NONTERMINATING) return VODA_E_UNSUPPORTED;
    // ---- apply: rename, updateJobStates, updatePodNodeName, history -----------------------------
    parallel(todo, [&](int i) {
      Replica& r = reps[(size_t)todo[(size_t)i]];
      const int base = joff[(size_t)i], nb = noff[(size_t)i], e0 = base + nb;
      r.rounds++;
      fnv(r.hist, ticks);
      fnv(r.hist, (int64_t)r.live.size());
      for (size_t q = 0; q < r.live.size(); q++) fnv(r.hist, hG[base + (int)q]);
      if (po.status[i] == VODA_ST_NONTERMINATING) {
        r.skipped++;
        updateJobStates(r);  // keep jobStates consistent with the released nodeStates
        fnv(r.hist, -1);
      } else {
        std::vector<NodeSt> fresh((size_t)r.nodes);
        std::vector<NodeSt> logical((size_t)r.nodes);
        for (int c = 0; c < r.nodes; c++) {
          logical[(size_t)c].total = r.nodeSt[(size_t)c].total;
          logical[(size_t)c].free = po.free_slots[nb + c];
        }
        for (int t = 0; t < po.lay_cnt[i]; t++) {
          NodeSt& n = logical[(size_t)po.lay_node[e0 + t]];
          const int job = r.live[(size_t)po.lay_job[e0 + t]];
          if (int* w = n.find(job)) *w = po.lay_workers[e0 + t];  // map assignment (:474,480)
          else n.jw.push_back({job, po.lay_workers[e0 + t]});
        }
        for (int c = 0; c < r.nodes; c++) fresh[(size_t)po.phys_of_logical[nb + c]] = std::move(logical[(size_t)c]);  // rename (:509-511)
        r.nodeSt.swap(fresh);
        updateJobStates(r);
        int64_t dw = 0, dl = 0;
        updatePodNodeName(r, &dw, &dl);
        r.migW += dw;
        r.migL += dl;
        for (int c = 0; c < r.nodes; c++) fnv(r.hist, po.phys_of_logical[nb + c]);
        fnv(r.hist, po.score[i]);
        fnv(r.hist, po.cross[i]);
        fnv(r.hist, dw);
        fnv(r.hist, dl);
      }
      for (int j : r.live) r.liveIndex[(size_t)j] = -1;
    });
  }
  for (int k = 0; k < R; k++) {
    const Replica& r = reps[(size_t)k];
    out->rounds[k] = r.rounds;
    out->done[k] = r.done;
    out->sim_time_sec[k] = r.now;
    out->jct_sum_sec[k] = r.jct;
    out->migrated_workers[k] = r.migW;
    out->migrated_launchers[k] = r.migL;
    out->placements_skipped[k] = r.skipped;
    for (int a = 0; a < 4; a++) out->actions[4 * k + a] = r.act[a];
    out->history[k] = r.hist;
  }
  out->ticks[0] = ticks;
  return VODA_OK;
}
