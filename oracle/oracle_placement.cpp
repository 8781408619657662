// oracle_placement.cpp — CPU restatement of the arithmetic core of PlacementManager.Place
// (heyfey/vodascheduler pkg/placement/placement_manager.go:306-332, after releaseSlots):
// bestFit (:415-487), scoreCandidates/score (:525-544), bindNodes (:492-522).
//
// TEST INFRASTRUCTURE, NOT PRODUCT (see voda_oracle.h).  Parity pin: SURVEY.md Appendix B
// KATs 12-14 for bestFit; the munkres step is parity-unpinned (see oracle_munkres.cpp).
// Everything the reference takes from Go map iteration order (request order among equal
// sizes, logical-node order, candidate order) is an explicit array of the batch.

#include <algorithm>
#include <cstdint>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "voda_oracle.h"

void voda_oracle_munkres_solve(int n, const int64_t* A, int32_t* colOfRow);

namespace {

// pkg/placement/types.go:9-12, :39-50
struct request {
  std::string job;
  long numWorkers;
};
struct nodeState {
  long totalSlots, freeSlots;
  std::unordered_map<std::string, long> jobNumWorkers;
};
struct Entry {
  int32_t node, job, workers;
};

// bestFit, placement_manager.go:415-487.  Returns false when the reference would loop for
// ever: the no-fit branch hands out the free slots of a node that has none,
// which changes nothing, so the `for requested > 0` loop can never end.
bool bestFit(const std::vector<long>& reqWorkers, std::vector<nodeState>& nodeList,
             std::vector<Entry>& created, long& crossNode) {
  std::vector<request> requests(reqWorkers.size());  // make([]request, len) then append (:416-419)
  for (size_t q = 0; q < reqWorkers.size(); q++)
    requests.push_back(request{"job-" + std::to_string(q), reqWorkers[q]});
  std::stable_sort(requests.begin(), requests.end(),
                   [](const request& a, const request& b) { return a.numWorkers > b.numWorkers; });
  long totalClusterSlots = 0;
  for (const auto& node : nodeList) totalClusterSlots += node.totalSlots;
  crossNode = 0;
  for (const auto& r : requests) {
    long requested = r.numWorkers;
    while (requested > 0) {
      if (totalClusterSlots == 0) return true;  // :455-457
      long bestIdx = -1, maxIdx = 0;
      for (size_t i = 0; i < nodeList.size(); i++) {
        const nodeState& node = nodeList[i];
        if (node.freeSlots >= requested) {
          if (bestIdx == -1)
            bestIdx = (long)i;
          else if (nodeList[(size_t)bestIdx].freeSlots > node.freeSlots)
            bestIdx = (long)i;
        }
        if (nodeList[(size_t)maxIdx].freeSlots < node.freeSlots) maxIdx = (long)i;
      }
      int32_t jobIdx = (int32_t)std::stol(r.job.substr(4));
      if (bestIdx == -1) {
        nodeState& mx = nodeList[(size_t)maxIdx];
        if (mx.freeSlots == 0) return false;  // guard: state would repeat for ever
        mx.jobNumWorkers[r.job] = mx.freeSlots;
        created.push_back(Entry{(int32_t)maxIdx, jobIdx, (int32_t)mx.freeSlots});
        requested -= mx.freeSlots;
        totalClusterSlots -= mx.freeSlots;
        mx.freeSlots = 0;
        crossNode++;
      } else {
        nodeState& bn = nodeList[(size_t)bestIdx];
        bn.jobNumWorkers[r.job] = r.numWorkers;  // the whole request, not `requested` (:479)
        created.push_back(Entry{(int32_t)bestIdx, jobIdx, (int32_t)r.numWorkers});
        requested -= r.numWorkers;
        bn.freeSlots -= r.numWorkers;
        totalClusterSlots -= r.numWorkers;
      }
    }
  }
  return true;
}

// score, placement_manager.go:534-544
int64_t score(const nodeState& position, const nodeState& candidate) {
  long s = 0;
  for (const auto& kv : position.jobNumWorkers) {
    auto it = candidate.jobNumWorkers.find(kv.first);
    long cand = it == candidate.jobNumWorkers.end() ? 0 : it->second;
    if (cand <= kv.second)
      s += cand;
    else
      s += kv.second;
  }
  return (int64_t)s;
}

void run(const voda_place_batch* b, const voda_place_out* out, int32_t r0, int32_t r1) {
  for (int32_t r = r0; r < r1; r++) {
    if (out->migrated_workers) out->migrated_workers[r] = 0;
    if (out->migrated_launchers) out->migrated_launchers[r] = 0;
    const int32_t n0 = b->node_off[r], n = b->node_off[r + 1] - n0;
    const int32_t q0 = b->req_off[r], m = b->req_off[r + 1] - q0;
    const int32_t e0 = q0 + n0;
    if (n <= 0) {  // no nodes: bestFit returns at once (totalClusterSlots == 0), size-0 matrix
      out->lay_cnt[r] = 0; out->score[r] = 0; out->cross[r] = 0; out->status[r] = VODA_ST_OK;
      continue;
    }
    // logical nodes: NewNodeState("TBD", n.totalSlots) (:317-320)
    std::vector<nodeState> logical((size_t)n), physical((size_t)n);
    for (int32_t i = 0; i < n; i++) {
      logical[(size_t)i].totalSlots = logical[(size_t)i].freeSlots = b->node_slots[n0 + i];
      for (int32_t e = b->prev_off[n0 + i]; e < b->prev_off[n0 + i + 1]; e++)
        physical[(size_t)i].jobNumWorkers["job-" + std::to_string(b->prev_job[e])] = b->prev_workers[e];
    }
    std::vector<long> req((size_t)m);
    for (int32_t q = 0; q < m; q++) req[(size_t)q] = b->req_workers[q0 + q];
    std::vector<Entry> created;
    long cross = 0;
    if (!bestFit(req, logical, created, cross)) {
      out->status[r] = VODA_ST_NONTERMINATING;
      out->lay_cnt[r] = 0; out->score[r] = 0; out->cross[r] = 0;
      continue;
    }
    // bindNodes (:492-522): row = logical node, column = physical node
    std::vector<int64_t> M((size_t)n * n);
    for (int32_t i = 0; i < n; i++)
      for (int32_t c = 0; c < n; c++) M[(size_t)i * n + c] = score(logical[(size_t)i], physical[(size_t)c]);
    std::vector<int32_t> colOfRow((size_t)n);
    voda_oracle_munkres_solve(n, M.data(), colOfRow.data());
    int64_t total = 0;
    for (int32_t i = 0; i < n; i++) {
      out->phys_of_logical[n0 + i] = colOfRow[(size_t)i];
      out->free_slots[n0 + i] = (int32_t)logical[(size_t)i].freeSlots;
      total += M[(size_t)i * n + colOfRow[(size_t)i]];
    }
    out->lay_cnt[r] = (int32_t)created.size();
    for (size_t t = 0; t < created.size(); t++) {
      out->lay_node[e0 + (int32_t)t] = created[t].node;
      out->lay_job[e0 + (int32_t)t] = created[t].job;
      out->lay_workers[e0 + (int32_t)t] = created[t].workers;
    }
    out->score[r] = total;
    out->cross[r] = (int32_t)cross;
    out->status[r] = VODA_ST_OK;
    if (b->pods_off && out->migrated_workers && out->migrated_launchers) {
      // rename (:509-511), updateJobStates (:548-566) with the nodes taken by physical index,
      // updatePodNodeName (:571-617) against the caller's pod table
      std::vector<int32_t> logicalOf((size_t)n);
      for (int32_t i = 0; i < n; i++) logicalOf[(size_t)colOfRow[(size_t)i]] = i;
      int64_t deletedWorkers = 0, deletedLaunchers = 0;
      for (int32_t q = 0; q < m; q++) {
        const std::string name = "job-" + std::to_string(q);
        std::vector<int32_t> newPods;  // node of worker pod idx
        bool present = false;
        for (int32_t c = 0; c < n; c++) {
          const auto& jw = logical[(size_t)logicalOf[(size_t)c]].jobNumWorkers;
          auto it = jw.find(name);
          if (it == jw.end()) continue;
          present = true;
          for (long k = 0; k < it->second; k++) newPods.push_back(c);
        }
        if (!present) continue;
        std::vector<int32_t> oldPods;
        for (int32_t t = b->pods_off[q0 + q]; t < b->pods_off[q0 + q + 1]; t++)
          for (int32_t k = 0; k < b->pods_slots[t]; k++) oldPods.push_back(b->pods_node[t]);
        long deleted = 0;
        for (size_t idx = 0; idx < newPods.size() && idx < oldPods.size(); idx++)
          if (newPods[idx] != oldPods[idx]) deleted++;
        deletedWorkers += deleted;
        if (deleted == (long)newPods.size()) deletedLaunchers++;
      }
      out->migrated_workers[r] = (int32_t)deletedWorkers;
      out->migrated_launchers[r] = (int32_t)deletedLaunchers;
    }
  }
}

}  // namespace

extern "C" int voda_oracle_place(const voda_place_batch* b, const voda_place_out* out,
                                 int32_t threads) {
  if (!b || !out || b->rounds < 0) return VODA_E_INVALID;
  if (threads <= 1 || b->rounds < 2) {
    run(b, out, 0, b->rounds);
    return VODA_OK;
  }
  std::vector<std::thread> pool;
  for (int32_t t = 0; t < threads; t++) {
    int32_t r0 = (int32_t)((int64_t)b->rounds * t / threads), r1 = (int32_t)((int64_t)b->rounds * (t + 1) / threads);
    pool.emplace_back(run, b, out, r0, r1);
  }
  for (auto& th : pool) th.join();
  return VODA_OK;
}
