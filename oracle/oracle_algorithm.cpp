// oracle_algorithm.cpp — structure-faithful CPU restatement of heyfey/vodascheduler
// pkg/algorithm (all eight policies behind NewAlgorithmFactory).
//
// TEST INFRASTRUCTURE, NOT PRODUCT (see voda_oracle.h).  Parity pin: SURVEY.md Appendix B
// known-answer tests only (the reference ships no tests for this path and cannot be built
// here: no Go toolchain).
//
// "Structure-faithful" means: the same containers the Go code uses (string-keyed hash maps
// for results, gains and speedup curves; fmt.Sprint(g) keys), the same loop nests, a stable
// sort wherever the reference calls sort.SliceStable, float32 arithmetic in the reference's
// evaluation order, and Go panics restated as a thrown status.  It is deliberately slow in
// the same places the reference is slow.  Paths below are relative to the reference tree.
//
// Build: g++ -O2 -fno-fast-math -ffp-contract=off (no FMA contraction: Go on amd64 rounds
// every float32 operation separately).

#include <cstring>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <memory>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "voda_oracle.h"

namespace {

// ---- Go-shaped types (pkg/common/trainingjob/trainingjob.go:17-66) ----------------------
struct Time {
  int64_t sec;
  int32_t nsec;
  // time.Time.Before on monotonic-free values: lexicographic (sec, nsec).
  bool Before(const Time& u) const { return sec < u.sec || (sec == u.sec && nsec < u.nsec); }
};
using SpeedupMap = std::unordered_map<std::string, float>;
struct JobInfo {
  float EstimatedRemainningTimeSec = 0.0f;  // trainingjob.go:63 (float32)
  SpeedupMap Speedup;                       // trainingjob.go:64 (map[string]float32)
};
struct JobConfig {
  long NumProc = 0, MinNumProc = 0, MaxNumProc = 0;  // Go int is 64-bit on amd64
};
struct TrainingJob {
  std::string Name;
  long Priority = 0;
  Time SubmitTime{0, 0};
  JobConfig Config;
  Time FirstStartTime{0, 0};  // Metrics.FirstStartTime
  std::shared_ptr<JobInfo> Info;
};
using ReadyJobs = std::vector<TrainingJob>;
using Result = std::unordered_map<std::string, long>;  // types.JobScheduleResult (types.go:61)

struct Panic {
  int status;
};

inline std::string Sprint(long g) { return std::to_string(g); }  // fmt.Sprint(int)

// Go map read: a missing key yields the zero value and does not insert.
inline float at(const SpeedupMap& m, const std::string& k) {
  auto it = m.find(k);
  return it == m.end() ? 0.0f : it->second;
}

// allTrue, pkg/algorithm/utils.go:9-16
bool allTrue(const std::unordered_map<std::string, bool>& sastified) {
  for (const auto& kv : sastified)
    if (!kv.second) return false;
  return true;
}

// validateResult, pkg/algorithm/utils.go:18-42.  The Go code ranges over a map, so WHICH
// message fires is random when several jobs offend; our tie rule (include/voda_accel.h) is
// the smallest status among the offending entries' first failing checks.
void validateResult(long totalGPU, const Result& result, const ReadyJobs& jobs) {
  std::unordered_map<std::string, long> jobMaxGPU, jobMinGPU;
  for (const auto& job : jobs) {
    jobMaxGPU[job.Name] = job.Config.MaxNumProc;
    jobMinGPU[job.Name] = job.Config.MinNumProc;
  }
  long allocatedGPU = 0;
  int worst = 0;
  for (const auto& kv : result) {
    long n = kv.second;
    int st = 0;
    if (n < 0)
      st = VODA_ST_NEGATIVE;
    else if (n > 0 && n < jobMinGPU[kv.first])
      st = VODA_ST_LESS_THAN_MIN;
    else if (n > jobMaxGPU[kv.first])
      st = VODA_ST_EXCEEDED_MAX;
    if (st && (!worst || st < worst)) worst = st;
    allocatedGPU += n;
  }
  if (worst) throw Panic{worst};
  if (allocatedGPU > totalGPU) throw Panic{VODA_ST_EXCEEDED_TOTAL};
}

// remove / index, pkg/algorithm/elastic_tiresias.go:174-190
void removeAt(ReadyJobs& slice, size_t s) { slice.erase(slice.begin() + (long)s); }
size_t indexOf(const ReadyJobs& jobs, const std::string& name) {
  for (size_t i = 0; i < jobs.size(); i++)
    if (jobs[i].Name == name) return i;
  throw Panic{VODA_ST_NOT_FOUND};
}

// calNextGain, elastic_tiresias.go:169-172 (one float32 subtraction)
float calNextGain(const SpeedupMap& speedup, long workers) {
  return at(speedup, Sprint(workers + 1)) - at(speedup, Sprint(workers));
}

// ---- FIFO, pkg/algorithm/fifo.go:25-52 ---------------------------------------------------
void FIFO_Schedule(ReadyJobs& jobs, long totalGPU, Result& result) {
  long freeGPU = totalGPU;
  std::stable_sort(jobs.begin(), jobs.end(), [](const TrainingJob& a, const TrainingJob& b) {
    return a.SubmitTime.Before(b.SubmitTime);
  });
  for (const auto& job : jobs) {
    result[job.Name] = 0;
    if (freeGPU >= job.Config.MinNumProc) {
      result[job.Name] = job.Config.MinNumProc;
      freeGPU -= job.Config.MinNumProc;
    }
  }
  validateResult(totalGPU, result, jobs);
}

// ---- SRJF, pkg/algorithm/srjf.go:25-52 ---------------------------------------------------
void SRJF_Schedule(ReadyJobs& jobs, long totalGPU, Result& result) {
  long freeGPU = totalGPU;
  std::stable_sort(jobs.begin(), jobs.end(), [](const TrainingJob& a, const TrainingJob& b) {
    return a.Info->EstimatedRemainningTimeSec < b.Info->EstimatedRemainningTimeSec;
  });
  for (const auto& job : jobs) {
    result[job.Name] = 0;
    if (freeGPU >= job.Config.MinNumProc) {
      result[job.Name] = job.Config.MinNumProc;
      freeGPU -= job.Config.MinNumProc;
    }
  }
  validateResult(totalGPU, result, jobs);
}

// ---- ElasticFIFO / ElasticSRJF, elastic_fifo.go:25-75 and elastic_srjf.go:25-72 ----------
// The two differ in the sort key and in one line of phase one (elastic_fifo.go:46-48 marks a
// job satisfied when min == max; elastic_srjf.go:42-47 has no such line).
void ElasticFill_Schedule(ReadyJobs& jobs, long totalGPU, Result& result, bool srjf) {
  std::unordered_map<std::string, bool> sastified;
  long freeGPU = totalGPU;
  if (srjf) {
    std::stable_sort(jobs.begin(), jobs.end(), [](const TrainingJob& a, const TrainingJob& b) {
      return a.Info->EstimatedRemainningTimeSec < b.Info->EstimatedRemainningTimeSec;
    });
  } else {
    std::stable_sort(jobs.begin(), jobs.end(), [](const TrainingJob& a, const TrainingJob& b) {
      return a.SubmitTime.Before(b.SubmitTime);
    });
  }
  // allocate the basic portion
  for (const auto& job : jobs) {
    result[job.Name] = 0;
    sastified[job.Name] = false;
    if (freeGPU >= job.Config.MinNumProc) {
      result[job.Name] = job.Config.MinNumProc;
      freeGPU -= job.Config.MinNumProc;
      if (!srjf && result[job.Name] == job.Config.MaxNumProc) sastified[job.Name] = true;
    } else {
      sastified[job.Name] = true;  // unable to allocate minGPU to the job
    }
  }
  // allocate the remaining GPUs, one per eligible job per pass
  while (freeGPU > 0 && !allTrue(sastified)) {
    for (const auto& job : jobs) {
      if (result[job.Name] < job.Config.MaxNumProc || !sastified[job.Name]) {
        result[job.Name] += 1;
        freeGPU -= 1;
        if (result[job.Name] == job.Config.MaxNumProc) sastified[job.Name] = true;
        if (freeGPU == 0) break;
      }
    }
  }
  validateResult(totalGPU, result, jobs);
}

// ---- Tiresias, pkg/algorithm/tiresias.go:51-99 --------------------------------------------
void Tiresias_Schedule(ReadyJobs& jobs, long totalGPU, Result& result, long queueNum) {
  long freeGPU = totalGPU;
  std::unordered_map<long, ReadyJobs> queues;
  for (long p = 0; p < queueNum; p++) queues[p] = ReadyJobs();
  for (const auto& job : jobs) queues[job.Priority].push_back(job);
  for (long p = 0; p < queueNum; p++) {
    auto& q = queues[p];
    std::stable_sort(q.begin(), q.end(), [](const TrainingJob& a, const TrainingJob& b) {
      return a.FirstStartTime.Before(b.FirstStartTime);
    });
  }
  for (long p = 0; p < queueNum; p++) {
    for (const auto& job : queues[p]) {
      result[job.Name] = 0;
      if (freeGPU >= job.Config.NumProc) {
        result[job.Name] = job.Config.NumProc;
        freeGPU -= job.Config.NumProc;
      }
    }
  }
  // Jobs whose priority is outside [0, queueNum) never get a result entry (tiresias.go:81-90
  // only visits the known queues); the ABI reports 0 for them.  validateResult ranges over
  // the result map, so such jobs are simply not validated.
  validateResult(totalGPU, result, jobs);
}

// ---- ElasticTiresias, pkg/algorithm/elastic_tiresias.go:37-159 ----------------------------
void ElasticTiresias_Schedule(ReadyJobs& jobs, long totalGPU, Result& result, long queueNum,
                              long compactionThreshold) {
  long freeGPU = totalGPU;
  long pendings = (long)jobs.size();
  std::unordered_map<std::string, float> gain;
  ReadyJobs oriJobs = jobs;  // :45-46

  std::unordered_map<long, ReadyJobs> queues;
  for (long p = 0; p < queueNum; p++) queues[p] = ReadyJobs();
  for (const auto& job : jobs) {  // :54-60
    queues[job.Priority].push_back(job);
    gain[job.Name] =
        at(job.Info->Speedup, Sprint(job.Config.MinNumProc)) / (float)job.Config.MinNumProc;
    result[job.Name] = 0;
  }
  for (long p = 0; p < queueNum; p++) {  // :65-69
    auto& q = queues[p];
    std::stable_sort(q.begin(), q.end(), [](const TrainingJob& a, const TrainingJob& b) {
      return a.FirstStartTime.Before(b.FirstStartTime);
    });
  }
  for (long p = 0; p < queueNum; p++) {  // :75-85 allocate the basic portion
    for (const auto& job : queues[p]) {
      if (freeGPU >= job.Config.NumProc) {
        result[job.Name] = job.Config.NumProc;
        freeGPU -= job.Config.NumProc;
        pendings -= 1;
        gain[job.Name] = calNextGain(job.Info->Speedup, result[job.Name]);
      }
    }
  }
  if (pendings > compactionThreshold) {  // :89-102 compaction
    for (long p = 1; p < queueNum; p++) {
      for (const auto& job : queues[p]) {
        if (result[job.Name] != 0) {
          freeGPU += (result[job.Name] - job.Config.MinNumProc);
          result[job.Name] = job.Config.MinNumProc;
          gain[job.Name] = calNextGain(job.Info->Speedup, result[job.Name]);
        }
      }
    }
  }
  // :107-115 remove the jobs that are already satisfied or cannot run
  std::vector<std::string> toRemove;
  for (const auto& job : jobs)
    if (result[job.Name] >= job.Config.MaxNumProc || freeGPU < job.Config.MinNumProc)
      toRemove.push_back(job.Name);
  for (const auto& name : toRemove) removeAt(jobs, indexOf(jobs, name));

  while (freeGPU > 0 && !jobs.empty()) {  // :119-152
    std::stable_sort(jobs.begin(), jobs.end(), [](const TrainingJob& a, const TrainingJob& b) {
      return a.Priority < b.Priority;
    });
    std::stable_sort(jobs.begin(), jobs.end(),
                     [&gain](const TrainingJob& a, const TrainingJob& b) {
                       return gain[a.Name] > gain[b.Name];
                     });
    TrainingJob job = jobs[0];
    if (gain[job.Name] <= 0) break;
    if (result[job.Name] == 0) {
      if (freeGPU >= job.Config.MinNumProc) {
        const float before = gain[job.Name];
        result[job.Name] = job.Config.MinNumProc;
        freeGPU -= job.Config.MinNumProc;
        gain[job.Name] = calNextGain(job.Info->Speedup, result[job.Name]);
        // NOT in the reference: with MinNumProc == 0 this branch changes nothing but the gain;
        // once the gain repeats, the Go loop never ends.  Reported as a status instead.
        const float after = gain[job.Name];
        if (job.Config.MinNumProc == 0 && std::memcmp(&before, &after, 4) == 0 && !(after <= 0))
          throw Panic{VODA_ST_NONTERMINATING};
      } else {
        removeAt(jobs, indexOf(jobs, job.Name));
      }
    } else {
      result[job.Name] += 1;
      freeGPU -= 1;
      gain[job.Name] = calNextGain(job.Info->Speedup, result[job.Name]);
      if (result[job.Name] >= job.Config.MaxNumProc) removeAt(jobs, indexOf(jobs, job.Name));
    }
  }
  validateResult(totalGPU, result, oriJobs);  // :157
}

// ---- FfDLOptimizer, pkg/algorithm/ffdl_optimizer.go:32-130 --------------------------------
void FfDL_body(ReadyJobs& jobs, long totalGPU, Result& result) {
  std::stable_sort(jobs.begin(), jobs.end(), [](const TrainingJob& a, const TrainingJob& b) {
    return a.SubmitTime.Before(b.SubmitTime);
  });
  long K = totalGPU;
  // make(ReadyJobs, 1, K+1) panics in the Go runtime when the capacity is below the length
  if (K + 1 < 1) throw Panic{VODA_ST_RUNTIME_PANIC};
  ReadyJobs feasibleJobs(1);  // index 0 is ignored (:54-56)
  for (long i = 0; i < K; i++) {
    if (i < (long)jobs.size())
      feasibleJobs.push_back(jobs[(size_t)i]);
    else
      break;
  }
  long J = (long)feasibleJobs.size() - 1;
  std::vector<std::vector<float>> P((size_t)J + 1);
  std::vector<std::vector<long>> SOL((size_t)J + 1);
  for (long i = 0; i <= J; i++) {
    P[(size_t)i].assign((size_t)K + 1, i == 0 ? 0.0f : -10000.0f);  // :78-87
    SOL[(size_t)i].assign((size_t)K + 1, 0);
  }
  for (long j = 1; j <= J; j++) {  // :91-105
    const TrainingJob& fj = feasibleJobs[(size_t)j];
    for (long k = 1; k <= K; k++) {
      for (long g = 1; g <= fj.Config.MaxNumProc; g++) {
        if (k - g >= 0) {
          float p = at(fj.Info->Speedup, Sprint(g)) + P[(size_t)j - 1][(size_t)(k - g)];
          if (p > P[(size_t)j][(size_t)k]) {
            P[(size_t)j][(size_t)k] = p;
            SOL[(size_t)j][(size_t)k] = g;
          }
        }
      }
    }
  }
  if (P[(size_t)J][(size_t)K] <= 0) throw Panic{VODA_ST_INFEASIBLE};  // :109-114
  long j = J, k = K;
  while (j > 0) {  // :116-125
    result[feasibleJobs[(size_t)j].Name] = SOL[(size_t)j][(size_t)k];
    k -= SOL[(size_t)j][(size_t)k];
    j -= 1;
  }
}

void FfDL_Schedule(ReadyJobs& jobs, long totalGPU, Result& result) {
  if (jobs.empty()) return;  // :33-37
  for (const auto& job : jobs) result[job.Name] = 0;
  // defer validateResult(totalGPU, result, jobs) (:42): it runs on the way out of a panic
  // too, and a panic raised by the deferred call replaces the one in flight.
  try {
    FfDL_body(jobs, totalGPU, result);
  } catch (const Panic&) {
    validateResult(totalGPU, result, jobs);
    throw;
  }
  validateResult(totalGPU, result, jobs);
}

// ---- AFS-L, pkg/algorithm/afsl.go:31-106 --------------------------------------------------
thread_local int64_t g_afsl_steps = 0, g_afsl_changes = 0, g_afsl_picks = 0;

double jobLength(const TrainingJob& job, long workers) {  // :94-100
  if (workers == 0) return std::numeric_limits<double>::infinity();
  return (double)(job.Info->EstimatedRemainningTimeSec / at(job.Info->Speedup, Sprint(workers)));
}

bool evaluate(const TrainingJob& j, const TrainingJob& jb, Result& result) {  // :102-106
  float left = (at(jb.Info->Speedup, Sprint(result[jb.Name] + 1)) -
                at(jb.Info->Speedup, Sprint(result[jb.Name]))) /
               at(jb.Info->Speedup, Sprint(result[jb.Name] + 1));
  float right = (at(j.Info->Speedup, Sprint(result[j.Name] + 1)) -
                 at(j.Info->Speedup, Sprint(result[j.Name]))) /
                at(j.Info->Speedup, Sprint(result[j.Name]));
  return left > right;
}

TrainingJob topPriority(const ReadyJobs& jobs, Result& result) {  // :72-92
  const TrainingJob* j = &jobs[0];
  for (size_t i = 1; i < jobs.size(); i++) {
    const TrainingJob* jb = &jobs[i];
    const TrainingJob* before = j;
    g_afsl_steps++;
    if (result[j->Name] == 0 && result[jb->Name] == 0) {
      if (j->Info->EstimatedRemainningTimeSec >= jb->Info->EstimatedRemainningTimeSec) j = jb;
    } else {
      double lenA = jobLength(*j, result[j->Name]);
      double lenB = jobLength(*jb, result[j->Name]);  // the champion's count for both (:82)
      if (lenA >= lenB) std::swap(j, jb);             // j := shorter, jb := longer
      if (evaluate(*j, *jb, result)) j = jb;
    }
    if (j != before) g_afsl_changes++;
  }
  return *j;
}

void AFSL_Schedule(ReadyJobs& jobs, long totalGPU, Result& result) {
  for (const auto& job : jobs) result[job.Name] = 0;
  ReadyJobs oriJobs = jobs;  // :36-38
  auto body = [&]() {
    std::stable_sort(jobs.begin(), jobs.end(), [](const TrainingJob& a, const TrainingJob& b) {
      return a.SubmitTime.Before(b.SubmitTime);
    });
    long freeGPU = totalGPU;
    while (freeGPU > 0 && !jobs.empty()) {  // :47-58
      TrainingJob job = topPriority(jobs, result);
      g_afsl_picks++;
      result[job.Name] += 1;
      freeGPU -= 1;
      if (result[job.Name] >= job.Config.MaxNumProc) removeAt(jobs, indexOf(jobs, job.Name));
    }
  };
  try {
    body();
  } catch (const Panic&) {
    validateResult(totalGPU, result, oriJobs);  // deferred at :39
    throw;
  }
  validateResult(totalGPU, result, oriJobs);
}

// ---- batch driver --------------------------------------------------------------------------
void run_rounds(const voda_alloc_batch* b, int32_t* out_gpus, int32_t* out_status, int32_t r0,
                int32_t r1) {
  const long W = b->speedup_stride;
  const long queueNum = b->queue_num > 0 ? b->queue_num : 2;                        // tiresias.go:24
  const long compaction = b->compaction_threshold >= 0 ? b->compaction_threshold : 10;  // elastic_tiresias.go:21
  for (int32_t r = r0; r < r1; r++) {
    const int32_t o = b->job_off[r], n = b->job_off[r + 1] - o;
    ReadyJobs jobs((size_t)n);
    for (int32_t i = 0; i < n; i++) {
      TrainingJob& t = jobs[(size_t)i];
      const int32_t f = o + i;
      t.Name = "job-" + std::to_string(i);
      t.Priority = b->priority ? b->priority[f] : 0;
      t.SubmitTime = Time{b->submit_sec ? b->submit_sec[f] : 0, b->submit_nsec ? b->submit_nsec[f] : 0};
      t.FirstStartTime = Time{b->first_sec ? b->first_sec[f] : 0, b->first_nsec ? b->first_nsec[f] : 0};
      t.Config.NumProc = b->num_proc ? b->num_proc[f] : 0;
      t.Config.MinNumProc = b->min_proc ? b->min_proc[f] : 0;
      t.Config.MaxNumProc = b->max_proc ? b->max_proc[f] : 0;
      t.Info = std::make_shared<JobInfo>();
      t.Info->EstimatedRemainningTimeSec = b->remaining_sec ? b->remaining_sec[f] : 0.0f;
      if (b->speedup)
        for (long g = 0; g < W; g++)  // dense table, or a pool of curves the jobs point into (speedup_off)
          t.Info->Speedup[Sprint(g)] = b->speedup[(b->speedup_off ? (size_t)b->speedup_off[f] : (size_t)f * (size_t)W) + (size_t)g];
    }
    ReadyJobs input = jobs;  // Schedule mutates its argument; keep names in input order
    Result result;
    int status = VODA_ST_OK;
    try {
      switch (b->algo) {
        case VODA_ALGO_FIFO: FIFO_Schedule(jobs, b->total_gpu[r], result); break;
        case VODA_ALGO_ELASTIC_FIFO: ElasticFill_Schedule(jobs, b->total_gpu[r], result, false); break;
        case VODA_ALGO_SRJF: SRJF_Schedule(jobs, b->total_gpu[r], result); break;
        case VODA_ALGO_ELASTIC_SRJF: ElasticFill_Schedule(jobs, b->total_gpu[r], result, true); break;
        case VODA_ALGO_TIRESIAS: Tiresias_Schedule(jobs, b->total_gpu[r], result, queueNum); break;
        case VODA_ALGO_ELASTIC_TIRESIAS:
          ElasticTiresias_Schedule(jobs, b->total_gpu[r], result, queueNum, compaction);
          break;
        case VODA_ALGO_FFDL_OPTIMIZER: FfDL_Schedule(jobs, b->total_gpu[r], result); break;
        case VODA_ALGO_AFSL: AFSL_Schedule(jobs, b->total_gpu[r], result); break;
        default: status = VODA_ST_NOT_FOUND;
      }
    } catch (const Panic& p) {
      status = p.status;
    }
    out_status[r] = status;
    for (int32_t i = 0; i < n; i++) {
      auto it = result.find(input[(size_t)i].Name);
      long v = it == result.end() ? 0 : it->second;
      if (v > INT32_MAX) v = INT32_MAX;
      if (v < INT32_MIN) v = INT32_MIN;
      out_gpus[o + i] = (int32_t)v;
    }
  }
}

}  // namespace

extern "C" int voda_oracle_allocate(const voda_alloc_batch* b, int32_t* out_gpus,
                                    int32_t* out_status, int32_t threads) {
  if (!b || !b->job_off || !b->total_gpu || b->rounds < 0) return VODA_E_INVALID;
  if (b->algo < 0 || b->algo >= VODA_ALGO_COUNT) return VODA_E_INVALID;
  if (threads <= 1 || b->rounds < 2) {
    g_afsl_steps = g_afsl_changes = g_afsl_picks = 0;
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

extern "C" void voda_oracle_afsl_counters(int64_t* fold_steps, int64_t* champion_changes,
                                          int64_t* picks) {
  if (fold_steps) *fold_steps = g_afsl_steps;
  if (champion_changes) *champion_changes = g_afsl_changes;
  if (picks) *picks = g_afsl_picks;
}
