// voda_host.hpp — C++ host side above the C ABI (include/voda_accel.h), mirroring the
// reference's Go interfaces for this path name for name, so a maintainer can read it next to
// the Go code.  (The reference is Go; no Go toolchain exists on this image, so the compiled
// host side is C++ — INTEGRATION.md carries the equivalent cgo shim as source.)
//
//   voda::algorithm::SchedulerAlgorithm / NewAlgorithmFactory   pkg/algorithm/types.go:16-47
//   voda::algorithm::TrainingJob, JobConfig, JobInfo            pkg/common/trainingjob/trainingjob.go:17-66
//   voda::munkres::NewMatrix / ComputeMunkresMax / RowCol       heyfey/munkres, used at
//                                                               pkg/placement/placement_manager.go:505-512
// Failure convention: where the Go code panics (utils.go:29-40, ffdl_optimizer.go:109-114)
// Schedule throws voda::Panic whose what() is the reference's message.
// Header-only; link with -lvoda_accel.  One Engine (voda_ctx) per thread.
#pragma once

#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "voda_accel.h"

namespace voda {

struct Panic : std::runtime_error {
  int status;
  Panic(int st, const std::string& msg) : std::runtime_error(msg), status(st) {}
};

// RAII voda_ctx
class Engine {
 public:
  explicit Engine(int device = 0) {
    if (voda_ctx_create(device, &ctx_) != VODA_OK) throw std::runtime_error(voda_last_error(nullptr));
  }
  ~Engine() { voda_ctx_destroy(ctx_); }
  Engine(const Engine&) = delete;
  Engine& operator=(const Engine&) = delete;
  voda_ctx* get() const { return ctx_; }

 private:
  voda_ctx* ctx_ = nullptr;
};

namespace algorithm {

struct Time {  // what time.Time.Before compares
  int64_t sec = 0;
  int32_t nsec = 0;
};
inline Time MaxTime() { return Time{INT64_MAX - 62135596800LL, 999999999}; }  // types.go:65

struct JobConfig {
  int NumProc = 1, MinNumProc = 1, MaxNumProc = 1, Epochs = 0;
};
struct JobMetrics {
  Time FirstStartTime = MaxTime();
};
struct JobInfo {
  std::string Name, Category, GpuType;
  float EstimatedRemainningTimeSec = 0.0f;
  std::map<std::string, float> Speedup;  // keyed by fmt.Sprint(g); missing key reads 0
};
struct TrainingJob {
  std::string Name;
  int Priority = 0;
  Time SubmitTime;
  JobConfig Config;
  JobMetrics Metrics;
  std::shared_ptr<JobInfo> Info;
};
using ReadyJobs = std::vector<TrainingJob>;
using JobScheduleResult = std::map<std::string, int>;  // types.JobScheduleResult

// NewBaseJobInfo, trainingjob.go:168-186: linear speedup "0".."33"
inline std::shared_ptr<JobInfo> NewBaseJobInfo(const std::string& name) {
  auto i = std::make_shared<JobInfo>();
  i->Name = name;
  i->Speedup["0"] = 0.0f;
  for (int g = 1; g <= 33; g++) i->Speedup[std::to_string(g)] = (float)g;
  return i;
}

class SchedulerAlgorithm {  // types.go:19-24
 public:
  virtual ~SchedulerAlgorithm() = default;
  virtual JobScheduleResult Schedule(const ReadyJobs& jobs, int totalGPU) = 0;
  virtual std::string GetName() const = 0;
  virtual bool NeedJobInfo() const = 0;
};

class Accel final : public SchedulerAlgorithm {
 public:
  Accel(int id, std::string name, std::string schedulerID, std::shared_ptr<Engine> eng)
      : id_(id), name_(std::move(name)), schedulerID_(std::move(schedulerID)), eng_(std::move(eng)) {}
  std::string GetName() const override { return name_; }
  bool NeedJobInfo() const override { return voda_algo_need_job_info(id_) != 0; }

  JobScheduleResult Schedule(const ReadyJobs& jobs, int totalGPU) override {
    const int n = (int)jobs.size();
    int W = 34;  // indices 0..33, trainingjob.go:13,173
    for (const auto& j : jobs)
      if (j.Info)
        for (const auto& kv : j.Info->Speedup) {
          try {
            int g = std::stoi(kv.first);
            if (std::to_string(g) == kv.first && g + 1 > W) W = g + 1;
          } catch (...) {
          }
        }
    std::vector<int32_t> off{0, n}, tot{totalGPU}, sn(n), fn(n), pr(n), np(n), mn(n), mx(n), out(n + 1);
    std::vector<int64_t> ss(n), fs(n);
    std::vector<float> rem(n), sp((size_t)n * W, 0.0f);
    for (int i = 0; i < n; i++) {
      const TrainingJob& j = jobs[i];
      ss[i] = j.SubmitTime.sec; sn[i] = j.SubmitTime.nsec;
      fs[i] = j.Metrics.FirstStartTime.sec; fn[i] = j.Metrics.FirstStartTime.nsec;
      pr[i] = j.Priority; np[i] = j.Config.NumProc; mn[i] = j.Config.MinNumProc; mx[i] = j.Config.MaxNumProc;
      if (j.Info) {
        rem[i] = j.Info->EstimatedRemainningTimeSec;
        for (const auto& kv : j.Info->Speedup) {
          try {
            int g = std::stoi(kv.first);
            if (g >= 0 && g < W && std::to_string(g) == kv.first) sp[(size_t)i * W + g] = kv.second;
          } catch (...) {
          }
        }
      } else if (NeedJobInfo()) {
        throw Panic(VODA_ST_RUNTIME_PANIC, "runtime error: invalid memory address or nil pointer dereference");
      }
    }
    voda_alloc_batch b{};
    b.algo = id_; b.rounds = 1; b.job_off = off.data(); b.total_gpu = tot.data();
    b.submit_sec = ss.data(); b.submit_nsec = sn.data(); b.first_sec = fs.data(); b.first_nsec = fn.data();
    b.priority = pr.data(); b.num_proc = np.data(); b.min_proc = mn.data(); b.max_proc = mx.data();
    b.remaining_sec = rem.data(); b.speedup = sp.data(); b.speedup_stride = W;
    b.queue_num = 2; b.compaction_threshold = 10;  // tiresias.go:24, elastic_tiresias.go:21
    int32_t status = -1;
    if (voda_allocate(eng_->get(), &b, out.data(), &status) != VODA_OK)
      throw std::runtime_error(voda_last_error(eng_->get()));
    if (status != VODA_ST_OK) throw Panic(status, voda_status_str(status));
    JobScheduleResult result;
    for (int i = 0; i < n; i++) result[jobs[i].Name] = out[i];
    return result;
  }

 private:
  int id_;
  std::string name_, schedulerID_;
  std::shared_ptr<Engine> eng_;
};

// types.go:26-47: an unknown name is errors.New("Not found")
inline std::unique_ptr<SchedulerAlgorithm> NewAlgorithmFactory(const std::string& algorithm,
                                                               const std::string& schedulerID,
                                                               std::shared_ptr<Engine> eng) {
  const int id = voda_algo_from_name(algorithm.c_str());
  if (id < 0) throw std::runtime_error("Not found");
  return std::make_unique<Accel>(id, algorithm, schedulerID, std::move(eng));
}

}  // namespace algorithm

namespace munkres {

struct Matrix {
  int n = 0;
  std::vector<int64_t> A;  // row-major, n*row + col
};
struct RowCol {
  int Row, Col;
};
inline Matrix NewMatrix(int n) { return Matrix{n, std::vector<int64_t>((size_t)n * n, 0)}; }

inline std::vector<RowCol> ComputeMunkresMax(const Matrix& m, Engine& eng) {
  std::vector<int32_t> cols((size_t)m.n + 1, -1);
  int32_t status = -1;
  if (voda_munkres_max(eng.get(), m.n, 1, m.A.data(), cols.data(), &status) != VODA_OK)
    throw std::runtime_error(voda_last_error(eng.get()));
  if (status != VODA_ST_OK) throw Panic(status, voda_status_str(status));
  std::vector<RowCol> res;
  for (int r = 0; r < m.n; r++) res.push_back(RowCol{r, cols[(size_t)r]});
  return res;
}

}  // namespace munkres
}  // namespace voda
