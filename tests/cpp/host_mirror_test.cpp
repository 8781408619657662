// Drives the C++ host mirror (include/voda_host.hpp) the way the reference's callers drive the
// Go interfaces (resource_allocator.go:89-106, placement_manager.go:505-512), on the KATs of
// SURVEY.md Appendix B/C.  Built and run by tests/test_cpp_host_gpu.py on a GPU box.
#include <cstdio>
#include <cstdlib>

#include "voda_host.hpp"

using namespace voda::algorithm;

static std::shared_ptr<JobInfo> info(const std::vector<float>& curve, float rem) {
  auto i = std::make_shared<JobInfo>();
  i->EstimatedRemainningTimeSec = rem;
  for (size_t g = 0; g < curve.size(); g++) i->Speedup[std::to_string(g)] = curve[g];
  return i;
}
static TrainingJob job(const char* name, int64_t submit, int64_t first, int prio, int np, int mn, int mx, float rem,
                       const std::vector<float>& curve) {
  TrainingJob j;
  j.Name = name; j.Priority = prio; j.SubmitTime = Time{submit, 0};
  j.Config = JobConfig{np, mn, mx, 0}; j.Metrics.FirstStartTime = Time{first, 0}; j.Info = info(curve, rem);
  return j;
}
#define CHECK(c) do { if (!(c)) { std::printf("FAILED %s:%d %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

int main() {
  const std::vector<float> C{0, 1, 1.9f, 2.7f, 3.4f, 4.0f, 4.5f, 4.9f, 5.2f, 5.4f};
  const std::vector<float> D{0, 1, 1.5f, 1.8f, 2.0f, 2.1f, 2.15f, 2.17f, 2.18f, 2.18f};
  std::vector<float> LIN(34);
  for (int g = 0; g < 34; g++) LIN[g] = (float)g;
  auto eng = std::make_shared<voda::Engine>(0);

  {  // KAT-1 ElasticFIFO
    ReadyJobs jobs{job("A", 1, 0, 0, 1, 1, 4, 30, LIN), job("B", 2, 0, 0, 1, 2, 2, 10, LIN), job("C", 3, 0, 0, 1, 2, 8, 20, LIN)};
    auto a = NewAlgorithmFactory("ElasticFIFO", "sched-0", eng);
    CHECK(a->GetName() == "ElasticFIFO" && !a->NeedJobInfo());
    auto r = a->Schedule(jobs, 8);
    CHECK(r["A"] == 3 && r["B"] == 2 && r["C"] == 3);
    // KAT-2 ElasticSRJF panics: exceeded job max gpu
    auto s = NewAlgorithmFactory("ElasticSRJF", "sched-0", eng);
    bool panicked = false;
    try { s->Schedule(jobs, 8); } catch (const voda::Panic& p) {
      panicked = p.status == VODA_ST_EXCEEDED_MAX && std::string(p.what()) == "Invalid GPU allocations: exceeded job max gpu";
    }
    CHECK(panicked);
  }
  {  // KAT-5 ElasticTiresias
    ReadyJobs jobs{job("A", 0, 5, 0, 2, 1, 4, 0, C), job("B", 0, 3, 1, 2, 1, 4, 0, D), job("C", 0, 4, 0, 4, 2, 6, 0, C),
                   job("D", 0, 1, 1, 1, 1, 2, 0, LIN)};
    auto r = NewAlgorithmFactory("ElasticTiresias", "s", eng)->Schedule(jobs, 8);
    CHECK(r["A"] == 2 && r["B"] == 1 && r["C"] == 4 && r["D"] == 1);
  }
  {  // KAT-8 FfDL and KAT-11/8 AFS-L
    ReadyJobs jobs{job("A", 1, 0, 0, 1, 1, 4, 100, C), job("B", 2, 0, 0, 1, 1, 4, 50, D), job("C", 3, 0, 0, 1, 1, 6, 200, C)};
    auto r = NewAlgorithmFactory("FfDLOptimizer", "s", eng)->Schedule(jobs, 8);
    CHECK(r["A"] == 4 && r["B"] == 1 && r["C"] == 3);
    auto q = NewAlgorithmFactory("AFS-L", "s", eng)->Schedule(jobs, 8);
    CHECK(q["A"] == 3 && q["B"] == 2 && q["C"] == 3);
  }
  {  // unknown policy
    bool notFound = false;
    try { NewAlgorithmFactory("Pollux", "s", eng); } catch (const std::runtime_error& e) { notFound = std::string(e.what()) == "Not found"; }
    CHECK(notFound);
  }
  {  // munkres tie-break KAT: [[4,0,0],[0,4,0],[4,4,0]] -> (0,2)(1,1)(2,0)
    auto m = voda::munkres::NewMatrix(3);
    m.A = {4, 0, 0, 0, 4, 0, 4, 4, 0};
    auto res = voda::munkres::ComputeMunkresMax(m, *eng);
    CHECK(res.size() == 3 && res[0].Col == 2 && res[1].Col == 1 && res[2].Col == 0);
  }
  std::printf("host mirror OK\n");
  return 0;
}
