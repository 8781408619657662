// hostpack.h — host-side helpers of the pipelined host path: a small persistent thread pool
// and the packer that strips a dense speedup table down to the entries FfDL can read.
//
// Why: end to end the FfDL path is PCIe-bound, and 136 of the 160 bytes per job are the dense
// 34-entry speedup row (pkg/common/trainingjob/trainingjob.go:64) of which the DP only reads
// g = 1 .. min(MaxNumProc, totalGPU, W-1) (ffdl_optimizer.go:93-97) — about 9 entries on the
// BASELINE workload.  Packing the rows on the host (a memory-bound gather spread over a few
// threads, overlapped with the previous chunk's DMA) cuts the bytes that cross PCIe ~2.7x.
// This is data movement only: no scheduling arithmetic happens on the CPU.
// MEASURED: it does not pay on the 16-core B200 host — the strided gather pulls nearly every
// cache line of the dense table through the CPU (about 35 GB/s with 8 threads) while the DMA
// engine moves the dense table at 55 GB/s — so the path is opt-in (VODA_PACK_ROWS=1).
#pragma once

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace voda {

class HostPool {
 public:
  explicit HostPool(int n) : n_(n < 1 ? 1 : n) {
    for (int i = 0; i < n_; i++) th_.emplace_back([this, i] { loop(i); });
  }
  ~HostPool() {
    {
      std::lock_guard<std::mutex> g(m_);
      stop_ = true;
      gen_++;
    }
    cv_.notify_all();
    for (auto& t : th_) t.join();
  }
  int size() const { return n_; }
  // run f(worker) on every worker, return when all are done
  void run(const std::function<void(int)>& f) {
    std::unique_lock<std::mutex> g(m_);
    job_ = &f;
    pending_ = n_;
    gen_++;
    cv_.notify_all();
    done_.wait(g, [this] { return pending_ == 0; });
    job_ = nullptr;
  }

 private:
  void loop(int id) {
    int seen = 0;
    for (;;) {
      const std::function<void(int)>* f;
      {
        std::unique_lock<std::mutex> g(m_);
        cv_.wait(g, [&] { return gen_ != seen; });
        seen = gen_;
        if (stop_) return;
        f = job_;
      }
      (*f)(id);
      {
        std::lock_guard<std::mutex> g(m_);
        if (--pending_ == 0) done_.notify_all();
      }
    }
  }
  int n_;
  std::vector<std::thread> th_;
  std::mutex m_;
  std::condition_variable cv_, done_;
  const std::function<void(int)>* job_ = nullptr;
  int gen_ = 0, pending_ = 0;
  bool stop_ = false;
};

// entries of job j's row that the FfDL DP can read: indices 0 .. min(max_proc, total, W-1)
inline int ffdl_row_len(int max_proc, int total, int W) {
  int g = std::min(std::min(max_proc, total), W - 1);
  if (g < 0) g = 0;
  return g + 1;
}

// Pass 1: per-worker totals.  Pass 2: offsets + copies.  Rounds [r0, r1) of the batch;
// off[] are the batch's absolute job offsets; sp_off (nj + 1 ints) and packed are outputs.
// Returns the number of floats packed.
inline size_t ffdl_pack_rows(HostPool& pool, const int32_t* off, const int32_t* total_gpu,
                             const int32_t* max_proc, const float* speedup, int W, int r0, int r1,
                             int32_t* sp_off, float* packed, size_t packed_cap) {
  const int T = pool.size();
  const int nr = r1 - r0;
  const int j0 = off[r0];
  std::vector<size_t> sum((size_t)T + 1, 0);
  auto range = [&](int w, int* a, int* b) {
    *a = r0 + (int)((long long)nr * w / T);
    *b = r0 + (int)((long long)nr * (w + 1) / T);
  };
  pool.run([&](int w) {
    int a, b;
    range(w, &a, &b);
    size_t s = 0;
    for (int r = a; r < b; r++) {
      const int tg = total_gpu[r];
      for (int j = off[r]; j < off[r + 1]; j++) s += (size_t)ffdl_row_len(max_proc[j], tg, W);
    }
    sum[(size_t)w + 1] = s;
  });
  for (int w = 0; w < T; w++) sum[(size_t)w + 1] += sum[(size_t)w];
  const size_t total = sum[(size_t)T];
  if (total > packed_cap) return total;  // caller grows the buffer and calls again
  pool.run([&](int w) {
    int a, b;
    range(w, &a, &b);
    size_t pos = sum[(size_t)w];
    for (int r = a; r < b; r++) {
      const int tg = total_gpu[r];
      for (int j = off[r]; j < off[r + 1]; j++) {
        const int len = ffdl_row_len(max_proc[j], tg, W);
        sp_off[j - j0] = (int32_t)pos;
        // rows are short (about 9 floats): copy in 16-byte steps, rounding the length up.  The
        // few floats written past `len` land where the NEXT row (written after this one, in
        // increasing order) or the slack at the end of the buffer goes; the source over-read
        // stays inside this job's own W-wide row because len <= W.
        const float* src = speedup + (size_t)j * (size_t)W;
        float* dst = packed + pos;
        // (the last row of a worker's range is copied exactly: the next floats belong to the
        // neighbouring worker, which writes them concurrently)
        const bool lastRow = (r == b - 1) && (j == off[r + 1] - 1);
        const int len4 = lastRow ? (len & ~3) : std::min((len + 3) & ~3, W);
        int i = 0;
        for (; i + 4 <= len4; i += 4) {
          float t0 = src[i], t1 = src[i + 1], t2 = src[i + 2], t3 = src[i + 3];
          dst[i] = t0; dst[i + 1] = t1; dst[i + 2] = t2; dst[i + 3] = t3;
        }
        for (; i < len; i++) dst[i] = src[i];
        pos += (size_t)len;
      }
    }
  });
  sp_off[off[r1] - j0] = (int32_t)total;
  return total;
}

}  // namespace voda
