// api.cu — the C ABI of include/voda_accel.h: contexts, host<->device staging, dispatch.
//
// Host-pointer entry points cut the batch into chunks of rounds and run them through a
// two-slot pipeline (two streams, two staging buffers): the H2D copy of chunk c+1 overlaps
// the kernel of chunk c and the D2H copy of chunk c-1.  Device-pointer entry points only
// enqueue kernels on the context's main stream.  Nothing here computes scheduling results on
// the CPU; without a CUDA device voda_ctx_create fails and no other entry point can be used.
#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <cstdlib>
#include <memory>

#include "common.cuh"
#include "hostpack.h"
#include "kernels.h"

using namespace voda;

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

struct voda_ctx {
  int device = 0;
  int sms = 0;
  cudaStream_t stream = nullptr;   // *_dev entry points
  cudaStream_t pipe[2] = {nullptr, nullptr};
  DevBuf scratch;                  // kernel scratch for the main stream
  DevBuf pipe_scratch[2];          // kernel scratch per pipeline slot
  DevBuf stage[2];                 // staged inputs + outputs per pipeline slot
  DevBuf curves;                   // the curve pool of a speedup_off batch (copied once per call)
  cudaEvent_t curves_ready = nullptr;
  void* hstage[2] = {nullptr, nullptr};   // pinned host staging per slot (packed speedup rows)
  size_t hstage_cap[2] = {0, 0};
  size_t direct_bytes = (size_t)16 << 10;  // VODA_DIRECT_BYTES: batches up to this size are read by the kernel
                                           // straight from the page-locked block (no copy engine at all)
  cudaEvent_t h2d_done[2] = {nullptr, nullptr};  // the slot's host staging has been consumed
  std::unique_ptr<HostPool> pool;  // host threads that pack rows (created on first use)
  bool pack_rows = false;          // VODA_PACK_ROWS=1 enables the FfDL row packing (measured slower: DESIGN.md)
  bool narrow_cols = true;         // VODA_NARROW_COLS=0: always ship the full-width speedup table
  int64_t launches = 0;
  int64_t h2d_bytes = 0, d2h_bytes = 0;  // bytes the last host-pointer call copied
  std::string err;
};

namespace {

thread_local std::string g_err_noctx;

int fail(voda_ctx* c, int code, const std::string& msg) {
  if (c) c->err = msg; else g_err_noctx = msg;
  return code;
}
int cuda_fail(voda_ctx* c, cudaError_t e, const char* what) {
  return fail(c, VODA_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(call)                                              \
  do {                                                        \
    cudaError_t e__ = (call);                                 \
    if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call); \
  } while (0)

int ensure(voda_ctx* ctx, DevBuf& b, size_t bytes, cudaStream_t user) {
  if (bytes <= b.cap) return VODA_OK;
  if (b.p) {
    CU(cudaStreamSynchronize(user));  // the old buffer may still be in use on that stream
    CU(cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
  }
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) {
    b.p = nullptr;
    return fail(ctx, VODA_E_NOMEM, std::string("cudaMalloc of scratch failed: ") + cudaGetErrorString(e));
  }
  b.cap = want;
  return VODA_OK;
}

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Host-pointer entry points enqueue copies and kernels that write into the CALLER's buffers.  If
// such a call has to return an error after its first enqueue, everything already in flight must
// land first: otherwise the caller may free or reuse memory that a D2H copy is still writing.
struct PipeGuard {
  voda_ctx* c;
  bool armed = true;
  explicit PipeGuard(voda_ctx* ctx) : c(ctx) {}
  ~PipeGuard() {
    if (!armed) return;
    cudaStreamSynchronize(c->pipe[0]);
    cudaStreamSynchronize(c->pipe[1]);
  }
};

// which arrays a policy reads (DESIGN.md "inputs per policy")
enum : unsigned {
  F_SUBMIT = 1, F_FIRST = 2, F_PRIO = 4, F_NUM = 8, F_MIN = 16, F_MAX = 32, F_REM = 64, F_SPEED = 128
};
unsigned algo_fields(int algo) {
  switch (algo) {
    case VODA_ALGO_FIFO:
    case VODA_ALGO_ELASTIC_FIFO: return F_SUBMIT | F_MIN | F_MAX;
    case VODA_ALGO_SRJF:
    case VODA_ALGO_ELASTIC_SRJF: return F_REM | F_MIN | F_MAX;
    case VODA_ALGO_TIRESIAS: return F_PRIO | F_FIRST | F_NUM | F_MIN | F_MAX;
    case VODA_ALGO_ELASTIC_TIRESIAS: return F_PRIO | F_FIRST | F_NUM | F_MIN | F_MAX | F_SPEED;
    case VODA_ALGO_FFDL_OPTIMIZER: return F_SUBMIT | F_MIN | F_MAX | F_SPEED;
    case VODA_ALGO_AFSL: return F_SUBMIT | F_MIN | F_MAX | F_REM | F_SPEED;
  }
  return 0;
}

int max_i32(const int32_t* v, int a, int b) {
  int m0 = INT32_MIN, m1 = INT32_MIN, m2 = INT32_MIN, m3 = INT32_MIN;
  int i = a;
  for (; i + 4 <= b; i += 4) {
    m0 = std::max(m0, v[i]);
    m1 = std::max(m1, v[i + 1]);
    m2 = std::max(m2, v[i + 2]);
    m3 = std::max(m3, v[i + 3]);
  }
  for (; i < b; i++) m0 = std::max(m0, v[i]);
  return std::max(std::max(m0, m1), std::max(m2, m3));
}

// Largest speedup-table index jobs [j0, j1) can make their policy read (inclusive), from the
// reference loops: FfDL g <= MaxNumProc (ffdl_optimizer.go:93-97); AFS-L S[r], S[r+1] with
// r <= MaxNumProc and len(jb, r_champion) with r_champion < the largest MaxNumProc
// (afsl.go:94-106), one more for the update that removes a job; ElasticTiresias S[min], S[min+1],
// S[NumProc+1], S[r+1] with r <= max(MaxNumProc, MinNumProc) (elastic_tiresias.go:58,80,96,139,147).
int speedup_index_cap(const voda_alloc_batch* b, int j0, int j1) {
  long cap = max_i32(b->max_proc, j0, j1);
  if (b->algo == VODA_ALGO_ELASTIC_TIRESIAS) {
    cap = std::max<long>(cap, max_i32(b->min_proc, j0, j1));
    cap = std::max<long>(cap, max_i32(b->num_proc, j0, j1));
  }
  cap = std::max<long>(cap, 1) + 2;
  return (int)std::min<long>(cap, INT32_MAX - 1);
}

int check_batch(voda_ctx* ctx, const voda_alloc_batch* b) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (!b) return fail(ctx, VODA_E_INVALID, "null batch");
  if (b->algo < 0 || b->algo >= VODA_ALGO_COUNT)
    return fail(ctx, VODA_E_INVALID, "unknown policy id (NewAlgorithmFactory: Not found)");
  if (b->rounds < 0) return fail(ctx, VODA_E_INVALID, "negative round count");
  if (b->rounds > 0 && (!b->job_off || !b->total_gpu)) return fail(ctx, VODA_E_INVALID, "job_off / total_gpu missing");
  const unsigned f = algo_fields(b->algo);
  if ((f & F_SUBMIT) && !b->submit_sec) return fail(ctx, VODA_E_INVALID, "policy needs submit_sec");
  if ((f & F_FIRST) && !b->first_sec) return fail(ctx, VODA_E_INVALID, "policy needs first_sec");
  if ((f & F_PRIO) && !b->priority) return fail(ctx, VODA_E_INVALID, "policy needs priority");
  if ((f & F_NUM) && !b->num_proc) return fail(ctx, VODA_E_INVALID, "policy needs num_proc");
  if ((f & F_MIN) && !b->min_proc) return fail(ctx, VODA_E_INVALID, "policy needs min_proc");
  if ((f & F_MAX) && !b->max_proc) return fail(ctx, VODA_E_INVALID, "policy needs max_proc");
  if ((f & F_REM) && !b->remaining_sec) return fail(ctx, VODA_E_INVALID, "policy needs remaining_sec");
  if ((f & F_SPEED) && (!b->speedup || b->speedup_stride < 1))
    return fail(ctx, VODA_E_INVALID, "policy needs the speedup table (stride >= 1)");
  if ((f & F_SPEED) && b->speedup_off && b->speedup_len < (int64_t)b->speedup_stride)
    return fail(ctx, VODA_E_INVALID, "speedup_off needs speedup_len (floats in the curve pool)");
  return VODA_OK;
}

AllocArgs make_args(const voda_alloc_batch* b) {
  AllocArgs a{};
  a.algo = b->algo;
  a.rounds = b->rounds;
  a.job_base = 0;
  a.job_off = b->job_off;
  a.total_gpu = b->total_gpu;
  a.submit_sec = b->submit_sec;
  a.submit_nsec = b->submit_nsec;
  a.first_sec = b->first_sec;
  a.first_nsec = b->first_nsec;
  a.priority = b->priority;
  a.num_proc = b->num_proc;
  a.min_proc = b->min_proc;
  a.max_proc = b->max_proc;
  a.remaining_sec = b->remaining_sec;
  a.speedup = b->speedup;
  a.sp_off = b->speedup_off;
  a.stride = b->speedup_stride;
  a.queue_num = b->queue_num > 0 ? b->queue_num : 2;                       // tiresias.go:24
  a.compaction = b->compaction_threshold >= 0 ? b->compaction_threshold : 10;  // elastic_tiresias.go:21
  a.max_jobs = b->max_jobs;
  a.max_total = b->max_total_gpu;
  a.max_gmax = b->max_max_proc > 0 ? b->max_max_proc : 0;
  return a;
}

// Shapes the kernels cannot take (a function of algo, max_jobs, max_total, stride only), checked
// before anything is enqueued.
int check_shape(voda_ctx* ctx, const AllocArgs& a) {
  switch (a.algo) {
    case VODA_ALGO_FFDL_OPTIMIZER:
      if (ffdl_smem_bytes(a) > 200 * 1024)
        return fail(ctx, VODA_E_UNSUPPORTED, "FfDL: total_gpu / jobs too large for one CTA's shared memory");
      return VODA_OK;
    case VODA_ALGO_FIFO:
    case VODA_ALGO_SRJF:
    case VODA_ALGO_ELASTIC_FIFO:
    case VODA_ALGO_ELASTIC_SRJF:
      if (fill_smem_bytes(a) > 200 * 1024)
        return fail(ctx, VODA_E_UNSUPPORTED, "too many jobs per round for one CTA's shared memory");
      return VODA_OK;
    case VODA_ALGO_TIRESIAS:
    case VODA_ALGO_ELASTIC_TIRESIAS:
      if (tiresias_smem_bytes(a, nullptr) > 220 * 1024)
        return fail(ctx, VODA_E_UNSUPPORTED, "too many jobs per round for one CTA's shared memory");
      return VODA_OK;
    case VODA_ALGO_AFSL:
      if (afsl_smem_bytes(a) > 220 * 1024)
        return fail(ctx, VODA_E_UNSUPPORTED, "too many jobs per round for one CTA's shared memory");
      return VODA_OK;
    default:
      return fail(ctx, VODA_E_UNSUPPORTED, "policy kernel not built into this library");
  }
}

// Enqueue the kernels of one policy over `a` on `stream`.
int dispatch_alloc(voda_ctx* ctx, const AllocArgs& a, DevBuf& scratch, cudaStream_t stream) {
  if (a.rounds == 0) return VODA_OK;
  int shape = check_shape(ctx, a);
  if (shape) return shape;
  switch (a.algo) {
    case VODA_ALGO_FFDL_OPTIMIZER: {
      // every round keeps its slab of stored DP rows until the backtrack kernel has read it: cap the scratch
      // at 16 GiB and walk larger batches in sub-batches (two launches each)
      const size_t per = (ffdl_scratch_per_round(a) + 255) & ~(size_t)255;
      const size_t budget = (size_t)16 << 30;
      int sub = (int)std::min<size_t>((size_t)a.rounds, std::max<size_t>(1, budget / per));
      int rc = ensure(ctx, scratch, per * (size_t)sub, stream);
      if (rc) return rc;
      for (int r0 = 0; r0 < a.rounds; r0 += sub) {
        const int count = std::min(sub, a.rounds - r0);
        CU(launch_ffdl(a, r0, count, ctx->sms, scratch.p, stream));
        ctx->launches += 2;
      }
      return VODA_OK;
    }
    case VODA_ALGO_FIFO:
    case VODA_ALGO_SRJF:
    case VODA_ALGO_ELASTIC_FIFO:
    case VODA_ALGO_ELASTIC_SRJF: {
      CU(launch_fill(a, ctx->sms, stream));
      ctx->launches += 1;
      return VODA_OK;
    }
    case VODA_ALGO_TIRESIAS:
    case VODA_ALGO_ELASTIC_TIRESIAS: {
      CU(launch_tiresias(a, ctx->sms, stream));
      ctx->launches += 1;
      return VODA_OK;
    }
    case VODA_ALGO_AFSL: {
      CU(launch_afsl(a, ctx->sms, stream));
      ctx->launches += 1;
      return VODA_OK;
    }
    default:
      return fail(ctx, VODA_E_UNSUPPORTED, "policy kernel not built into this library");
  }
}

}  // namespace

extern "C" {

const char* voda_version(void) { return "vodascheduler_b200 0.2 (sm_100a)"; }

size_t voda_struct_size(int32_t which) {
  switch (which) {
    case 0: return sizeof(voda_alloc_batch);
    case 1: return sizeof(voda_place_batch);
    case 2: return sizeof(voda_place_out);
    case 3: return sizeof(voda_curve_batch);
    case 4: return sizeof(voda_tick_batch);
    case 5: return sizeof(voda_replay_batch);
    case 6: return sizeof(voda_replay_out);
  }
  return 0;
}

int voda_ctx_create(int device, voda_ctx** out) {
  if (!out) return fail(nullptr, VODA_E_INVALID, "null out pointer");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(nullptr, VODA_E_CUDA,
                std::string("no CUDA device: this library has no CPU path (") +
                    (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") + ")");
  if (device < 0 || device >= count) return fail(nullptr, VODA_E_INVALID, "device index out of range");
  voda_ctx* ctx = new voda_ctx();
  ctx->device = device;
  auto bail = [&](cudaError_t err, const char* what) {
    int rc = cuda_fail(nullptr, err, what);
    delete ctx;
    return rc;
  };
  if ((e = cudaSetDevice(device)) != cudaSuccess) return bail(e, "cudaSetDevice");
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
  ctx->sms = prop.multiProcessorCount;
  if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
  for (int i = 0; i < 2; i++)
    if ((e = cudaStreamCreateWithFlags(&ctx->pipe[i], cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
  for (int i = 0; i < 2; i++)
    if ((e = cudaEventCreateWithFlags(&ctx->h2d_done[i], cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if ((e = cudaEventCreateWithFlags(&ctx->curves_ready, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
  if (const char* v = std::getenv("VODA_PACK_ROWS")) ctx->pack_rows = std::atoi(v) != 0;
  if (const char* v = std::getenv("VODA_NARROW_COLS")) ctx->narrow_cols = std::atoi(v) != 0;
  if (const char* v = std::getenv("VODA_DIRECT_BYTES")) ctx->direct_bytes = (size_t)std::max(0, std::atoi(v));
  *out = ctx;
  return VODA_OK;
}

void voda_ctx_destroy(voda_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  DevBuf* bufs[] = {&ctx->scratch, &ctx->pipe_scratch[0], &ctx->pipe_scratch[1], &ctx->stage[0], &ctx->stage[1],
                    &ctx->curves};
  if (ctx->curves_ready) cudaEventDestroy(ctx->curves_ready);
  for (DevBuf* b : bufs)
    if (b->p) cudaFree(b->p);
  for (int i = 0; i < 2; i++) {
    if (ctx->hstage[i]) cudaFreeHost(ctx->hstage[i]);
    if (ctx->h2d_done[i]) cudaEventDestroy(ctx->h2d_done[i]);
  }
  ctx->pool.reset();
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  for (int i = 0; i < 2; i++)
    if (ctx->pipe[i]) cudaStreamDestroy(ctx->pipe[i]);
  delete ctx;
}

void* voda_ctx_stream(voda_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int voda_ctx_sync(voda_ctx* ctx) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  CU(cudaSetDevice(ctx->device));
  CU(cudaStreamSynchronize(ctx->stream));
  CU(cudaStreamSynchronize(ctx->pipe[0]));
  CU(cudaStreamSynchronize(ctx->pipe[1]));
  return VODA_OK;
}

int64_t voda_ctx_launch_count(voda_ctx* ctx) { return ctx ? ctx->launches : 0; }

void voda_ctx_copied_bytes(voda_ctx* ctx, int64_t* h2d, int64_t* d2h) {
  if (h2d) *h2d = ctx ? ctx->h2d_bytes : 0;
  if (d2h) *d2h = ctx ? ctx->d2h_bytes : 0;
}

const char* voda_last_error(voda_ctx* ctx) { return ctx ? ctx->err.c_str() : g_err_noctx.c_str(); }

const char* voda_status_str(int32_t s) {
  switch (s) {
    case VODA_ST_OK: return "ok";
    case VODA_ST_NEGATIVE: return "Invalid GPU allocations: can't be negative";
    case VODA_ST_LESS_THAN_MIN: return "Invalid GPU allocations: less than job min gpu";
    case VODA_ST_EXCEEDED_MAX: return "Invalid GPU allocations: exceeded job max gpu";
    case VODA_ST_EXCEEDED_TOTAL: return "Invalid GPU allocations: exceeded total GPUs";
    case VODA_ST_INFEASIBLE: return "infeasible";
    case VODA_ST_NOT_FOUND: return "Not found";
    case VODA_ST_NONTERMINATING: return "reference would not terminate";
    case VODA_ST_BOUND: return "round exceeds the declared max_jobs / max_total_gpu";
    case VODA_ST_RUNTIME_PANIC: return "runtime error: makeslice: cap out of range";
  }
  return "unknown status";
}

static const char* kAlgoNames[VODA_ALGO_COUNT] = {"FIFO", "ElasticFIFO", "SRJF", "ElasticSRJF",
                                                  "Tiresias", "ElasticTiresias", "FfDLOptimizer", "AFS-L"};

int voda_algo_from_name(const char* name) {
  if (!name) return -1;
  for (int i = 0; i < VODA_ALGO_COUNT; i++)
    if (std::strcmp(name, kAlgoNames[i]) == 0) return i;
  return -1;
}
const char* voda_algo_name(int32_t algo) {
  return (algo >= 0 && algo < VODA_ALGO_COUNT) ? kAlgoNames[algo] : "";
}
int voda_algo_need_job_info(int32_t algo) {
  switch (algo) {
    case VODA_ALGO_SRJF:
    case VODA_ALGO_ELASTIC_SRJF:
    case VODA_ALGO_ELASTIC_TIRESIAS:
    case VODA_ALGO_FFDL_OPTIMIZER:
    case VODA_ALGO_AFSL: return 1;
  }
  return 0;
}

void* voda_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
  return p;
}
void voda_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

int voda_allocate_dev(voda_ctx* ctx, const voda_alloc_batch* b, int32_t* d_out_gpus, int32_t* d_out_status) {
  int rc = check_batch(ctx, b);
  if (rc) return rc;
  if (b->rounds == 0) return VODA_OK;
  if (!d_out_gpus || !d_out_status) return fail(ctx, VODA_E_INVALID, "null output");
  if (b->max_jobs <= 0 || b->max_total_gpu < 0)
    return fail(ctx, VODA_E_INVALID, "voda_allocate_dev needs max_jobs (and max_total_gpu) declared");
  CU(cudaSetDevice(ctx->device));
  AllocArgs a = make_args(b);
  a.out_gpus = d_out_gpus;
  a.out_status = d_out_status;
  return dispatch_alloc(ctx, a, ctx->scratch, ctx->stream);
}

int voda_allocate(voda_ctx* ctx, const voda_alloc_batch* b, int32_t* out_gpus, int32_t* out_status) {
  int rc = check_batch(ctx, b);
  if (rc) return rc;
  if (b->rounds == 0) return VODA_OK;
  if (!out_gpus || !out_status) return fail(ctx, VODA_E_INVALID, "null output");
  CU(cudaSetDevice(ctx->device));
  const int R = b->rounds;
  const int32_t* off = b->job_off;
  if (off[0] != 0) return fail(ctx, VODA_E_INVALID, "job_off[0] must be 0");
  int max_jobs = b->max_jobs, max_total = b->max_total_gpu;
  if (max_jobs <= 0 || max_total <= 0) {
    int mj = 0, mt = 0;
    for (int r = 0; r < R; r++) {
      int n = off[r + 1] - off[r];
      if (n < 0) return fail(ctx, VODA_E_INVALID, "job_off must ascend");
      mj = std::max(mj, n);
      mt = std::max(mt, b->total_gpu[r]);
    }
    if (max_jobs <= 0) max_jobs = std::max(mj, 1);
    if (max_total <= 0) max_total = mt;
  }
  const unsigned f = algo_fields(b->algo);
  const size_t W = (f & F_SPEED) ? (size_t)b->speedup_stride : 0;
  {  // refuse the shapes no kernel takes before anything is enqueued
    AllocArgs probe = make_args(b);
    probe.max_jobs = max_jobs;
    probe.max_total = max_total;
    rc = check_shape(ctx, probe);
    if (rc) return rc;
  }
  // bytes staged per job / per round
  size_t per_job = 4 /*out*/;
  if (f & F_SUBMIT) per_job += 8 + (b->submit_nsec ? 4 : 0);
  if (f & F_FIRST) per_job += 8 + (b->first_nsec ? 4 : 0);
  if (f & F_PRIO) per_job += 4;
  if (f & F_NUM) per_job += 4;
  if (f & F_MIN) per_job += 4;
  if (f & F_MAX) per_job += 4;
  if (f & F_REM) per_job += 4;
  per_job += (b->speedup_off && (f & F_SPEED)) ? 4 : 4 * W;  // a pool of curves: only the offset per job
  // chunking: a chunk must hold enough rounds to occupy every SM (one CTA per round, about a
  // dozen resident CTAs per SM) and at least ~16 MiB so the copies run at PCIe speed; several
  // chunks per batch let the H2D copy of one overlap the kernels of the previous one
  const size_t total_bytes = per_job * (size_t)off[R] + 12 * (size_t)R;
  const size_t bytes_per_round = std::max<size_t>(1, total_bytes / (size_t)R);
  size_t rounds_per_sz = std::max<size_t>((size_t)ctx->sms * 12, ((size_t)16 << 20) / bytes_per_round + 1);
  if (b->algo == VODA_ALGO_AFSL) {
    // an AFS-L round is one long dependent chain: a chunk that is not a whole number of waves of the
    // launch leaves SMs idle for a full round's time
    AllocArgs probe = make_args(b);
    probe.max_jobs = max_jobs;
    probe.max_total = max_total;
    const size_t wave = (size_t)afsl_wave_rounds(probe, ctx->sms);
    if (wave > 0) rounds_per_sz = std::max<size_t>(1, rounds_per_sz / wave) * wave;  // rounded down: the first chunk's copy is exposed
  }
  if (rounds_per_sz > (size_t)R) rounds_per_sz = (size_t)R;
  const int rounds_per = (int)rounds_per_sz;
  ctx->h2d_bytes = 0;
  ctx->d2h_bytes = 0;
  PipeGuard guard(ctx);  // any error return below first waits for what is already in flight
  // a pool of shared curves crosses PCIe once per call; the chunks only stage the jobs' offsets
  const bool pool = (f & F_SPEED) && b->speedup_off != nullptr;
  if (pool) {
    rc = ensure(ctx, ctx->curves, 4 * (size_t)b->speedup_len, ctx->pipe[0]);
    if (rc) return rc;
    CU(cudaStreamSynchronize(ctx->pipe[1]));  // nothing of an earlier call still reads the old pool
    CU(cudaMemcpyAsync(ctx->curves.p, b->speedup, 4 * (size_t)b->speedup_len, cudaMemcpyHostToDevice, ctx->pipe[0]));
    CU(cudaEventRecord(ctx->curves_ready, ctx->pipe[0]));
    CU(cudaStreamWaitEvent(ctx->pipe[1], ctx->curves_ready, 0));
    ctx->h2d_bytes += 4 * (int64_t)b->speedup_len;
  }
  for (int c = 0, r0 = 0; r0 < R; c++, r0 += rounds_per) {
    const int r1 = std::min(R, r0 + rounds_per);
    const int slot = c & 1;
    cudaStream_t s = ctx->pipe[slot];
    const int j0 = off[r0], j1 = off[r1];
    const size_t nj = (size_t)(j1 - j0), nr = (size_t)(r1 - r0);
    // carve the staging buffer
    size_t pos = 0;
    auto carve = [&](size_t bytes) { size_t at = pos; pos += align256(bytes ? bytes : 1); return at; };
    struct Copy { size_t at; const void* src; size_t bytes; };
    std::vector<Copy> copies;
    auto in = [&](const void* base, size_t elem, size_t first, size_t count) -> size_t {
      size_t at = carve(elem * count);
      if (count) copies.push_back({at, (const char*)base + elem * first, elem * count});
      return at;
    };
    const size_t o_off = in(off, 4, (size_t)r0, nr + 1);
    const size_t o_tot = in(b->total_gpu, 4, (size_t)r0, nr);
    size_t o_ss = 0, o_sn = 0, o_fs = 0, o_fn = 0, o_pr = 0, o_np = 0, o_mn = 0, o_mx = 0, o_rm = 0, o_sp = 0;
    if (f & F_SUBMIT) { o_ss = in(b->submit_sec, 8, j0, nj); if (b->submit_nsec) o_sn = in(b->submit_nsec, 4, j0, nj); }
    if (f & F_FIRST) { o_fs = in(b->first_sec, 8, j0, nj); if (b->first_nsec) o_fn = in(b->first_nsec, 4, j0, nj); }
    if (f & F_PRIO) o_pr = in(b->priority, 4, j0, nj);
    if (f & F_NUM) o_np = in(b->num_proc, 4, j0, nj);
    if (f & F_MIN) o_mn = in(b->min_proc, 4, j0, nj);
    if (f & F_MAX) o_mx = in(b->max_proc, 4, j0, nj);
    if (f & F_REM) o_rm = in(b->remaining_sec, 4, j0, nj);
    // FfDL, opt-in (VODA_PACK_ROWS=1): strip the dense rows down to the entries the DP can read
    // (hostpack.h).  It cuts the PCIe bytes 2.7x but the strided host gather reads almost as
    // many bytes from host DRAM as the DMA engine would have moved, more slowly: measured
    // 10.1 ms vs 8.4 ms per 10k rounds on the 16-core box, so it is off by default
    const bool packed = !pool && ctx->pack_rows && b->algo == VODA_ALGO_FFDL_OPTIMIZER && W >= 8 && nj > 0;
    size_t o_spoff = 0, packed_floats = 0, Wc = W;
    bool narrow = false;
    if (packed) {
      if (!ctx->pool) {
        int t = (int)std::thread::hardware_concurrency() / 2;
        if (const char* v = std::getenv("VODA_HOST_THREADS")) t = std::atoi(v);
        ctx->pool.reset(new HostPool(std::max(1, std::min(t, 8))));
      }
      const size_t need = 4 * (nj + 1) + 4 * W * nj + 256;
      if (ctx->hstage_cap[slot] < need) {
        CU(cudaStreamSynchronize(s));
        if (ctx->hstage[slot]) CU(cudaFreeHost(ctx->hstage[slot]));
        ctx->hstage[slot] = nullptr;
        ctx->hstage_cap[slot] = 0;
        CU(cudaHostAlloc(&ctx->hstage[slot], need + need / 8, cudaHostAllocMapped));
        ctx->hstage_cap[slot] = need + need / 8;
      } else {
        CU(cudaEventSynchronize(ctx->h2d_done[slot]));  // the DMA of chunk c-2 has read this slot
      }
      int32_t* h_off = (int32_t*)ctx->hstage[slot];
      float* h_sp = (float*)((char*)ctx->hstage[slot] + align256(4 * (nj + 1)));
      packed_floats = ffdl_pack_rows(*ctx->pool, off, b->total_gpu, b->max_proc, b->speedup, (int)W, r0, r1,
                                     h_off, h_sp, W * nj);
      o_spoff = carve(4 * (nj + 1));
      copies.push_back({o_spoff, h_off, 4 * (nj + 1)});
      o_sp = carve(4 * packed_floats);
      copies.push_back({o_sp, h_sp, 4 * packed_floats});
    } else if (pool) {
      o_spoff = in(b->speedup_off, 4, j0, nj);
    } else if (f & F_SPEED) {
      // Only the leading columns of the speedup table that the policy can index in this chunk
      // cross PCIe: index g never exceeds a bound set by the chunk's largest MaxNumProc (and, for
      // ElasticTiresias, MinNumProc / NumProc), so the device copy is a table of Wc <= W columns
      // (2-D DMA: row width 4*Wc out of a host pitch of 4*W).  A kernel reading g >= Wc would see
      // the Go map miss (0) — which is what the full table yields there only if g >= W, so Wc is
      // used only when the bound proves no such read happens.
      Wc = W;
      if (ctx->narrow_cols && nj > 0) {
        const size_t cap = (size_t)speedup_index_cap(b, j0, j1) + 1;
        if (cap * 4 <= W * 3) Wc = cap;  // not worth a strided DMA for less than a quarter saved
      }
      if (Wc == W) {
        o_sp = in(b->speedup, 4 * W, j0, nj);
      } else {
        o_sp = carve(4 * Wc * nj);
        narrow = true;
      }
    }
    const size_t o_og = carve(4 * nj), o_os = carve(4 * nr);
    rc = ensure(ctx, ctx->stage[slot], pos, s);
    if (rc) return rc;
    char* base = (char*)ctx->stage[slot].p;
    // A small batch — the reference's own use: ONE round per Schedule() call — is latency-bound: a
    // dozen separate copies cost more than the kernel.  Its arrays are gathered into one page-locked
    // block with the device layout and cross in ONE copy each way.
    const bool gathered = R == (int)nr && !packed && !narrow && !pool && pos <= ((size_t)256 << 10);
    bool direct = false;
    if (gathered) {
      if (ctx->hstage_cap[slot] < pos) {
        CU(cudaStreamSynchronize(s));
        if (ctx->hstage[slot]) CU(cudaFreeHost(ctx->hstage[slot]));
        ctx->hstage[slot] = nullptr;
        ctx->hstage_cap[slot] = 0;
        CU(cudaHostAlloc(&ctx->hstage[slot], (size_t)512 << 10, cudaHostAllocMapped));
        ctx->hstage_cap[slot] = (size_t)512 << 10;
      }
      char* hs = (char*)ctx->hstage[slot];
      for (const Copy& cp : copies) {
        std::memcpy(hs + cp.at, cp.src, cp.bytes);
        ctx->h2d_bytes += (int64_t)cp.bytes;
      }
      // The smallest batches (one 64-job round is 11 KB) skip the copy engine altogether: the kernel
      // reads the page-locked block over PCIe and writes its results into it, so the call is one
      // launch and one wait (64 jobs: E-Tiresias 35 -> 24 us, E-SRJF 34 -> 22 us per call).  Not for
      // AFS-L, whose kernel re-reads its curves from global memory (71 -> 93 us), nor above 16 KB
      // (160 jobs: slower by 10 %).
      direct = pos <= ctx->direct_bytes && b->algo != VODA_ALGO_AFSL;
      if (direct) {
        void* dv = nullptr;
        CU(cudaHostGetDevicePointer(&dv, hs, 0));
        base = (char*)dv;
      } else {
        CU(cudaMemcpyAsync(base, hs, o_og, cudaMemcpyHostToDevice, s));
      }
    } else {
      for (const Copy& cp : copies) {
        CU(cudaMemcpyAsync(base + cp.at, cp.src, cp.bytes, cudaMemcpyHostToDevice, s));
        ctx->h2d_bytes += (int64_t)cp.bytes;
      }
    }
    if (narrow) {
      CU(cudaMemcpy2DAsync(base + o_sp, 4 * Wc, b->speedup + (size_t)j0 * W, 4 * W, 4 * Wc, nj,
                           cudaMemcpyHostToDevice, s));
      ctx->h2d_bytes += (int64_t)(4 * Wc * nj);
    }
    if (packed) CU(cudaEventRecord(ctx->h2d_done[slot], s));
    AllocArgs a = make_args(b);
    if (narrow) a.stride = (int32_t)Wc;
    a.rounds = (int)nr;
    a.job_base = j0;
    a.max_jobs = max_jobs;
    a.max_total = max_total;
    // FfDL keeps a quarter of its DP rows when no job of the chunk can take more than 32 GPUs: the
    // host arrays are here, so the bound is the chunk's own
    if (b->algo == VODA_ALGO_FFDL_OPTIMIZER && a.max_gmax == 0 && nj > 0)
      a.max_gmax = std::max(1, max_i32(b->max_proc, j0, j1));
    a.job_off = (const int32_t*)(base + o_off);
    a.total_gpu = (const int32_t*)(base + o_tot);
    a.submit_sec = (f & F_SUBMIT) ? (const int64_t*)(base + o_ss) : nullptr;
    a.submit_nsec = ((f & F_SUBMIT) && b->submit_nsec) ? (const int32_t*)(base + o_sn) : nullptr;
    a.first_sec = (f & F_FIRST) ? (const int64_t*)(base + o_fs) : nullptr;
    a.first_nsec = ((f & F_FIRST) && b->first_nsec) ? (const int32_t*)(base + o_fn) : nullptr;
    a.priority = (f & F_PRIO) ? (const int32_t*)(base + o_pr) : nullptr;
    a.num_proc = (f & F_NUM) ? (const int32_t*)(base + o_np) : nullptr;
    a.min_proc = (f & F_MIN) ? (const int32_t*)(base + o_mn) : nullptr;
    a.max_proc = (f & F_MAX) ? (const int32_t*)(base + o_mx) : nullptr;
    a.remaining_sec = (f & F_REM) ? (const float*)(base + o_rm) : nullptr;
    a.speedup = pool ? (const float*)ctx->curves.p : (f & F_SPEED) ? (const float*)(base + o_sp) : nullptr;
    a.sp_off = (packed || pool) ? (const int32_t*)(base + o_spoff) : nullptr;
    a.out_gpus = (int32_t*)(base + o_og);
    a.out_status = (int32_t*)(base + o_os);
    rc = dispatch_alloc(ctx, a, ctx->pipe_scratch[slot], s);
    if (rc) return rc;
    ctx->d2h_bytes += (int64_t)(4 * nj + 4 * nr);
    if (gathered) {  // outputs come back as one block and are handed out after the wait
      char* hs = (char*)ctx->hstage[slot];
      if (!direct) CU(cudaMemcpyAsync(hs + o_og, base + o_og, pos - o_og, cudaMemcpyDeviceToHost, s));
      CU(cudaStreamSynchronize(s));
      if (nj) std::memcpy(out_gpus + j0, hs + o_og, 4 * nj);
      std::memcpy(out_status + r0, hs + o_os, 4 * nr);
      guard.armed = false;  // the only chunk, and its stream has been waited for
      return VODA_OK;
    }
    if (nj) CU(cudaMemcpyAsync(out_gpus + j0, base + o_og, 4 * nj, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(out_status + r0, base + o_os, 4 * nr, cudaMemcpyDeviceToHost, s));
  }
  CU(cudaStreamSynchronize(ctx->pipe[0]));
  CU(cudaStreamSynchronize(ctx->pipe[1]));
  guard.armed = false;
  return VODA_OK;
}

int voda_munkres_max_dev(voda_ctx* ctx, int32_t n, int32_t batch, const int64_t* d_A, int32_t* d_out,
                         int32_t* d_status) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (n < 0 || batch < 0) return fail(ctx, VODA_E_INVALID, "negative size");
  if (n > 128) return fail(ctx, VODA_E_UNSUPPORTED, "munkres: n > 128 does not fit one CTA's shared memory");
  if (batch == 0) return VODA_OK;
  if (!d_status || (n > 0 && (!d_A || !d_out))) return fail(ctx, VODA_E_INVALID, "null buffer");
  CU(cudaSetDevice(ctx->device));
  if (n == 0) {  // NewMatrix(0): nothing to assign
    CU(cudaMemsetAsync(d_status, 0, 4 * (size_t)batch, ctx->stream));
    return VODA_OK;
  }
  int rc = ensure(ctx, ctx->scratch, munkres_scratch_bytes(n, batch, ctx->sms), ctx->stream);
  if (rc) return rc;
  int launched = 0;
  CU(launch_munkres(n, batch, d_A, d_out, d_status, ctx->scratch.p, ctx->sms, ctx->stream, &launched));
  ctx->launches += launched;
  return VODA_OK;
}

int voda_munkres_max(voda_ctx* ctx, int32_t n, int32_t batch, const int64_t* A, int32_t* out, int32_t* status) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (n < 0 || batch < 0) return fail(ctx, VODA_E_INVALID, "negative size");
  if (n > 128) return fail(ctx, VODA_E_UNSUPPORTED, "munkres: n > 128 does not fit one CTA's shared memory");
  if (batch == 0) return VODA_OK;
  if (!status || (n > 0 && (!A || !out))) return fail(ctx, VODA_E_INVALID, "null buffer");
  if (n == 0) {
    for (int b = 0; b < batch; b++) status[b] = VODA_ST_OK;
    return VODA_OK;
  }
  CU(cudaSetDevice(ctx->device));
  const size_t per = (size_t)n * n * 8;
  ctx->h2d_bytes = (int64_t)(per * (size_t)batch);
  ctx->d2h_bytes = (int64_t)(4 * (size_t)batch * ((size_t)n + 1));
  PipeGuard guard(ctx);
  // chunks big enough to occupy every SM (32 one-warp CTAs per SM), at least ~8 MiB
  size_t per_chunk = std::max<size_t>((size_t)ctx->sms * 32, (8u << 20) / per + 1);
  if (per_chunk > (size_t)batch) per_chunk = (size_t)batch;
  for (size_t c = 0, b0 = 0; b0 < (size_t)batch; c++, b0 += per_chunk) {
    const size_t nb = std::min(per_chunk, (size_t)batch - b0);
    const int slot = (int)(c & 1);
    cudaStream_t s = ctx->pipe[slot];
    const size_t o_A = 0, o_out = align256(per * nb), o_st = o_out + align256(4 * nb * n),
                 o_need = o_st + align256(4 * nb), total = o_need + munkres_scratch_bytes(n, (int)nb, ctx->sms);
    int rc = ensure(ctx, ctx->stage[slot], total, s);
    if (rc) return rc;
    char* base = (char*)ctx->stage[slot].p;
    CU(cudaMemcpyAsync(base + o_A, A + b0 * (size_t)n * n, per * nb, cudaMemcpyHostToDevice, s));
    int launched = 0;
    CU(launch_munkres(n, (int)nb, (const int64_t*)(base + o_A), (int32_t*)(base + o_out), (int32_t*)(base + o_st),
                      base + o_need, ctx->sms, s, &launched));
    ctx->launches += launched;
    CU(cudaMemcpyAsync(out + b0 * n, base + o_out, 4 * nb * n, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(status + b0, base + o_st, 4 * nb, cudaMemcpyDeviceToHost, s));
  }
  CU(cudaStreamSynchronize(ctx->pipe[0]));
  CU(cudaStreamSynchronize(ctx->pipe[1]));
  guard.armed = false;
  return VODA_OK;
}

static int check_place(voda_ctx* ctx, const voda_place_batch* b, const voda_place_out* o) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (!b || !o) return fail(ctx, VODA_E_INVALID, "null batch / out");
  if (b->rounds < 0) return fail(ctx, VODA_E_INVALID, "negative round count");
  if (b->rounds > 0 && (!b->node_off || !b->req_off || !b->prev_off || !o->status || !o->lay_cnt || !o->score || !o->cross))
    return fail(ctx, VODA_E_INVALID, "null array");
  return VODA_OK;
}

static PlaceArgs make_place_args(const voda_place_batch* b, const voda_place_out* o) {
  PlaceArgs a{};
  a.rounds = b->rounds;
  a.node_off = b->node_off; a.node_slots = b->node_slots; a.req_off = b->req_off; a.req_workers = b->req_workers;
  a.prev_off = b->prev_off; a.prev_job = b->prev_job; a.prev_workers = b->prev_workers;
  a.max_nodes = b->max_nodes; a.max_requests = b->max_requests;
  const bool pods = b->pods_off && o->migrated_workers && o->migrated_launchers;
  a.pods_off = pods ? b->pods_off : nullptr; a.pods_node = pods ? b->pods_node : nullptr;
  a.pods_slots = pods ? b->pods_slots : nullptr;
  a.migrated_workers = pods ? o->migrated_workers : nullptr;
  a.migrated_launchers = pods ? o->migrated_launchers : nullptr;
  a.phys_of_logical = o->phys_of_logical; a.free_slots = o->free_slots; a.lay_cnt = o->lay_cnt;
  a.lay_node = o->lay_node; a.lay_job = o->lay_job; a.lay_workers = o->lay_workers;
  a.score = reinterpret_cast<long long*>(o->score); a.cross = o->cross; a.status = o->status;
  return a;
}

static int dispatch_place(voda_ctx* ctx, const PlaceArgs& a, DevBuf& scratch, cudaStream_t stream, int first = 0,
                          int last = -1) {
  if (last < 0) last = a.rounds;
  if (last <= first) return VODA_OK;
  if (a.max_nodes > 128) return fail(ctx, VODA_E_UNSUPPORTED, "placement: more than 128 nodes per round");
  if (place_smem_bytes(a) > 220 * 1024)
    return fail(ctx, VODA_E_UNSUPPORTED, "placement: too many requests per round for one CTA's shared memory");
  int rc = ensure(ctx, scratch, place_scratch_bytes(a, ctx->sms, last - first), stream);
  if (rc) return rc;
  int launched = 0;
  CU(launch_place(a, ctx->sms, scratch.p, stream, &launched, first, last));
  ctx->launches += launched;
  return VODA_OK;
}

int voda_place_dev(voda_ctx* ctx, const voda_place_batch* b, const voda_place_out* o) {
  int rc = check_place(ctx, b, o);
  if (rc) return rc;
  if (b->max_nodes <= 0 || b->max_requests < 0)
    return fail(ctx, VODA_E_INVALID, "voda_place_dev needs max_nodes / max_requests declared");
  CU(cudaSetDevice(ctx->device));
  PlaceArgs a = make_place_args(b, o);
  if (a.max_requests < 1) a.max_requests = 1;
  return dispatch_place(ctx, a, ctx->scratch, ctx->stream);
}

int voda_place(voda_ctx* ctx, const voda_place_batch* b, const voda_place_out* o) {
  int rc = check_place(ctx, b, o);
  if (rc) return rc;
  const int R = b->rounds;
  if (R == 0) return VODA_OK;
  CU(cudaSetDevice(ctx->device));
  const int nodes = b->node_off[R], reqs = b->req_off[R], prevs = b->prev_off[nodes];
  int max_nodes = b->max_nodes, max_req = b->max_requests;
  if (max_nodes <= 0 || max_req <= 0) {
    int mn = 0, mr = 0;
    for (int r = 0; r < R; r++) {
      mn = std::max(mn, b->node_off[r + 1] - b->node_off[r]);
      mr = std::max(mr, b->req_off[r + 1] - b->req_off[r]);
    }
    if (max_nodes <= 0) max_nodes = std::max(mn, 1);
    if (max_req <= 0) max_req = std::max(mr, 1);
  }
  // placement batches are small (a few KB per round): one staged copy, one kernel
  cudaStream_t s = ctx->pipe[0];
  PipeGuard guard(ctx);  // an error return below first waits for what is already in flight
  size_t pos = 0;
  auto carve = [&](size_t bytes) { size_t at = pos; pos += align256(bytes ? bytes : 1); return at; };
  const size_t cap = (size_t)nodes + (size_t)reqs;
  const size_t i_no = carve(4 * (size_t)(R + 1)), i_ns = carve(4 * (size_t)nodes), i_ro = carve(4 * (size_t)(R + 1)),
               i_rw = carve(4 * (size_t)reqs), i_po = carve(4 * (size_t)(nodes + 1)), i_pj = carve(4 * (size_t)prevs),
               i_pw = carve(4 * (size_t)prevs);
  const bool pods = b->pods_off && o->migrated_workers && o->migrated_launchers;
  const int podRuns = pods ? b->pods_off[reqs] : 0;
  if (pods && podRuns > 0 && (!b->pods_node || !b->pods_slots)) return fail(ctx, VODA_E_INVALID, "pods_node / pods_slots missing");
  const size_t i_qo = carve(pods ? 4 * (size_t)(reqs + 1) : 0), i_qn = carve(4 * (size_t)podRuns), i_qs = carve(4 * (size_t)podRuns);
  const size_t o_mw = carve(pods ? 4 * (size_t)R : 0), o_ml = carve(pods ? 4 * (size_t)R : 0);
  const size_t o_ph = carve(4 * (size_t)nodes), o_fr = carve(4 * (size_t)nodes), o_lc = carve(4 * (size_t)R),
               o_ln = carve(4 * cap), o_lj = carve(4 * cap), o_lw = carve(4 * cap), o_sc = carve(8 * (size_t)R),
               o_cr = carve(4 * (size_t)R), o_st = carve(4 * (size_t)R);
  rc = ensure(ctx, ctx->stage[0], pos, s);
  if (rc) return rc;
  char* base = (char*)ctx->stage[0].p;
  // a small batch (one Place() per call) is latency-bound: its arrays are gathered into one
  // page-locked block and cross in one copy each way instead of seven in and nine out
  const bool gathered = pos <= ((size_t)256 << 10);
  if (gathered && ctx->hstage_cap[0] < pos) {
    CU(cudaStreamSynchronize(s));
    if (ctx->hstage[0]) CU(cudaFreeHost(ctx->hstage[0]));
    ctx->hstage[0] = nullptr;
    ctx->hstage_cap[0] = 0;
    CU(cudaHostAlloc(&ctx->hstage[0], (size_t)512 << 10, cudaHostAllocMapped));
    ctx->hstage_cap[0] = (size_t)512 << 10;
  }
  char* hs = (char*)ctx->hstage[0];
  // chunks of at least two waves of the solve kernel, a quarter of the batch at most
  int parts = 4;
  if (const char* v = std::getenv("VODA_PLACE_CHUNKS")) parts = std::max(1, std::atoi(v));
  const int chunkRounds = std::max(2 * 10 * ctx->sms, (R + parts - 1) / parts);
  const bool chunked = !gathered && !pods && R >= 2 * chunkRounds;
  ctx->h2d_bytes = 0;
  ctx->d2h_bytes = 0;
  auto h2d = [&](size_t at, const void* src, size_t bytes) -> cudaError_t {
    if (!bytes) return cudaSuccess;
    ctx->h2d_bytes += (int64_t)bytes;
    if (gathered) {
      std::memcpy(hs + at, src, bytes);
      return cudaSuccess;
    }
    return cudaMemcpyAsync(base + at, src, bytes, cudaMemcpyHostToDevice, s);
  };
  CU(h2d(i_no, b->node_off, 4 * (size_t)(R + 1)));
  CU(h2d(i_ro, b->req_off, 4 * (size_t)(R + 1)));
  CU(h2d(i_po, b->prev_off, 4 * (size_t)(nodes + 1)));
  if (!chunked) {
    CU(h2d(i_ns, b->node_slots, 4 * (size_t)nodes));
    CU(h2d(i_rw, b->req_workers, 4 * (size_t)reqs));
    CU(h2d(i_pj, b->prev_job, 4 * (size_t)prevs));
    CU(h2d(i_pw, b->prev_workers, 4 * (size_t)prevs));
  }
  if (pods) {
    CU(h2d(i_qo, b->pods_off, 4 * (size_t)(reqs + 1)));
    CU(h2d(i_qn, b->pods_node, 4 * (size_t)podRuns));
    CU(h2d(i_qs, b->pods_slots, 4 * (size_t)podRuns));
  }
  PlaceArgs a{};
  a.rounds = R;
  a.node_off = (const int32_t*)(base + i_no); a.node_slots = (const int32_t*)(base + i_ns);
  a.req_off = (const int32_t*)(base + i_ro); a.req_workers = (const int32_t*)(base + i_rw);
  a.prev_off = (const int32_t*)(base + i_po); a.prev_job = (const int32_t*)(base + i_pj);
  a.prev_workers = (const int32_t*)(base + i_pw);
  a.max_nodes = max_nodes; a.max_requests = max_req;
  if (pods) {
    a.pods_off = (const int32_t*)(base + i_qo); a.pods_node = (const int32_t*)(base + i_qn);
    a.pods_slots = (const int32_t*)(base + i_qs);
    a.migrated_workers = (int32_t*)(base + o_mw); a.migrated_launchers = (int32_t*)(base + o_ml);
  }
  a.phys_of_logical = (int32_t*)(base + o_ph); a.free_slots = (int32_t*)(base + o_fr); a.lay_cnt = (int32_t*)(base + o_lc);
  a.lay_node = (int32_t*)(base + o_ln); a.lay_job = (int32_t*)(base + o_lj); a.lay_workers = (int32_t*)(base + o_lw);
  a.score = (long long*)(base + o_sc); a.cross = (int32_t*)(base + o_cr); a.status = (int32_t*)(base + o_st);
  if (chunked) {
    // Large batches: the three offset arrays have crossed (above, on stream 0); the data of every chunk
    // of rounds goes to its place in the same device layout on one of two streams, so that the copies of
    // one chunk run under the kernels of the other.
    CU(cudaEventRecord(ctx->h2d_done[0], ctx->pipe[0]));
    CU(cudaStreamWaitEvent(ctx->pipe[1], ctx->h2d_done[0], 0));
    int c = 0;
    for (int r0 = 0; r0 < R; r0 += chunkRounds, c++) {
      const int r1 = std::min(R, r0 + chunkRounds);
      cudaStream_t cs = ctx->pipe[c & 1];
      const size_t n0 = (size_t)b->node_off[r0], n1 = (size_t)b->node_off[r1], q0 = (size_t)b->req_off[r0], q1 = (size_t)b->req_off[r1];
      const size_t p0 = (size_t)b->prev_off[n0], p1 = (size_t)b->prev_off[n1];
      auto in = [&](size_t at, const int32_t* src, size_t i0, size_t i1) -> cudaError_t {
        if (i1 <= i0) return cudaSuccess;
        ctx->h2d_bytes += (int64_t)(4 * (i1 - i0));
        return cudaMemcpyAsync(base + at + 4 * i0, src + i0, 4 * (i1 - i0), cudaMemcpyHostToDevice, cs);
      };
      CU(in(i_ns, b->node_slots, n0, n1));
      CU(in(i_rw, b->req_workers, q0, q1));
      CU(in(i_pj, b->prev_job, p0, p1));
      CU(in(i_pw, b->prev_workers, p0, p1));
      rc = dispatch_place(ctx, a, ctx->pipe_scratch[c & 1], cs, r0, r1);
      if (rc) return rc;
      auto out = [&](void* dst, size_t at, size_t width, size_t i0, size_t i1) -> cudaError_t {
        if (!dst || i1 <= i0) return cudaSuccess;
        ctx->d2h_bytes += (int64_t)(width * (i1 - i0));
        return cudaMemcpyAsync((char*)dst + width * i0, base + at + width * i0, width * (i1 - i0), cudaMemcpyDeviceToHost, cs);
      };
      CU(out(o->phys_of_logical, o_ph, 4, n0, n1));
      CU(out(o->free_slots, o_fr, 4, n0, n1));
      CU(out(o->lay_cnt, o_lc, 4, (size_t)r0, (size_t)r1));
      CU(out(o->lay_node, o_ln, 4, q0 + n0, q1 + n1));
      CU(out(o->lay_job, o_lj, 4, q0 + n0, q1 + n1));
      CU(out(o->lay_workers, o_lw, 4, q0 + n0, q1 + n1));
      CU(out(o->score, o_sc, 8, (size_t)r0, (size_t)r1));
      CU(out(o->cross, o_cr, 4, (size_t)r0, (size_t)r1));
      CU(out(o->status, o_st, 4, (size_t)r0, (size_t)r1));
    }
    CU(cudaStreamSynchronize(ctx->pipe[0]));
    CU(cudaStreamSynchronize(ctx->pipe[1]));
    guard.armed = false;
    return VODA_OK;
  }
  if (gathered) CU(cudaMemcpyAsync(base, hs, o_mw, cudaMemcpyHostToDevice, s));  // every input, one copy
  rc = dispatch_place(ctx, a, ctx->pipe_scratch[0], s);
  if (rc) return rc;
  if (gathered) {  // every output, one copy, handed out after the wait
    CU(cudaMemcpyAsync(hs + o_mw, base + o_mw, pos - o_mw, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
  }
  auto d2h = [&](void* dst, size_t at, size_t bytes) -> cudaError_t {
    if (!bytes || !dst) return cudaSuccess;
    ctx->d2h_bytes += (int64_t)bytes;
    if (gathered) {
      std::memcpy(dst, hs + at, bytes);
      return cudaSuccess;
    }
    return cudaMemcpyAsync(dst, base + at, bytes, cudaMemcpyDeviceToHost, s);
  };
  CU(d2h(o->phys_of_logical, o_ph, 4 * (size_t)nodes));
  CU(d2h(o->free_slots, o_fr, 4 * (size_t)nodes));
  CU(d2h(o->lay_cnt, o_lc, 4 * (size_t)R));
  CU(d2h(o->lay_node, o_ln, 4 * cap));
  CU(d2h(o->lay_job, o_lj, 4 * cap));
  CU(d2h(o->lay_workers, o_lw, 4 * cap));
  CU(d2h(o->score, o_sc, 8 * (size_t)R));
  CU(d2h(o->cross, o_cr, 4 * (size_t)R));
  CU(d2h(o->status, o_st, 4 * (size_t)R));
  if (pods) {
    CU(d2h(o->migrated_workers, o_mw, 4 * (size_t)R));
    CU(d2h(o->migrated_launchers, o_ml, 4 * (size_t)R));
  }
  CU(cudaStreamSynchronize(s));
  guard.armed = false;
  return VODA_OK;
}

static int check_curves(voda_ctx* ctx, const voda_curve_batch* b, const void* o1, const void* o2, const void* o3,
                        const void* st) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (!b) return fail(ctx, VODA_E_INVALID, "null batch");
  if (b->jobs < 0 || b->width < 2) return fail(ctx, VODA_E_INVALID, "curves: jobs >= 0 and width >= 2 required");
  if (b->jobs > 0 && (!b->rec_off || !o1 || !o2 || !o3 || !st)) return fail(ctx, VODA_E_INVALID, "null buffer");
  return VODA_OK;
}

int voda_build_curves_dev(voda_ctx* ctx, const voda_curve_batch* b, double* d_time, double* d_speedup,
                          double* d_eff, int32_t* d_status) {
  int rc = check_curves(ctx, b, d_time, d_speedup, d_eff, d_status);
  if (rc || b->jobs == 0) return rc;
  CU(cudaSetDevice(ctx->device));
  CU(launch_curves(b->jobs, b->rec_off, b->workers, b->epoch_time_sec, b->width, d_time, d_speedup, d_eff, d_status,
                   ctx->stream));
  ctx->launches += 1;
  return VODA_OK;
}

int voda_build_curves(voda_ctx* ctx, const voda_curve_batch* b, double* out_time, double* out_speedup,
                      double* out_eff, int32_t* out_status) {
  int rc = check_curves(ctx, b, out_time, out_speedup, out_eff, out_status);
  if (rc || b->jobs == 0) return rc;
  if (b->rec_off[0] != 0) return fail(ctx, VODA_E_INVALID, "rec_off[0] must be 0");
  const size_t J = (size_t)b->jobs, W = (size_t)b->width;
  const size_t recs = (size_t)b->rec_off[J];
  if (recs > 0 && (!b->workers || !b->epoch_time_sec)) return fail(ctx, VODA_E_INVALID, "null record arrays");
  CU(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->pipe[0];
  PipeGuard guard(ctx);  // an error return below first waits for what is already in flight
  const size_t o_off = 0, o_w = align256(4 * (J + 1)), o_t = o_w + align256(4 * recs), o_et = o_t + align256(8 * recs),
               o_sp = o_et + align256(8 * J * W), o_ef = o_sp + align256(8 * J * W), o_st = o_ef + align256(8 * J * W),
               total = o_st + align256(4 * J);
  rc = ensure(ctx, ctx->stage[0], total, s);
  if (rc) return rc;
  char* base = (char*)ctx->stage[0].p;
  CU(cudaMemcpyAsync(base + o_off, b->rec_off, 4 * (J + 1), cudaMemcpyHostToDevice, s));
  if (recs) {
    CU(cudaMemcpyAsync(base + o_w, b->workers, 4 * recs, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(base + o_t, b->epoch_time_sec, 8 * recs, cudaMemcpyHostToDevice, s));
  }
  CU(launch_curves(b->jobs, (const int32_t*)(base + o_off), (const int32_t*)(base + o_w), (const double*)(base + o_t),
                   b->width, (double*)(base + o_et), (double*)(base + o_sp), (double*)(base + o_ef),
                   (int32_t*)(base + o_st), s));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(out_time, base + o_et, 8 * J * W, cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(out_speedup, base + o_sp, 8 * J * W, cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(out_eff, base + o_ef, 8 * J * W, cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(out_status, base + o_st, 4 * J, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  guard.armed = false;
  return VODA_OK;
}

static int check_tick(voda_ctx* ctx, const voda_tick_batch* b) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (!b) return fail(ctx, VODA_E_INVALID, "null batch");
  if (b->jobs < 0) return fail(ctx, VODA_E_INVALID, "negative job count");
  if (b->queue_num < 0 || b->queue_num > VODA_TICK_MAX_QUEUES)
    return fail(ctx, VODA_E_UNSUPPORTED, "tick: queue_num outside [0, VODA_TICK_MAX_QUEUES]");
  if (b->jobs > 0 && (!b->status || !b->num_gpu || !b->priority || !b->last_update_ns || !b->running_ns || !b->gpu_ns ||
                      !b->total_ns || !b->waiting_ns || !b->last_running_ns || !b->last_gpu_ns || !b->last_waiting_ns ||
                      !b->changed))
    return fail(ctx, VODA_E_INVALID, "tick: null array");
  return VODA_OK;
}

static TickArgs tick_args(const voda_tick_batch* b) {
  TickArgs a{};
  a.jobs = b->jobs;
  a.tiresias = b->tiresias;
  a.queue_num = b->queue_num;
  a.promote_knob = b->promote_knob;
  a.now_ns = b->now_ns;
  for (int q = 0; q < VODA_TICK_MAX_QUEUES; q++) a.thr[q] = b->thresholds_sec[q];
  a.status = b->status;
  a.num_gpu = b->num_gpu;
  a.priority = b->priority;
  a.last_update_ns = (long long*)b->last_update_ns;
  a.running_ns = (long long*)b->running_ns;
  a.gpu_ns = (long long*)b->gpu_ns;
  a.total_ns = (long long*)b->total_ns;
  a.waiting_ns = (long long*)b->waiting_ns;
  a.last_running_ns = (long long*)b->last_running_ns;
  a.last_gpu_ns = (long long*)b->last_gpu_ns;
  a.last_waiting_ns = (long long*)b->last_waiting_ns;
  a.changed = b->changed;
  return a;
}

int voda_tick_dev(voda_ctx* ctx, const voda_tick_batch* b) {
  int rc = check_tick(ctx, b);
  if (rc || b->jobs == 0) return rc;
  CU(cudaSetDevice(ctx->device));
  CU(launch_tick(tick_args(b), ctx->stream));
  ctx->launches += 1;
  return VODA_OK;
}

int voda_tick(voda_ctx* ctx, const voda_tick_batch* b) {
  int rc = check_tick(ctx, b);
  if (rc || b->jobs == 0) return rc;
  CU(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->pipe[0];
  PipeGuard guard(ctx);  // an error return below first waits for what is already in flight
  const size_t J = (size_t)b->jobs, a4 = align256(4 * J), a8 = align256(8 * J);
  rc = ensure(ctx, ctx->stage[0], 4 * a4 + 8 * a8, s);
  if (rc) return rc;
  char* base = (char*)ctx->stage[0].p;
  int32_t* d4[4];
  for (int i = 0; i < 4; i++) d4[i] = (int32_t*)(base + i * a4);
  int64_t* d8[8];
  for (int i = 0; i < 8; i++) d8[i] = (int64_t*)(base + 4 * a4 + i * a8);
  int64_t* h8[8] = {b->last_update_ns, b->running_ns, b->gpu_ns, b->total_ns, b->waiting_ns, b->last_running_ns,
                    b->last_gpu_ns, b->last_waiting_ns};
  CU(cudaMemcpyAsync(d4[0], b->status, 4 * J, cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(d4[1], b->num_gpu, 4 * J, cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(d4[2], b->priority, 4 * J, cudaMemcpyHostToDevice, s));
  for (int i = 0; i < 8; i++) CU(cudaMemcpyAsync(d8[i], h8[i], 8 * J, cudaMemcpyHostToDevice, s));
  voda_tick_batch d = *b;
  d.status = d4[0];
  d.num_gpu = d4[1];
  d.priority = d4[2];
  d.changed = d4[3];
  d.last_update_ns = d8[0];
  d.running_ns = d8[1];
  d.gpu_ns = d8[2];
  d.total_ns = d8[3];
  d.waiting_ns = d8[4];
  d.last_running_ns = d8[5];
  d.last_gpu_ns = d8[6];
  d.last_waiting_ns = d8[7];
  CU(launch_tick(tick_args(&d), s));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(b->priority, d4[2], 4 * J, cudaMemcpyDeviceToHost, s));
  CU(cudaMemcpyAsync(b->changed, d4[3], 4 * J, cudaMemcpyDeviceToHost, s));
  for (int i = 0; i < 8; i++) CU(cudaMemcpyAsync(h8[i], d8[i], 8 * J, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  guard.armed = false;
  return VODA_OK;
}

int voda_compare_results_dev(voda_ctx* ctx, int32_t jobs, const int32_t* d_old, const int32_t* d_new,
                             const int32_t* d_status, int32_t* d_action) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (jobs < 0) return fail(ctx, VODA_E_INVALID, "negative job count");
  if (jobs == 0) return VODA_OK;
  if (!d_old || !d_new || !d_status || !d_action) return fail(ctx, VODA_E_INVALID, "null buffer");
  CU(cudaSetDevice(ctx->device));
  CU(launch_compare(jobs, d_old, d_new, d_status, d_action, ctx->stream));
  ctx->launches += 1;
  return VODA_OK;
}

int voda_compare_results(voda_ctx* ctx, int32_t jobs, const int32_t* old_gpu, const int32_t* new_gpu,
                         const int32_t* status, int32_t* action) {
  if (!ctx) return fail(nullptr, VODA_E_INVALID, "null context");
  if (jobs < 0) return fail(ctx, VODA_E_INVALID, "negative job count");
  if (jobs == 0) return VODA_OK;
  if (!old_gpu || !new_gpu || !status || !action) return fail(ctx, VODA_E_INVALID, "null buffer");
  CU(cudaSetDevice(ctx->device));
  cudaStream_t s = ctx->pipe[0];
  PipeGuard guard(ctx);  // an error return below first waits for what is already in flight
  const size_t J = (size_t)jobs, a4 = align256(4 * J);
  int rc = ensure(ctx, ctx->stage[0], 4 * a4, s);
  if (rc) return rc;
  char* base = (char*)ctx->stage[0].p;
  CU(cudaMemcpyAsync(base, old_gpu, 4 * J, cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(base + a4, new_gpu, 4 * J, cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(base + 2 * a4, status, 4 * J, cudaMemcpyHostToDevice, s));
  CU(launch_compare(jobs, (const int32_t*)base, (const int32_t*)(base + a4), (const int32_t*)(base + 2 * a4),
                    (int32_t*)(base + 3 * a4), s));
  ctx->launches += 1;
  CU(cudaMemcpyAsync(action, base + 3 * a4, 4 * J, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  guard.armed = false;
  return VODA_OK;
}

}  // extern "C"
