/*
 * voda_accel.h — C ABI of the B200 engine for Voda's allocation + placement hot path.
 *
 * Every entry point is a plain `extern "C"` function over pointers and sizes (no torch or
 * C++ types), so the reference's Go services can bind it through cgo (INTEGRATION.md shows
 * the shim).  Each declaration cites the reference interface it stands in for; paths are
 * relative to the reference tree (heyfey/vodascheduler).
 *
 * Conventions
 *   - The caller owns every buffer.  `voda_*` (no suffix) entry points take HOST pointers
 *     and do the host<->device copies themselves (pinned memory from voda_host_alloc makes
 *     those copies asynchronous); `voda_*_dev` entry points take DEVICE pointers and only
 *     enqueue kernels on the context's stream.
 *   - Return value = infrastructure outcome (VODA_OK or a negative VODA_E_*; text from
 *     voda_last_error).  Per-round reference outcome goes to out_status[round]
 *     (VODA_ST_*): the reference reports failure by panicking (pkg/algorithm/utils.go:29-40,
 *     ffdl_optimizer.go:109-114), and voda_status_str() returns the same message text.
 *   - A voda_ctx owns one device, its streams and its scratch memory.  It is NOT thread
 *     safe: use one context per goroutine / OS thread.  There is no global mutable state.
 *   - There is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef VODA_ACCEL_H
#define VODA_ACCEL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- infrastructure return codes -------------------------------------------------- */
#define VODA_OK             0
#define VODA_E_INVALID     -1   /* bad argument (NULL array the policy needs, negative size ...) */
#define VODA_E_CUDA        -2   /* a CUDA runtime call failed; see voda_last_error            */
#define VODA_E_NOMEM       -3   /* scratch / staging allocation failed                         */
#define VODA_E_UNSUPPORTED -4   /* shape outside what the kernels handle       */
#define VODA_E_TOO_LARGE   -5   /* request exceeds a fixed capacity (no-allocation entry point)*/

/* ---- per-round status: the reference's panic messages ------------------------------ */
#define VODA_ST_OK              0
#define VODA_ST_NEGATIVE        1  /* "Invalid GPU allocations: can't be negative"      utils.go:29-31 */
#define VODA_ST_LESS_THAN_MIN   2  /* "Invalid GPU allocations: less than job min gpu"  utils.go:32-34 */
#define VODA_ST_EXCEEDED_MAX    3  /* "Invalid GPU allocations: exceeded job max gpu"   utils.go:35-37 */
#define VODA_ST_EXCEEDED_TOTAL  4  /* "Invalid GPU allocations: exceeded total GPUs"    utils.go:40-42 */
#define VODA_ST_INFEASIBLE      5  /* "infeasible"                              ffdl_optimizer.go:109-114 */
#define VODA_ST_NOT_FOUND       6  /* "Not found"                  types.go:45, elastic_tiresias.go:188-189 */
#define VODA_ST_NONTERMINATING  7  /* guard: the reference loop would never end (bestFit, placement_manager.go:437-485;
                                      ElasticTiresias greedy loop with MinNumProc == 0, elastic_tiresias.go:136-143:
                                      the pick changes neither the job's count nor the free GPUs nor its gain) */
#define VODA_ST_BOUND           8  /* round exceeds the max_jobs / max_total_gpu / max_max_proc the caller declared */
#define VODA_ST_RUNTIME_PANIC   9  /* the Go runtime itself would panic (makeslice with negative total) */

/* ---- policies: the eight names of NewAlgorithmFactory, pkg/algorithm/types.go:26-47 - */
#define VODA_ALGO_FIFO             0
#define VODA_ALGO_ELASTIC_FIFO     1
#define VODA_ALGO_SRJF             2
#define VODA_ALGO_ELASTIC_SRJF     3
#define VODA_ALGO_TIRESIAS         4
#define VODA_ALGO_ELASTIC_TIRESIAS 5
#define VODA_ALGO_FFDL_OPTIMIZER   6
#define VODA_ALGO_AFSL             7
#define VODA_ALGO_COUNT            8

typedef struct voda_ctx voda_ctx;

/* Create / destroy a context bound to CUDA device `device`.  Fails (VODA_E_CUDA) when no
 * device is present — there is no CPU path behind this library. */
int  voda_ctx_create(int device, voda_ctx** out);
void voda_ctx_destroy(voda_ctx* ctx);
/* The cudaStream_t (as void*) that *_dev entry points launch on; time with events on it. */
void* voda_ctx_stream(voda_ctx* ctx);
/* Block until everything enqueued on the context has finished. */
int  voda_ctx_sync(voda_ctx* ctx);
/* Number of kernels of this library launched through the context since it was created. */
int64_t voda_ctx_launch_count(voda_ctx* ctx);
/* Bytes the LAST host-pointer call on this context (voda_allocate, voda_munkres_max) copied host->device and
 * device->host — what actually crossed PCIe, e.g. only the speedup columns the policy can index. */
void voda_ctx_copied_bytes(voda_ctx* ctx, int64_t* h2d_bytes, int64_t* d2h_bytes);
const char* voda_last_error(voda_ctx* ctx);
const char* voda_status_str(int32_t status);
const char* voda_version(void);
/* sizeof() of a public struct as this library was compiled, for bindings that mirror the layouts by hand
 * (cgo does not need it; the ctypes mirror checks itself against it): which = 0 voda_alloc_batch,
 * 1 voda_place_batch, 2 voda_place_out, 3 voda_curve_batch, 4 voda_tick_batch, 5 voda_replay_batch,
 * 6 voda_replay_out; anything else -> 0. */
size_t voda_struct_size(int32_t which);

/* Name <-> id of a policy.  Replaces the string switch of NewAlgorithmFactory
 * (pkg/algorithm/types.go:26-47): unknown name -> -1 (the reference returns
 * errors.New("Not found")).  voda_algo_need_job_info mirrors NeedJobInfo()
 * (srjf.go:58-60, elastic_srjf.go:78-80, elastic_tiresias.go:165-167,
 * ffdl_optimizer.go:136-138, afsl.go:68-70; false for the others). */
int         voda_algo_from_name(const char* name);
const char* voda_algo_name(int32_t algo);
int         voda_algo_need_job_info(int32_t algo);

/* Page-locked host memory for the host-pointer entry points (optional but recommended). */
void* voda_host_alloc(size_t bytes);
void  voda_host_free(void* p);

/*
 * A batch of R independent scheduling rounds, struct-of-arrays.  Round r owns jobs
 * [job_off[r], job_off[r+1]) in the INPUT order of the slice handed to Schedule() — input
 * order is part of the input because every sort in the reference is stable.
 * Field sources: pkg/common/trainingjob/trainingjob.go:17-66.
 *   submit_*     TrainingJob.SubmitTime            (Time.Before == lexicographic (sec, nsec))
 *   first_*      TrainingJob.Metrics.FirstStartTime
 *   priority     TrainingJob.Priority
 *   num/min/max  TrainingJob.Config.{NumProc,MinNumProc,MaxNumProc}
 *   remaining    TrainingJob.Info.EstimatedRemainningTimeSec (float32)
 *   speedup      TrainingJob.Info.Speedup as a dense table: speedup[job*stride + g] is the
 *                value under key fmt.Sprint(g); a key the map lacks is 0.0f, and so is every
 *                g outside [0, stride).
 *   speedup_off  OPTIONAL: the table as a POOL of curves that jobs point into — the curve of job j is
 *                the stride entries starting at speedup[speedup_off[j]] (an offset in floats), and many
 *                jobs may share one curve.  What-if rounds are mostly the same jobs under other totals,
 *                subsets or remaining times (the map TrainingJob.Info.Speedup belongs to the job, not to
 *                the round): their curves then cross PCIe once per call instead of once per round.
 *                speedup_len = floats in speedup[].  NULL: the dense layout above.
 * Arrays a policy never reads may be NULL ("inputs per policy").
 *   max_max_proc OPTIONAL, read by the FfDL Optimizer only.  Its dynamic program keeps rows of the
 *                table P (ffdl_optimizer.go:54-60) in device scratch for the backtrack; when no job
 *                can take more than 32 GPUs it keeps every fourth row and recomputes the rest.  A
 *                caller of the _dev / _static entry points who declares the bound (<= 32) lets
 *                the scratch be SIZED for that: a quarter of the memory per round.  A round
 *                that breaks a declared bound reports VODA_ST_BOUND.  voda_allocate works the bound
 *                out from the host arrays itself.
 */
typedef struct {
  int32_t        algo;                 /* VODA_ALGO_*                                        */
  int32_t        rounds;               /* R                                                  */
  const int32_t* job_off;              /* [R+1] ascending, job_off[0] == 0                   */
  const int32_t* total_gpu;            /* [R]  second argument of Schedule()                 */
  const int64_t* submit_sec;           /* [jobs] */
  const int32_t* submit_nsec;          /* [jobs] */
  const int64_t* first_sec;            /* [jobs] */
  const int32_t* first_nsec;           /* [jobs] */
  const int32_t* priority;             /* [jobs] */
  const int32_t* num_proc;             /* [jobs] */
  const int32_t* min_proc;             /* [jobs] */
  const int32_t* max_proc;             /* [jobs] */
  const float*   remaining_sec;        /* [jobs] */
  const float*   speedup;              /* [jobs * speedup_stride] */
  int32_t        speedup_stride;       /* W; the reference's default tables have 34 entries  */
  int32_t        queue_num;            /* TiresiasQueueNum, tiresias.go:17-24; 0 -> 2        */
  int32_t        compaction_threshold; /* ElasticTiresiasCompactionThreshold,
                                          elastic_tiresias.go:18-22; negative -> 10          */
  int32_t        max_jobs;             /* upper bound of jobs per round; 0 -> computed       */
  int32_t        max_total_gpu;        /* upper bound of total_gpu;      0 -> computed       */
  const int32_t* speedup_off;          /* [jobs] or NULL                                     */
  int64_t        speedup_len;          /* floats in speedup[] when speedup_off is given      */
  int32_t        max_max_proc;         /* upper bound of the GPUs one job can take, min(MaxNumProc,
                                          total_gpu); 0 -> not declared (see below)          */
  int32_t        reserved0;            /* 0                                                  */
} voda_alloc_batch;

/*
 * Replaces SchedulerAlgorithm.Schedule(ReadyJobs, int) types.JobScheduleResult
 * (pkg/algorithm/types.go:19-24; call site pkg/allocator/allocator/resource_allocator.go:106)
 * for a whole batch of rounds.  out_gpus[j] is the reference's result[job.Name] for the job
 * at flat index j (every input job gets an entry, 0 when unallocated); out_status[r] is
 * VODA_ST_OK or the panic the reference would raise for round r, in which case out_gpus of
 * that round holds the allocation the reference had computed when it panicked.
 */
int voda_allocate(voda_ctx* ctx, const voda_alloc_batch* host_batch,
                  int32_t* out_gpus /*[jobs]*/, int32_t* out_status /*[R]*/);
/* Same, every pointer (inside the struct too) is a device pointer; max_jobs and
 * max_total_gpu must be given.  Asynchronous on voda_ctx_stream(). */
int voda_allocate_dev(voda_ctx* ctx, const voda_alloc_batch* dev_batch,
                      int32_t* d_out_gpus, int32_t* d_out_status);

/*
 * FfDL Optimizer rounds for callers that cannot allocate device memory.  This entry point has
 * no counterpart in the reference.  It computes exactly what voda_allocate_dev computes for
 * algo == VODA_ALGO_FFDL_OPTIMIZER (every pointer, inside the struct too, is a device pointer;
 * max_jobs and max_total_gpu must be declared; `algo` is ignored), but it never calls
 * cudaMalloc / cudaFree: the kernels' scratch (stored rows of the DP table) is a fixed 32 MiB array inside the library
 * image (held by every GPU that loads the library until the process exits) instead of a
 * buffer the context grows on demand.  Use it when the GPU's memory already belongs to
 * someone else's allocator, or when the calling thread cannot afford the device-wide stall
 * of an allocation.  No voda_ctx is needed: it runs on the CURRENT device, on `stream`
 * (a cudaStream_t, NULL = default stream), asynchronously.
 *   - Shapes whose per-round scratch, about J * (ceil(max_total_gpu/8) * 32 + 4 * speedup_stride + 16)
 *     bytes with J = min(max_jobs, max_total_gpu), exceeds 32 MiB are refused with VODA_E_TOO_LARGE and a
 *     message in errbuf.  With max_max_proc declared (<= 32) the first term counts J/4 rows, not J.
 *   - All calls on a GPU share the array (it keeps the previous call's bytes; nothing needs
 *     clearing: the backtrack only uses what the same round's forward pass wrote).  Calls that overlap on
 *     the GPU race on it: serialise them.
 *   - The first call on each device looks up the array's device address; make that call
 *     before any stream capture on the device.
 * errbuf (may be NULL) receives a NUL-terminated message when the result is not VODA_OK.
 */
int voda_ffdl_allocate_static(void* stream, const voda_alloc_batch* dev_batch,
                              int32_t* d_out_gpus, int32_t* d_out_status,
                              char* errbuf, size_t errbuf_len);

/*
 * Replaces m := munkres.NewMatrix(n); m.A = A; munkres.ComputeMunkresMax(m)
 * (github.com/heyfey/munkres@d4f16a25250d, not vendored; call site
 * pkg/placement/placement_manager.go:505-512) for `batch` independent n x n matrices.
 * A is row-major int64, A[b][row*n + col].  out_col_of_row[b][row] is the column of the
 * RowCol{Row,Col} pair whose Row == row (the reference returns pairs in row order).
 * 1 <= n <= 128.  Any int64 values: the reference forms MaxInt64 - A[i][j] and subtracts row minima in
 * Go's wrapping arithmetic, which is the ordinary maximum-weight assignment when every row holds
 * scores of one sign and something else (still well defined) when a row mixes negative and
 * non-negative ones; this library computes the reference's answer in both cases.
 */
int voda_munkres_max(voda_ctx* ctx, int32_t n, int32_t batch, const int64_t* A,
                     int32_t* out_col_of_row /*[batch*n]*/, int32_t* out_status /*[batch]*/);
int voda_munkres_max_dev(voda_ctx* ctx, int32_t n, int32_t batch, const int64_t* d_A,
                         int32_t* d_out_col_of_row, int32_t* d_out_status);

/*
 * A batch of R independent placement rounds: the arithmetic core of
 * PlacementManager.Place (pkg/placement/placement_manager.go:306-332) after releaseSlots,
 * i.e. bestFit (:415-487) + scoreCandidates/score (:525-544) + ComputeMunkresMax (:505-507)
 * + the rename of logical nodes (:509-511).  Everything the reference takes from Go map
 * iteration order is an explicit array here.
 *   node_off[r]..node_off[r+1]   the round's nodes.  node_slots[i] = totalSlots of LOGICAL
 *                                node i (order of the list built at :317-320); column c of
 *                                the score matrix is PHYSICAL node c (order of :494-497).
 *   req_off / req_workers        the round's requests in jobRequests iteration order;
 *                                request q of the round is "job q".
 *   prev_off / prev_job /        current (post-release) contents of the physical nodes, CSR
 *   prev_workers                 over the flat node index: entries of node i are
 *                                [prev_off[i], prev_off[i+1]), prev_job is a round-local job
 *                                index, prev_workers its jobNumWorkers value on that node.
 *                                A job appears AT MOST ONCE in a node's list (the reference keeps
 *                                a Go map per node: repeated keys cannot exist there); the values
 *                                may be anything, negative ones included.
 *   pods_off / pods_node /       OPTIONAL (all three or none): the pod -> node table the round starts
 *   pods_slots                   from (pm.podNodeName, :571-617), run-length coded per request: the
 *                                runs of request q (flat index req_off[r] + q) are [pods_off[q],
 *                                pods_off[q+1]), run t puts the next pods_slots[t] worker pods of the job
 *                                (pod index ascending) on physical node pods_node[t] — the job's
 *                                nodeNumSlotsList as updateJobStates left it BEFORE this round's
 *                                releaseSlots.  With it the kernel also counts the pods the round moves.
 */
typedef struct {
  int32_t        rounds;
  const int32_t* node_off;      /* [R+1] */
  const int32_t* node_slots;    /* [nodes] */
  const int32_t* req_off;       /* [R+1] */
  const int32_t* req_workers;   /* [requests] */
  const int32_t* prev_off;      /* [nodes+1] */
  const int32_t* prev_job;      /* [prev entries] */
  const int32_t* prev_workers;  /* [prev entries] */
  int32_t        max_nodes;     /* upper bound of nodes per round (<= 128); 0 -> computed    */
  int32_t        max_requests;  /* upper bound of requests per round;       0 -> computed    */
  const int32_t* pods_off;      /* [requests+1] or NULL */
  const int32_t* pods_node;     /* [pod runs] */
  const int32_t* pods_slots;    /* [pod runs] */
} voda_place_batch;

/*
 * Outputs, per round r with n nodes and m requests (entry base e0 = req_off[r] + node_off[r],
 * entry capacity m + n — bestFit creates at most one entry per request plus one per node):
 *   out_phys_of_logical[node_off[r] + i]  physical column bound to logical node i (:509-511)
 *   out_free[node_off[r] + i]             freeSlots of logical node i after bestFit (can be
 *                                         negative: reference quirk, placement_manager.go:479-482)
 *   out_lay_cnt[r]                        number of (node, job, workers) entries bestFit made
 *   out_lay_node/job/workers[e0 + t]      entry t in creation order: logical node index,
 *                                         round-local job index, jobNumWorkers value
 *   out_score[r]                          total overlap score of the chosen binding (:512)
 *   out_cross[r]                          jobs split across nodes (crossNode, :431,476)
 *   out_status[r]                         VODA_ST_OK / VODA_ST_NONTERMINATING / VODA_ST_BOUND;
 *                                         the other outputs of a round are unspecified
 *                                         unless its status is VODA_ST_OK
 *   out_migrated_workers[r]               OPTIONAL, needs pods_*: worker pods whose node changes
 *   out_migrated_launchers[r]             (deletedWorkers, :592-596) and jobs ALL of whose workers
 *                                         move, so that their launcher is deleted too
 *                                         (deletedLaunchers, :606-609) — updateJobStates +
 *                                         updatePodNodeName (:548-617) on the new binding, with a
 *                                         job's nodes taken by physical index (the reference takes
 *                                         them in Go map order: any order is one it could produce)
 */
typedef struct {
  int32_t* phys_of_logical;  /* [nodes] */
  int32_t* free_slots;       /* [nodes] */
  int32_t* lay_cnt;          /* [R] */
  int32_t* lay_node;         /* [requests + nodes] */
  int32_t* lay_job;          /* [requests + nodes] */
  int32_t* lay_workers;      /* [requests + nodes] */
  int64_t* score;            /* [R] */
  int32_t* cross;            /* [R] */
  int32_t* status;           /* [R] */
  int32_t* migrated_workers;    /* [R] or NULL */
  int32_t* migrated_launchers;  /* [R] or NULL */
} voda_place_out;

int voda_place(voda_ctx* ctx, const voda_place_batch* host_batch, const voda_place_out* host_out);
int voda_place_dev(voda_ctx* ctx, const voda_place_batch* dev_batch, const voda_place_out* dev_out);

/* ---- speedup-curve builder ---------------------------------------------------------------
 * Replaces the arithmetic of python/metrics_collector/metrics_collector.py:131-167
 * (_update_time_metric, _update_speedup, _update_efficiency) for a batch of jobs: the curves
 * the allocation policies read as Info.Speedup.  Job j owns the records
 * [rec_off[j], rec_off[j+1]) of its metrics CSV, in file order: workers[i] = the `workers`
 * column, epoch_time_sec[i] = the `epoch_time_sec` column.  Outputs, float64, dense over the
 * worker count k = 0 .. width-1 (the dictionaries' keys "0" .. str(width-1)), starting from
 * CreateBaseJobInfo's linear defaults (pkg/common/mongo/mongo.go:76-100):
 *   out_epoch_time[j*width + k]   mean epoch time with k workers (pandas: pairwise sum / count),
 *                                 1.0 where the job has no record with k workers, 0 at k = 0
 *   out_speedup[j*width + k]      epoch_time[1] / epoch_time[k], 0 at k = 0
 *   out_efficiency[j*width + k]   speedup[k] / k, 0 at k = 0
 *   out_status[j]                 VODA_ST_OK, or VODA_ST_BOUND when a record's worker count is
 *                                 outside [0, width) (the outputs of that job ignore such records)
 * Bit-identical to the reference run through pandas / numpy (tests/golden/curve_kats.json).
 */
typedef struct {
  int32_t jobs;
  int32_t width;                 /* 34 = keys "0" .. "33" of CreateBaseJobInfo */
  const int32_t* rec_off;        /* [jobs + 1] */
  const int32_t* workers;        /* [records] */
  const double* epoch_time_sec;  /* [records] */
} voda_curve_batch;

int voda_build_curves(voda_ctx* ctx, const voda_curve_batch* host_batch, double* out_epoch_time,
                      double* out_speedup, double* out_efficiency, int32_t* out_status);
int voda_build_curves_dev(voda_ctx* ctx, const voda_curve_batch* dev_batch, double* d_out_epoch_time,
                          double* d_out_speedup, double* d_out_efficiency, int32_t* d_out_status);

/* ---- scheduler bookkeeping between rounds ---------------------------------------------------
 * voda_tick replaces the loop body of Scheduler.updateTimeMetrics
 * (pkg/scheduler/scheduler/scheduler.go:769-803) for `jobs` jobs at once — the jobs of many
 * what-if replicas back to back: duration accounting by job status and, when `tiresias` is set
 * (s.Algorithm is "Tiresias" or "ElasticTiresias"), the demote / promote rules
 * (pkg/algorithm/tiresias.go:27-36 thresholds and knob, :109-119 helpers).  Durations are
 * time.Duration values: int64 nanoseconds.  Arrays marked io are updated in place.
 * voda_compare_results replaces Scheduler.compareResults (scheduler.go:448-480): per job the
 * action the scheduler takes when its GPU count goes from old_gpu to new_gpu.
 */
#define VODA_JOB_SUBMITTED 0 /* types.JobSubmitted .. JobFailed, pkg/common/types/types.go:34-46 */
#define VODA_JOB_WAITING 1
#define VODA_JOB_RUNNING 2
#define VODA_JOB_COMPLETED 3
#define VODA_JOB_FAILED 4

#define VODA_ACT_NONE 0
#define VODA_ACT_HALT 1      /* halts     (n > new, new == 0, job not completed / failed) */
#define VODA_ACT_SCALE_IN 2  /* scaleIns  (n > new > 0) */
#define VODA_ACT_SCALE_OUT 3 /* scaleOuts (0 < n < new) */
#define VODA_ACT_START 4     /* starts    (n == 0 < new) */

#define VODA_TICK_MAX_QUEUES 8

typedef struct {
  int32_t jobs;
  int32_t tiresias;          /* 0 / 1 */
  int32_t queue_num;         /* TiresiasQueueNum (2); <= VODA_TICK_MAX_QUEUES */
  int32_t promote_knob;      /* TiresiasPromoteKnob (8) */
  int64_t now_ns;            /* time.Now() of this tick, unix nanoseconds */
  double thresholds_sec[VODA_TICK_MAX_QUEUES]; /* TiresiasThresholdsSec[0 .. queue_num-1] */
  const int32_t* status;     /* [jobs] VODA_JOB_* */
  const int32_t* num_gpu;    /* [jobs] s.JobNumGPU[job] */
  int32_t* priority;         /* io [jobs] */
  int64_t* last_update_ns;   /* io [jobs] Metrics.LastUpdateTime */
  int64_t* running_ns;       /* io RunningDuration */
  int64_t* gpu_ns;           /* io GpuDuration */
  int64_t* total_ns;         /* io TotalDuration */
  int64_t* waiting_ns;       /* io WaitingDuration */
  int64_t* last_running_ns;  /* io LastRunningDuration */
  int64_t* last_gpu_ns;      /* io LastGpuDuration */
  int64_t* last_waiting_ns;  /* io LastWaitingDuration */
  int32_t* changed;          /* out [jobs]: 1 where the job sets priorityChanged */
} voda_tick_batch;

/* device pointers; asynchronous on the context's stream */
int voda_tick_dev(voda_ctx* ctx, const voda_tick_batch* dev_batch);
int voda_compare_results_dev(voda_ctx* ctx, int32_t jobs, const int32_t* d_old_gpu, const int32_t* d_new_gpu,
                             const int32_t* d_status, int32_t* d_action);
/* host pointers (staged through the context; returns when the outputs are in place) */
int voda_tick(voda_ctx* ctx, const voda_tick_batch* host_batch);
int voda_compare_results(voda_ctx* ctx, int32_t jobs, const int32_t* old_gpu, const int32_t* new_gpu,
                         const int32_t* status, int32_t* action);

/* ---- trace replay: many what-if replicas of the scheduler's round loop ------------------------
 * BASELINE configs[4].  One replica = one synthetic trace driven through the reference's round:
 * resched() (pkg/scheduler/scheduler/scheduler.go:326-364) = ElasticSRJF.Schedule over the ready
 * jobs, then PlacementManager.Place of the result (releaseSlots, bestFit, bindNodes,
 * updateJobStates, updatePodNodeName: pkg/placement/placement_manager.go:306-617), triggered by job
 * arrivals and completions (scheduler.go:648,670,887,955) and rate-limited to one round per
 * rate_limit_sec (scheduler.go:212,297-311: 30 s).  A replica is sequential in simulated time, so
 * the parallelism is ACROSS replicas: at every tick the rounds of all replicas that need one go to
 * the GPU as one allocation batch and one placement batch (voda_allocate, voda_compare_results,
 * voda_place on this context); the bookkeeping between the kernels (releaseSlots, updateJobStates,
 * updatePodNodeName) runs on the host, replicas spread over host threads.
 *
 * Where the reference iterates a Go map (random order) this library uses a defined order: nodes by
 * physical index; the requests of a round = the replica's live jobs in arrival order, zero
 * allocations included (JobScheduleResult holds every ready job); a node's jobs in the order
 * bestFit created them; a job's nodeNumSlotsList by physical node index (so worker pod i of a job
 * sits on the i-th slot of that order, and a scale-down releases the highest node first, :356-359).
 *
 * Model of the cluster between rounds: a job holding w GPUs progresses at S[w] work-seconds per
 * second (float64: remaining -= dt * S[min(w, W-1)]); it completes at the first tick that finds
 * remaining <= 0; an idle cluster jumps to the next arrival.
 * A round whose bestFit would never return in the reference (VODA_ST_NONTERMINATING) keeps the
 * previous bindings (only updateJobStates runs) and is counted in placements_skipped.
 */
typedef struct {
  int32_t        replicas;
  const int32_t* job_off;        /* [replicas+1] jobs of replica k are [job_off[k], job_off[k+1]) */
  const double*  arrival_sec;    /* [jobs] ascending inside a replica */
  const double*  work_sec;       /* [jobs] serial work, seconds at one GPU */
  const int32_t* min_proc;       /* [jobs] */
  const int32_t* max_proc;       /* [jobs] */
  const float*   speedup;        /* [jobs * speedup_stride] */
  int32_t        speedup_stride;
  const int32_t* nodes;          /* [replicas] nodes of the replica's cluster (<= 128) */
  const int32_t* slots;          /* [replicas] slots per node */
  const double*  rate_limit_sec; /* [replicas] */
  int64_t        max_ticks;      /* stop after this many ticks (<= 0: run to completion) */
  int32_t        host_threads;   /* bookkeeping threads (<= 0: one per four replicas, at most 16) */
} voda_replay_batch;

/* per replica k */
typedef struct {
  int64_t* rounds;              /* allocation+placement rounds run */
  int64_t* done;                /* jobs completed */
  double*  sim_time_sec;        /* simulated clock at the end */
  double*  jct_sum_sec;         /* sum of job completion times */
  int64_t* migrated_workers;    /* worker pods whose node changed (deletedWorkers, :592-596) */
  int64_t* migrated_launchers;  /* jobs all of whose workers moved (:606-609) */
  int64_t* placements_skipped;  /* rounds with VODA_ST_NONTERMINATING placement */
  int64_t* actions;             /* [replicas*4] halts, scale-ins, scale-outs, starts (compareResults) */
  uint64_t* history;            /* FNV-1a over every round's (tick, allocations, binding, score, cross,
                                   migrated workers, migrated launchers) */
  int64_t* ticks;               /* [1] ticks of the whole run */
} voda_replay_out;

int voda_replay(voda_ctx* ctx, const voda_replay_batch* batch, const voda_replay_out* out);

#ifdef __cplusplus
}
#endif
#endif /* VODA_ACCEL_H */
