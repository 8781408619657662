"""ctypes mirror of include/voda_accel.h (struct layouts, constants, prototypes).

Pure declarations: nothing here computes.  Shared by the product binding (_lib.py) and by the
test-only checker binding kept outside this package, so both see the same struct layout.
"""
import ctypes as C

OK, E_INVALID, E_CUDA, E_NOMEM, E_UNSUPPORTED, E_TOO_LARGE = 0, -1, -2, -3, -4, -5

ST_OK, ST_NEGATIVE, ST_LESS_THAN_MIN, ST_EXCEEDED_MAX, ST_EXCEEDED_TOTAL = 0, 1, 2, 3, 4
ST_INFEASIBLE, ST_NOT_FOUND, ST_NONTERMINATING, ST_BOUND, ST_RUNTIME_PANIC = 5, 6, 7, 8, 9

# pkg/algorithm/types.go:26-47 — the eight factory names, in ABI id order
ALGO_NAMES = ("FIFO", "ElasticFIFO", "SRJF", "ElasticSRJF", "Tiresias", "ElasticTiresias",
              "FfDLOptimizer", "AFS-L")
(ALGO_FIFO, ALGO_ELASTIC_FIFO, ALGO_SRJF, ALGO_ELASTIC_SRJF, ALGO_TIRESIAS,
 ALGO_ELASTIC_TIRESIAS, ALGO_FFDL_OPTIMIZER, ALGO_AFSL) = range(8)

i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
f32p = C.POINTER(C.c_float)


class AllocBatchStruct(C.Structure):
    _fields_ = [
        ("algo", C.c_int32), ("rounds", C.c_int32),
        ("job_off", C.c_void_p), ("total_gpu", C.c_void_p),
        ("submit_sec", C.c_void_p), ("submit_nsec", C.c_void_p),
        ("first_sec", C.c_void_p), ("first_nsec", C.c_void_p),
        ("priority", C.c_void_p), ("num_proc", C.c_void_p),
        ("min_proc", C.c_void_p), ("max_proc", C.c_void_p),
        ("remaining_sec", C.c_void_p), ("speedup", C.c_void_p),
        ("speedup_stride", C.c_int32), ("queue_num", C.c_int32),
        ("compaction_threshold", C.c_int32), ("max_jobs", C.c_int32),
        ("max_total_gpu", C.c_int32),
        ("speedup_off", C.c_void_p), ("speedup_len", C.c_int64),
        ("max_max_proc", C.c_int32), ("reserved0", C.c_int32),
    ]


class PlaceBatchStruct(C.Structure):
    _fields_ = [
        ("rounds", C.c_int32),
        ("node_off", C.c_void_p), ("node_slots", C.c_void_p),
        ("req_off", C.c_void_p), ("req_workers", C.c_void_p),
        ("prev_off", C.c_void_p), ("prev_job", C.c_void_p), ("prev_workers", C.c_void_p),
        ("max_nodes", C.c_int32), ("max_requests", C.c_int32),
        ("pods_off", C.c_void_p), ("pods_node", C.c_void_p), ("pods_slots", C.c_void_p),
    ]


class PlaceOutStruct(C.Structure):
    _fields_ = [
        ("phys_of_logical", C.c_void_p), ("free_slots", C.c_void_p), ("lay_cnt", C.c_void_p),
        ("lay_node", C.c_void_p), ("lay_job", C.c_void_p), ("lay_workers", C.c_void_p),
        ("score", C.c_void_p), ("cross", C.c_void_p), ("status", C.c_void_p),
        ("migrated_workers", C.c_void_p), ("migrated_launchers", C.c_void_p),
    ]


# every symbol include/voda_accel.h declares: name -> (restype, argtypes)
class CurveBatchStruct(C.Structure):  # voda_curve_batch
    _fields_ = [("jobs", C.c_int32), ("width", C.c_int32), ("rec_off", C.c_void_p), ("workers", C.c_void_p),
                ("epoch_time_sec", C.c_void_p)]


class ReplayBatchStruct(C.Structure):  # voda_replay_batch
    _fields_ = [("replicas", C.c_int32), ("job_off", C.c_void_p), ("arrival_sec", C.c_void_p), ("work_sec", C.c_void_p),
                ("min_proc", C.c_void_p), ("max_proc", C.c_void_p), ("speedup", C.c_void_p),
                ("speedup_stride", C.c_int32), ("nodes", C.c_void_p), ("slots", C.c_void_p),
                ("rate_limit_sec", C.c_void_p), ("max_ticks", C.c_int64), ("host_threads", C.c_int32)]


class ReplayOutStruct(C.Structure):  # voda_replay_out
    _fields_ = [("rounds", C.c_void_p), ("done", C.c_void_p), ("sim_time_sec", C.c_void_p), ("jct_sum_sec", C.c_void_p),
                ("migrated_workers", C.c_void_p), ("migrated_launchers", C.c_void_p), ("placements_skipped", C.c_void_p),
                ("actions", C.c_void_p), ("history", C.c_void_p), ("ticks", C.c_void_p)]


TICK_MAX_QUEUES = 8
JOB_SUBMITTED, JOB_WAITING, JOB_RUNNING, JOB_COMPLETED, JOB_FAILED = range(5)
ACT_NONE, ACT_HALT, ACT_SCALE_IN, ACT_SCALE_OUT, ACT_START = range(5)


class TickBatchStruct(C.Structure):  # voda_tick_batch
    _fields_ = [("jobs", C.c_int32), ("tiresias", C.c_int32), ("queue_num", C.c_int32), ("promote_knob", C.c_int32),
                ("now_ns", C.c_int64), ("thresholds_sec", C.c_double * TICK_MAX_QUEUES),
                ("status", C.c_void_p), ("num_gpu", C.c_void_p), ("priority", C.c_void_p),
                ("last_update_ns", C.c_void_p), ("running_ns", C.c_void_p), ("gpu_ns", C.c_void_p),
                ("total_ns", C.c_void_p), ("waiting_ns", C.c_void_p), ("last_running_ns", C.c_void_p),
                ("last_gpu_ns", C.c_void_p), ("last_waiting_ns", C.c_void_p), ("changed", C.c_void_p)]


PROTOTYPES = {
    "voda_ctx_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "voda_ctx_destroy": (None, [C.c_void_p]),
    "voda_ctx_stream": (C.c_void_p, [C.c_void_p]),
    "voda_ctx_sync": (C.c_int, [C.c_void_p]),
    "voda_ctx_launch_count": (C.c_int64, [C.c_void_p]),
    "voda_ctx_copied_bytes": (None, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "voda_last_error": (C.c_char_p, [C.c_void_p]),
    "voda_status_str": (C.c_char_p, [C.c_int32]),
    "voda_version": (C.c_char_p, []),
    "voda_struct_size": (C.c_size_t, [C.c_int32]),
    "voda_algo_from_name": (C.c_int, [C.c_char_p]),
    "voda_algo_name": (C.c_char_p, [C.c_int32]),
    "voda_algo_need_job_info": (C.c_int, [C.c_int32]),
    "voda_host_alloc": (C.c_void_p, [C.c_size_t]),
    "voda_host_free": (None, [C.c_void_p]),
    "voda_allocate": (C.c_int, [C.c_void_p, C.POINTER(AllocBatchStruct), C.c_void_p, C.c_void_p]),
    "voda_allocate_dev": (C.c_int, [C.c_void_p, C.POINTER(AllocBatchStruct), C.c_void_p, C.c_void_p]),
    "voda_ffdl_allocate_static": (C.c_int, [C.c_void_p, C.POINTER(AllocBatchStruct), C.c_void_p, C.c_void_p,
                                            C.c_char_p, C.c_size_t]),
    "voda_munkres_max": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "voda_munkres_max_dev": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "voda_place": (C.c_int, [C.c_void_p, C.POINTER(PlaceBatchStruct), C.POINTER(PlaceOutStruct)]),
    "voda_place_dev": (C.c_int, [C.c_void_p, C.POINTER(PlaceBatchStruct), C.POINTER(PlaceOutStruct)]),
    "voda_tick": (C.c_int, [C.c_void_p, C.POINTER(TickBatchStruct)]),
    "voda_tick_dev": (C.c_int, [C.c_void_p, C.POINTER(TickBatchStruct)]),
    "voda_compare_results": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "voda_compare_results_dev": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "voda_replay": (C.c_int, [C.c_void_p, C.POINTER(ReplayBatchStruct), C.POINTER(ReplayOutStruct)]),
    "voda_build_curves": (C.c_int, [C.c_void_p, C.POINTER(CurveBatchStruct), C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "voda_build_curves_dev": (C.c_int, [C.c_void_p, C.POINTER(CurveBatchStruct), C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p]),
}


def bind(lib, prototypes=PROTOTYPES):
    """Attach restype/argtypes; raises AttributeError when a declared symbol is missing."""
    for name, (res, args) in prototypes.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib
