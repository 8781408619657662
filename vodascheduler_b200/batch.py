"""Struct-of-arrays batches handed to the C ABI (include/voda_accel.h).

A batch is R independent rounds flattened job-major; the arrays are exactly the fields the
reference policies read from TrainingJob (pkg/common/trainingjob/trainingjob.go:17-66).  This
module only lays data out; it does no scheduling arithmetic.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi

_ALLOC_ARRAYS = (
    ("job_off", np.int32), ("total_gpu", np.int32),
    ("submit_sec", np.int64), ("submit_nsec", np.int32),
    ("first_sec", np.int64), ("first_nsec", np.int32),
    ("priority", np.int32), ("num_proc", np.int32),
    ("min_proc", np.int32), ("max_proc", np.int32),
    ("remaining_sec", np.float32), ("speedup", np.float32),
)

_PLACE_OPTIONAL = (("pods_off", np.int32), ("pods_node", np.int32), ("pods_slots", np.int32))

_PLACE_ARRAYS = (
    ("node_off", np.int32), ("node_slots", np.int32), ("req_off", np.int32),
    ("req_workers", np.int32), ("prev_off", np.int32), ("prev_job", np.int32),
    ("prev_workers", np.int32),
)


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor (device or pinned host)


class AllocBatch:
    """R scheduling rounds for one policy.  Arrays are numpy (host) or torch (device)."""

    def __init__(self, algo, job_off, total_gpu, *, submit_sec=None, submit_nsec=None,
                 first_sec=None, first_nsec=None, priority=None, num_proc=None, min_proc=None,
                 max_proc=None, remaining_sec=None, speedup=None, speedup_stride=0,
                 queue_num=0, compaction_threshold=-1, max_jobs=0, max_total_gpu=0, speedup_off=None,
                 max_max_proc=0):
        if isinstance(algo, str):
            algo = _abi.ALGO_NAMES.index(algo)
        self.algo = int(algo)
        loc = locals()
        for name, dt in _ALLOC_ARRAYS:
            v = loc[name]
            if v is not None and isinstance(v, (np.ndarray, list, tuple)):
                v = np.ascontiguousarray(np.asarray(v, dtype=dt).reshape(-1))
            setattr(self, name, v)
        # optional: the speedup table as a pool of curves, speedup_off[j] = first float of job j's curve
        self.speedup_off = None if speedup_off is None else (
            np.ascontiguousarray(np.asarray(speedup_off, dtype=np.int32).reshape(-1))
            if isinstance(speedup_off, (np.ndarray, list, tuple)) else speedup_off)
        self.speedup_stride = int(speedup_stride)
        self.queue_num = int(queue_num)
        self.compaction_threshold = int(compaction_threshold)
        self.max_jobs = int(max_jobs)
        self.max_total_gpu = int(max_total_gpu)
        self.max_max_proc = int(max_max_proc)  # declared bound of the GPUs one job can take; 0: not declared

    @property
    def rounds(self):
        return int(self.job_off.shape[0]) - 1

    @property
    def jobs(self):
        return int(self.job_off[-1]) if isinstance(self.job_off, np.ndarray) else int(self._jobs)

    def struct(self):
        s = _abi.AllocBatchStruct()
        s.algo, s.rounds = self.algo, self.rounds
        for name, _ in _ALLOC_ARRAYS:
            setattr(s, name, _ptr(getattr(self, name)))
        s.speedup_off = _ptr(getattr(self, "speedup_off", None))
        sp = getattr(self, "speedup", None)
        s.speedup_len = 0 if sp is None else int(sp.shape[0] if isinstance(sp, np.ndarray) else sp.numel())
        s.speedup_stride = self.speedup_stride
        s.queue_num = self.queue_num
        s.compaction_threshold = self.compaction_threshold
        s.max_jobs = self.max_jobs
        s.max_total_gpu = self.max_total_gpu
        s.max_max_proc = getattr(self, "max_max_proc", 0)
        return s

    def with_algo(self, algo):
        b = AllocBatch.__new__(AllocBatch)
        b.__dict__.update(self.__dict__)
        b.algo = _abi.ALGO_NAMES.index(algo) if isinstance(algo, str) else int(algo)
        return b

    def slice_rounds(self, r0, r1):
        """Rounds [r0, r1) as a new host batch (used to shard rounds across ranks)."""
        j0, j1 = int(self.job_off[r0]), int(self.job_off[r1])
        kw = {}
        for name, _ in _ALLOC_ARRAYS[2:]:
            v = getattr(self, name)
            if v is None:
                kw[name] = None
            elif name == "speedup":
                # a pool of curves stays whole (the jobs' offsets are sliced); a dense table is cut
                kw[name] = v if getattr(self, "speedup_off", None) is not None else \
                    v[j0 * self.speedup_stride:j1 * self.speedup_stride]
            else:
                kw[name] = v[j0:j1]
        if getattr(self, "speedup_off", None) is not None:
            kw["speedup_off"] = self.speedup_off[j0:j1]
        return AllocBatch(self.algo, self.job_off[r0:r1 + 1] - j0, self.total_gpu[r0:r1],
                          speedup_stride=self.speedup_stride, queue_num=self.queue_num,
                          compaction_threshold=self.compaction_threshold,
                          max_jobs=self.max_jobs, max_total_gpu=self.max_total_gpu,
                          max_max_proc=getattr(self, "max_max_proc", 0), **kw)

    def bounds(self):
        n = np.diff(self.job_off)
        return (int(n.max()) if n.size else 0,
                int(max(0, self.total_gpu.max())) if self.total_gpu.size else 0)

    def to_device(self, device):
        """Copy every array to `device` as torch tensors (plumbing only: device memory)."""
        import torch
        b = AllocBatch.__new__(AllocBatch)
        b.__dict__.update(self.__dict__)
        b._jobs = self.jobs
        mj, mt = self.bounds()
        b.max_jobs = self.max_jobs or mj
        b.max_total_gpu = self.max_total_gpu or mt
        if not getattr(self, "max_max_proc", 0) and self.max_proc is not None and self.max_proc.size:
            b.max_max_proc = max(1, int(self.max_proc.max()))  # declared like the two bounds above
        for name, _ in _ALLOC_ARRAYS:
            v = getattr(self, name)
            setattr(b, name, None if v is None else torch.from_numpy(v).to(device))
        so = getattr(self, "speedup_off", None)
        b.speedup_off = None if so is None else torch.from_numpy(so).to(device)
        return b

    def nbytes(self):
        return sum(getattr(self, n).nbytes for n, _ in _ALLOC_ARRAYS
                   if isinstance(getattr(self, n), np.ndarray))


class PlaceBatch:
    """R placement rounds (see voda_place_batch in include/voda_accel.h)."""

    def __init__(self, node_off, node_slots, req_off, req_workers, prev_off, prev_job,
                 prev_workers, max_nodes=0, max_requests=0, pods_off=None, pods_node=None, pods_slots=None):
        loc = locals()
        for name, dt in _PLACE_ARRAYS + _PLACE_OPTIONAL:
            v = loc[name]
            if isinstance(v, (np.ndarray, list, tuple)):
                v = np.ascontiguousarray(np.asarray(v, dtype=dt).reshape(-1))
            setattr(self, name, v)
        self.max_nodes = int(max_nodes)
        self.max_requests = int(max_requests)

    @property
    def rounds(self):
        return int(self.node_off.shape[0]) - 1

    @property
    def nodes(self):
        return int(self.node_off[-1])

    @property
    def requests(self):
        return int(self.req_off[-1])

    def struct(self):
        s = _abi.PlaceBatchStruct()
        s.rounds = self.rounds
        for name, _ in _PLACE_ARRAYS + _PLACE_OPTIONAL:
            setattr(s, name, _ptr(getattr(self, name, None)))
        s.max_nodes, s.max_requests = self.max_nodes, self.max_requests
        return s

    def to_device(self, device):
        import torch
        b = PlaceBatch.__new__(PlaceBatch)
        b.__dict__.update(self.__dict__)
        b._nodes, b._requests = self.nodes, self.requests
        b.max_nodes = self.max_nodes or int(np.diff(self.node_off).max(initial=0))
        b.max_requests = self.max_requests or int(np.diff(self.req_off).max(initial=0))
        for name, _ in _PLACE_ARRAYS:
            setattr(b, name, torch.from_numpy(getattr(self, name)).to(device))
        for name, _ in _PLACE_OPTIONAL:
            v = getattr(self, name, None)
            setattr(b, name, None if v is None else torch.from_numpy(v).to(device))
        return b


class PlaceOut:
    """Caller-owned output arrays of voda_place (numpy)."""

    def __init__(self, rounds, nodes, requests):
        cap = max(1, nodes + requests)
        self.phys_of_logical = np.full(max(1, nodes), -1, np.int32)
        self.free_slots = np.zeros(max(1, nodes), np.int32)
        self.lay_cnt = np.zeros(max(1, rounds), np.int32)
        self.lay_node = np.zeros(cap, np.int32)
        self.lay_job = np.zeros(cap, np.int32)
        self.lay_workers = np.zeros(cap, np.int32)
        self.score = np.zeros(max(1, rounds), np.int64)
        self.cross = np.zeros(max(1, rounds), np.int32)
        self.status = np.full(max(1, rounds), -1, np.int32)
        self.migrated_workers = np.zeros(max(1, rounds), np.int32)
        self.migrated_launchers = np.zeros(max(1, rounds), np.int32)

    def struct(self):
        s = _abi.PlaceOutStruct()
        for name, _ in _abi.PlaceOutStruct._fields_:
            setattr(s, name, _ptr(getattr(self, name)))
        return s
