"""Host-side mirror of the reference's allocation interface (pkg/algorithm/types.go:16-47).

Same names and meaning as the Go code so callers (and tests) read like the reference:

    algo = NewAlgorithmFactory("ElasticTiresias", "scheduler-0")   # types.go:26-47
    algo.GetName(); algo.NeedJobInfo()
    result = algo.Schedule(ready_jobs, total_gpu)                   # {job name: GPUs}

`Schedule` flattens the jobs BY INDEX into the struct-of-arrays batch of
include/voda_accel.h, runs the round on the GPU through the C ABI and rebuilds the
`JobScheduleResult` map (pkg/common/types/types.go:61).  Where the reference panics
(pkg/algorithm/utils.go:29-40, ffdl_optimizer.go:109-114) this raises SchedulePanic carrying the
same message.  Nothing here computes an allocation on the CPU.

Differences, on purpose: the Go code sorts the caller's slice in place (elastic_fifo.go:31); the
caller never reads it afterwards (resource_allocator.go:106-110), so this mirror leaves the list
alone.  `ScheduleBatch` is the batched form the GPU is built for.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, NamedTuple, Optional, Sequence, Tuple

import numpy as np

from . import _abi
from .batch import AllocBatch

MAX_NUM_GPU = 32  # trainingjob.go:13


class Time(NamedTuple):
    """time.Time reduced to what Time.Before compares: (seconds, nanoseconds)."""
    sec: int
    nsec: int = 0


# types.MaxTime, pkg/common/types/types.go:65 — FirstStartTime of a job that never started
MaxTime = Time((1 << 63) - 62135596801, 999999999)


@dataclass
class JobConfig:  # trainingjob.go:33-38
    NumProc: int = 1
    MinNumProc: int = 1
    MaxNumProc: int = 1
    Epochs: int = 0


@dataclass
class JobMetrics:  # trainingjob.go:41-55 (only the field the policies read)
    FirstStartTime: Time = MaxTime


@dataclass
class JobInfo:  # trainingjob.go:58-66
    Name: str = ""
    Category: str = ""
    GpuType: str = ""
    EstimatedRemainningTimeSec: float = 0.0
    Speedup: Dict[str, float] = field(default_factory=dict)
    Efficiency: Dict[str, float] = field(default_factory=dict)


def NewBaseJobInfo(name: str, category: str = "", gpuType: str = "") -> JobInfo:
    """Linear-speedup default, trainingjob.go:168-186: keys "0".."33"."""
    speedup = {"0": 0.0}
    efficiency = {"0": 0.0}
    for i in range(1, MAX_NUM_GPU + 2):
        speedup[str(i)] = float(i)
        efficiency[str(i)] = 1.0
    return JobInfo(name, category, gpuType, 0.0, speedup, efficiency)


@dataclass
class TrainingJob:  # trainingjob.go:17-31
    Name: str
    Priority: int = 0
    SubmitTime: Time = Time(0, 0)
    Config: JobConfig = field(default_factory=JobConfig)
    Metrics: JobMetrics = field(default_factory=JobMetrics)
    Info: Optional[JobInfo] = None


ReadyJobs = List[TrainingJob]
JobScheduleResult = Dict[str, int]


class SchedulePanic(RuntimeError):
    """The reference would panic with this message (status code in .status)."""

    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status = status


def _table_width(jobs: Sequence[TrainingJob]) -> int:
    width = MAX_NUM_GPU + 2
    for j in jobs:
        if j.Info is not None:
            for k in j.Info.Speedup:
                try:
                    width = max(width, int(k) + 1)
                except ValueError:
                    pass  # a key fmt.Sprint(int) can never produce is never looked up
    return width


def pack_rounds(algo: int, rounds: Sequence[Tuple[Sequence[TrainingJob], int]], queue_num=0,
                compaction_threshold=-1) -> AllocBatch:
    """ReadyJobs of several rounds -> one AllocBatch (input order preserved)."""
    width = max([_table_width(jobs) for jobs, _ in rounds] or [MAX_NUM_GPU + 2])
    n = sum(len(jobs) for jobs, _ in rounds)
    off = np.zeros(len(rounds) + 1, np.int32)
    tot = np.zeros(len(rounds), np.int32)
    cols = {k: np.zeros(n, dt) for k, dt in (
        ("submit_sec", np.int64), ("submit_nsec", np.int32), ("first_sec", np.int64), ("first_nsec", np.int32),
        ("priority", np.int32), ("num_proc", np.int32), ("min_proc", np.int32), ("max_proc", np.int32),
        ("remaining_sec", np.float32))}
    S = np.zeros((n, width), np.float32)
    i = 0
    for r, (jobs, total) in enumerate(rounds):
        tot[r] = total
        for j in jobs:
            cols["submit_sec"][i], cols["submit_nsec"][i] = j.SubmitTime
            cols["first_sec"][i], cols["first_nsec"][i] = j.Metrics.FirstStartTime
            cols["priority"][i] = j.Priority
            cols["num_proc"][i] = j.Config.NumProc
            cols["min_proc"][i] = j.Config.MinNumProc
            cols["max_proc"][i] = j.Config.MaxNumProc
            if j.Info is not None:
                cols["remaining_sec"][i] = j.Info.EstimatedRemainningTimeSec
                for k, v in j.Info.Speedup.items():
                    try:
                        g = int(k)
                    except ValueError:
                        continue
                    if 0 <= g < width and str(g) == k:
                        S[i, g] = v
            i += 1
        off[r + 1] = i
    return AllocBatch(algo, off, tot, speedup=S.reshape(-1), speedup_stride=width, queue_num=queue_num,
                      compaction_threshold=compaction_threshold, **cols)


class SchedulerAlgorithm:
    """type SchedulerAlgorithm interface, types.go:19-24."""

    _ctx = None  # one engine context per process for the single-round convenience path

    def __init__(self, algorithm: str, schedulerID: str, ctx=None):
        self.algorithm = algorithm
        self.schedulerID = schedulerID
        self._id = _abi.ALGO_NAMES.index(algorithm)
        self._own_ctx = ctx

    def GetName(self) -> str:
        return self.algorithm

    def NeedJobInfo(self) -> bool:
        from ._lib import load_library
        return bool(load_library().voda_algo_need_job_info(self._id))

    def _context(self):
        if self._own_ctx is not None:
            return self._own_ctx
        if SchedulerAlgorithm._ctx is None:
            from ._lib import Context
            SchedulerAlgorithm._ctx = Context(0)
        return SchedulerAlgorithm._ctx

    def ScheduleBatch(self, rounds: Sequence[Tuple[Sequence[TrainingJob], int]]):
        """Many independent rounds at once -> list of (result map | SchedulePanic)."""
        from ._lib import status_str
        from .algorithm_defaults import TiresiasQueueNum, ElasticTiresiasCompactionThreshold
        if self.NeedJobInfo():
            for jobs, _ in rounds:
                for j in jobs:
                    if j.Info is None:  # the Go code dereferences a nil *JobInfo here
                        raise SchedulePanic(_abi.ST_RUNTIME_PANIC, "runtime error: invalid memory address or nil pointer dereference")
        b = pack_rounds(self._id, rounds, TiresiasQueueNum, ElasticTiresiasCompactionThreshold)
        gpus, status = self._context().allocate(b)
        out = []
        for r, (jobs, _) in enumerate(rounds):
            if status[r] != 0:
                out.append(SchedulePanic(int(status[r]), status_str(status[r])))
                continue
            o = int(b.job_off[r])
            res = {j.Name: int(gpus[o + i]) for i, j in enumerate(jobs)}
            if self.algorithm == "Tiresias":
                # tiresias.go:81-90 only writes jobs whose Priority names an existing queue
                res = {j.Name: res[j.Name] for j in jobs if 0 <= j.Priority < TiresiasQueueNum}
            out.append(res)
        return out

    def Schedule(self, jobs: ReadyJobs, totalGPU: int) -> JobScheduleResult:
        res = self.ScheduleBatch([(jobs, totalGPU)])[0]
        if isinstance(res, SchedulePanic):
            raise res
        return res


def NewAlgorithmFactory(algorithm: str, schedulerID: str, ctx=None) -> SchedulerAlgorithm:
    """types.go:26-47: one of the eight names, anything else is errors.New("Not found")."""
    if algorithm not in _abi.ALGO_NAMES:
        raise KeyError("Not found")
    return SchedulerAlgorithm(algorithm, schedulerID, ctx)
