"""The allocator's wire format: the JSON the scheduler POSTs to
`/allocation` and the JSON map it gets back, plus the batch form the GPU is built for.

Mirrors pkg/allocator/allocator:
  AllocationRequest {SchedulerID, NumGpu, AlgorithmName, ReadyJobs}      types.go:5-10 (no json
      tags, so Go's default field names are the keys)
  ReadyJobs = []TrainingJob with the json tags of pkg/common/trainingjob/trainingjob.go:17-66
      (job_name, priority, submit_timestamp, config{num_proc,min_num_proc,max_num_proc},
       time_metrics{first_start_timestamp}, info{estimate_remainning_time_seconds, speedup})
  allocateResource(req) -> types.JobScheduleResult                        resource_allocator.go:76-111
Times are RFC 3339 strings with up to nanosecond fractions (Go's time.Time JSON form).

`allocate_resource` is the one-request path; `allocate_resource_batch` decodes many requests,
groups them by policy and sends each group to the GPU as one batch (a `POST /allocation/batch`
body is simply a JSON array of AllocationRequest objects).  Decoding is host work; every
allocation is computed by the CUDA engine.
"""
from __future__ import annotations

import datetime as _dt
import json

import numpy as np
import re
from typing import Dict, List, Sequence, Tuple, Union

from . import _abi
from .algorithm import (JobConfig, JobInfo, JobMetrics, NewAlgorithmFactory, NewBaseJobInfo, SchedulePanic,
                        Time, TrainingJob)

_RFC3339 = re.compile(
    r"^(\d{4})-(\d{2})-(\d{2})[Tt](\d{2}):(\d{2}):(\d{2})(?:\.(\d{1,9}))?(Z|z|[+-]\d{2}:\d{2})$")
_EPOCH = _dt.datetime(1970, 1, 1)
# Go's zero time.Time, what an absent timestamp decodes to: January 1, year 1, 00:00 UTC
ZERO_TIME = Time(-62135596800, 0)


def parse_time(s: str) -> Time:
    """RFC 3339 -> (unix seconds, nanoseconds), the two numbers Time.Before compares."""
    m = _RFC3339.match(s)
    if not m:
        raise ValueError(f"parsing time {s!r}: not RFC 3339")
    y, mo, d, hh, mm, ss, frac, zone = m.groups()
    t = _dt.datetime(int(y), int(mo), int(d), int(hh), int(mm), int(ss)) - _EPOCH
    sec = t.days * 86400 + t.seconds
    if zone not in ("Z", "z"):
        sign = 1 if zone[0] == "+" else -1
        sec -= sign * (int(zone[1:3]) * 3600 + int(zone[4:6]) * 60)
    nsec = int((frac or "0").ljust(9, "0"))
    return Time(sec, nsec)


def format_time(t: Time) -> str:
    base = _EPOCH + _dt.timedelta(seconds=t.sec)
    frac = f".{t.nsec:09d}".rstrip("0") if t.nsec else ""
    return (f"{base.year:04d}-{base.month:02d}-{base.day:02d}T{base.hour:02d}:{base.minute:02d}:{base.second:02d}"
            + frac + "Z")


def parse_float32(text: str) -> float:
    """A JSON number as Go's encoding/json stores it into a float32 field: the decimal text rounded
    ONCE to the nearest float32 (strconv.ParseFloat(s, 32)).  Rounding to float64 first and to float32
    afterwards differs when the float64 lands exactly half-way between two float32 values while the
    decimal itself is a hair to one side; that case is resolved with exact rational arithmetic."""
    x = float(text)
    with np.errstate(over="ignore"):
        f = np.float32(x)
    if not np.isfinite(f) or float(f) == x:
        return float(f)
    lo, hi = np.float32(np.nextafter(x, -np.inf)), np.float32(np.nextafter(x, np.inf))
    if lo != hi:  # x is a float32 midpoint: the side is decided by the exact decimal
        from fractions import Fraction
        exact, mid = Fraction(text), Fraction(x)
        if exact > mid:
            return float(hi)
        if exact < mid:
            return float(lo)
    return float(f)


def job_from_json(o: dict) -> TrainingJob:
    """One element of ReadyJobs (unknown keys are ignored, like encoding/json does)."""
    cfg = o.get("config") or {}
    tm = o.get("time_metrics")
    info = o.get("info")
    ji = None
    if info is not None:  # `info` is omitempty: absent means a nil *JobInfo
        ji = JobInfo(Name=info.get("job_name", ""), Category=info.get("job_category", ""),
                     GpuType=info.get("gpu_type", ""),
                     EstimatedRemainningTimeSec=float(info.get("estimate_remainning_time_seconds", 0.0)),
                     Speedup={str(k): float(v) for k, v in (info.get("speedup") or {}).items()},
                     Efficiency={str(k): float(v) for k, v in (info.get("efficiency") or {}).items()})
    first = ZERO_TIME
    if tm is not None and tm.get("first_start_timestamp"):
        first = parse_time(tm["first_start_timestamp"])
    return TrainingJob(
        Name=o.get("job_name", ""), Priority=int(o.get("priority", 0)),
        SubmitTime=parse_time(o["submit_timestamp"]) if o.get("submit_timestamp") else ZERO_TIME,
        Config=JobConfig(int(cfg.get("num_proc", 0)), int(cfg.get("min_num_proc", 0)),
                         int(cfg.get("max_num_proc", 0)), int(cfg.get("epochs", 0))),
        Metrics=JobMetrics(first), Info=ji)


def job_to_json(j: TrainingJob) -> dict:
    o = {"job_name": j.Name, "priority": j.Priority, "submit_timestamp": format_time(j.SubmitTime),
         "config": {"num_proc": j.Config.NumProc, "min_num_proc": j.Config.MinNumProc,
                    "max_num_proc": j.Config.MaxNumProc, "epochs": j.Config.Epochs},
         "time_metrics": {"first_start_timestamp": format_time(j.Metrics.FirstStartTime)}}
    if j.Info is not None:
        o["info"] = {"job_name": j.Info.Name, "job_category": j.Info.Category, "gpu_type": j.Info.GpuType,
                     "estimate_remainning_time_seconds": j.Info.EstimatedRemainningTimeSec,
                     "speedup": dict(j.Info.Speedup), "efficiency": dict(j.Info.Efficiency)}
    return o


class AllocationRequest:
    """types.go:5-10."""

    def __init__(self, SchedulerID: str, NumGpu: int, AlgorithmName: str, ReadyJobs: List[TrainingJob]):
        self.SchedulerID, self.NumGpu, self.AlgorithmName, self.ReadyJobs = SchedulerID, NumGpu, AlgorithmName, ReadyJobs

    @staticmethod
    def from_json(body: Union[str, bytes, dict]) -> "AllocationRequest":
        # float fields of the request are float32 in Go (trainingjob.go:63-66): parse them as such
        o = json.loads(body, parse_float=parse_float32) if isinstance(body, (str, bytes, bytearray)) else body
        return AllocationRequest(o.get("SchedulerID", ""), int(o.get("NumGpu", 0)), o.get("AlgorithmName", ""),
                                 [job_from_json(j) for j in (o.get("ReadyJobs") or [])])

    def to_json(self) -> str:
        return json.dumps({"SchedulerID": self.SchedulerID, "NumGpu": self.NumGpu,
                           "AlgorithmName": self.AlgorithmName,
                           "ReadyJobs": [job_to_json(j) for j in self.ReadyJobs]})


def _fill_info(req: AllocationRequest, info_lookup) -> None:
    """The job_info store is outside the path.  In the reference getJobsInfo assigns to the range
    copy (resource_allocator.go:121,134), so only `info` that arrived in the JSON reaches Schedule;
    a caller that owns a store passes `info_lookup(job) -> JobInfo | None` and it is applied by
    index, None meaning the linear NewBaseJobInfo default (resource_allocator.go:122-133)."""
    if info_lookup is None:
        return
    for j in req.ReadyJobs:
        got = info_lookup(j)
        j.Info = got if got is not None else NewBaseJobInfo(j.Name)


def allocate_resource(body: Union[str, bytes, dict, AllocationRequest], ctx=None, info_lookup=None
                      ) -> Dict[str, int]:
    """resource_allocator.go:76-111 for one request.  Raises KeyError("Not found") for an unknown
    policy name (the handler answers 400) and SchedulePanic where Schedule would panic."""
    req = body if isinstance(body, AllocationRequest) else AllocationRequest.from_json(body)
    algo = NewAlgorithmFactory(req.AlgorithmName, req.SchedulerID, ctx)
    if algo.NeedJobInfo():
        _fill_info(req, info_lookup)
    return algo.Schedule(req.ReadyJobs, req.NumGpu)


def handle_allocation(body: Union[str, bytes], ctx=None, info_lookup=None) -> Tuple[int, str]:
    """allocateResourceHandler, resource_allocator.go:41-74: (HTTP status, JSON body).  A body that
    does not decode and an unknown policy answer 400; a panic inside Schedule is not caught there
    (net/http drops the connection), so SchedulePanic propagates to the caller here too."""
    try:
        req = AllocationRequest.from_json(body)
        res = allocate_resource(req, ctx, info_lookup)
    except (ValueError, KeyError, TypeError, AttributeError) as e:  # json / time / "Not found"
        return 400, json.dumps({"error": str(e.args[0]) if e.args else str(e)})
    return 200, json.dumps(res)


def allocate_resource_batch(bodies: Sequence[Union[str, bytes, dict, AllocationRequest]], ctx=None,
                            info_lookup=None) -> List[Union[Dict[str, int], Exception]]:
    """Many requests at once: one GPU batch per policy.  The i-th result is the allocation map,
    or the exception the single-request path would have raised for that request."""
    reqs = [b if isinstance(b, AllocationRequest) else AllocationRequest.from_json(b) for b in bodies]
    out: List[Union[Dict[str, int], Exception]] = [None] * len(reqs)  # type: ignore[list-item]
    groups: Dict[str, List[int]] = {}
    for i, r in enumerate(reqs):
        if r.AlgorithmName not in _abi.ALGO_NAMES:
            out[i] = KeyError("Not found")
        else:
            groups.setdefault(r.AlgorithmName, []).append(i)
    for name, idx in groups.items():
        algo = NewAlgorithmFactory(name, reqs[idx[0]].SchedulerID, ctx)
        ok = []
        for i in idx:  # a nil Info only breaks the request it is in
            if algo.NeedJobInfo():
                _fill_info(reqs[i], info_lookup)
            if algo.NeedJobInfo() and any(j.Info is None for j in reqs[i].ReadyJobs):
                out[i] = SchedulePanic(_abi.ST_RUNTIME_PANIC, "runtime error: invalid memory address or nil pointer dereference")
            else:
                ok.append(i)
        if ok:
            res = algo.ScheduleBatch([(reqs[i].ReadyJobs, reqs[i].NumGpu) for i in ok])
            for i, r in zip(ok, res):
                out[i] = r
    return out
