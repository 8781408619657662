"""The allocator endpoint through its wire format: AllocationRequest JSON in, JSON map out
(resource_allocator.go:41-111), single and batched, against the known-answer tests."""
import json

import pytest

from helpers import load_kats
from vodascheduler_b200 import wire
from vodascheduler_b200.algorithm import NewBaseJobInfo, SchedulePanic, Time

pytestmark = pytest.mark.gpu
KATS = load_kats()


def _body(kat, with_info=True):
    jobs = []
    for j in kat["jobs"]:
        o = {"job_name": j["name"], "priority": j["priority"],
             "submit_timestamp": wire.format_time(Time(j["submit"])),
             "config": {"num_proc": j["num_proc"], "min_num_proc": j["min"], "max_num_proc": j["max"]},
             "time_metrics": {"first_start_timestamp": wire.format_time(Time(j["first"]))}}
        if with_info:
            o["info"] = {"job_name": j["name"], "estimate_remainning_time_seconds": j["remaining"],
                         "speedup": {str(g): v for g, v in enumerate(j["speedup"])}}
        jobs.append(o)
    return json.dumps({"SchedulerID": "sched-0", "NumGpu": kat["total"], "AlgorithmName": kat["algo"],
                       "ReadyJobs": jobs})


def _wire_kats():
    # first-start stamps beyond year 9999 (types.MaxTime) cannot be marshalled by Go's time.Time
    return [k for k in KATS["alloc"] if all(0 <= j["first"] < 253402300800 for j in k["jobs"])]


@pytest.mark.parametrize("kat", _wire_kats(), ids=lambda k: f"KAT-{k['kat']}-{k['algo']}")
def test_handler_kats(ctx, kat):
    if kat["expect"] is None:
        with pytest.raises(SchedulePanic) as e:
            wire.handle_allocation(_body(kat), ctx)
        assert e.value.status == kat["status"]
    else:
        code, body = wire.handle_allocation(_body(kat), ctx)
        assert code == 200 and json.loads(body) == kat["expect"]


def test_batch_equals_single(ctx):
    kats = _wire_kats()
    bodies = [_body(k) for k in kats] + ['{"AlgorithmName":"Gandiva","NumGpu":1,"ReadyJobs":[]}']
    res = wire.allocate_resource_batch(bodies, ctx)
    assert len(res) == len(bodies)
    for k, r in zip(kats, res):
        if k["expect"] is None:
            assert isinstance(r, SchedulePanic) and r.status == k["status"]
        else:
            assert r == k["expect"]
    assert isinstance(res[-1], KeyError)


def test_nil_info_panics_only_its_own_request(ctx):
    """getJobsInfo writes to the range copy (resource_allocator.go:121,134): a job whose JSON has no
    `info` reaches an info-reading policy with a nil pointer."""
    k = next(k for k in _wire_kats() if k["algo"] == "ElasticSRJF" and k["expect"] is not None)
    with pytest.raises(SchedulePanic, match="nil pointer"):
        wire.allocate_resource(_body(k, with_info=False), ctx)
    res = wire.allocate_resource_batch([_body(k, with_info=False), _body(k)], ctx)
    assert isinstance(res[0], SchedulePanic) and res[1] == k["expect"]
    # Elastic-FIFO never reads Info
    f = next(k for k in _wire_kats() if k["algo"] == "ElasticFIFO" and k["expect"] is not None)
    assert wire.allocate_resource(_body(f, with_info=False), ctx) == f["expect"]


def test_info_lookup_fills_by_index(ctx):
    k = next(k for k in _wire_kats() if k["algo"] == "FfDLOptimizer" and k["expect"] is not None)
    seen = []

    def lookup(job):
        seen.append(job.Name)
        return None  # not in the store -> linear default

    res = wire.allocate_resource(_body(k, with_info=False), ctx, info_lookup=lookup)
    assert seen == [j["name"] for j in k["jobs"]]
    body = json.loads(_body(k, with_info=False))
    for j in body["ReadyJobs"]:
        base = NewBaseJobInfo(j["job_name"])
        j["info"] = {"speedup": base.Speedup, "estimate_remainning_time_seconds": 0.0}
    assert res == wire.allocate_resource(body, ctx)
