"""The host-side mirrors of the reference interfaces (SchedulerAlgorithm / NewAlgorithmFactory,
munkres.ComputeMunkresMax, PlacementManager.Place) driven the way the reference's callers
drive them (resource_allocator.go:89-106, placement_manager.go:505-512, scheduler.go:355-358)."""
import pytest

from helpers import load_kats
from vodascheduler_b200 import ALGO_NAMES
from vodascheduler_b200.algorithm import (JobConfig, JobInfo, JobMetrics, MaxTime, NewAlgorithmFactory,
                                          NewBaseJobInfo, SchedulePanic, Time, TrainingJob)
from vodascheduler_b200.placement import PlacementManager, munkres

pytestmark = pytest.mark.gpu
KATS = load_kats()


def _jobs(kat):
    out = []
    for j in kat["jobs"]:
        info = JobInfo(Name=j["name"], EstimatedRemainningTimeSec=j["remaining"],
                       Speedup={str(g): float(v) for g, v in enumerate(j["speedup"])})
        out.append(TrainingJob(j["name"], j["priority"], Time(j["submit"]),
                               JobConfig(j["num_proc"], j["min"], j["max"]), JobMetrics(Time(j["first"])), info))
    return out


def test_factory_names_and_flags(ctx):
    for name in ALGO_NAMES:
        a = NewAlgorithmFactory(name, "sched-0", ctx)
        assert a.GetName() == name
    assert not NewAlgorithmFactory("FIFO", "s", ctx).NeedJobInfo()
    assert NewAlgorithmFactory("AFS-L", "s", ctx).NeedJobInfo()
    with pytest.raises(KeyError, match="Not found"):
        NewAlgorithmFactory("Gandiva", "s", ctx)


@pytest.mark.parametrize("kat", KATS["alloc"], ids=lambda k: f"KAT-{k['kat']}-{k['algo']}")
def test_schedule_kats(ctx, kat):
    algo = NewAlgorithmFactory(kat["algo"], "sched-0", ctx)
    if kat["expect"] is None:
        with pytest.raises(SchedulePanic) as e:
            algo.Schedule(_jobs(kat), kat["total"])
        assert e.value.status == kat["status"]
        want = "would not terminate" if kat["status"] == 7 else "Invalid GPU allocations"
        assert want in str(e.value)
    else:
        assert algo.Schedule(_jobs(kat), kat["total"]) == kat["expect"]


def test_default_job_info_is_linear(ctx):
    """NewBaseJobInfo (trainingjob.go:168-186): linear curves, so FfDL's tie rule decides:
    the DP keeps the smallest g on ties, which hands the surplus to the EARLIEST job."""
    jobs = [TrainingJob(f"j{i}", 0, Time(i), JobConfig(1, 1, 4), JobMetrics(MaxTime), NewBaseJobInfo(f"j{i}"))
            for i in range(3)]
    res = NewAlgorithmFactory("FfDLOptimizer", "s", ctx).Schedule(jobs, 6)
    assert res == {"j0": 4, "j1": 1, "j2": 1}


def test_compute_munkres_max(ctx):
    for kat in KATS["munkres"]:
        n = len(kat["A"])
        m = munkres.NewMatrix(n)
        m.A = [v for row in kat["A"] for v in row]
        res = munkres.ComputeMunkresMax(m, ctx)
        assert [rc.Row for rc in res] == list(range(n))
        assert [rc.Col for rc in res] == kat["cols"]
        assert sum(m.A[n * rc.Row + rc.Col] for rc in res) == kat["score"]


def test_place_keeps_jobs_where_they_are(ctx):
    """Two rounds with the same requests: the second one must migrate nothing, because the
    Hungarian binding maps every logical node back onto the physical node it came from."""
    pm = PlacementManager({"gpu1": 8, "gpu2": 8, "gpu3": 8}, ctx)
    assert pm.Place({"A": 6, "B": 6, "C": 4, "D": 2}) == []  # nothing existed, nothing to delete
    used = sum(sum(nd.jobNumWorkers.values()) for nd in pm.nodeStates.values())
    assert used == 18
    before = dict(pm.podNodeName)
    assert pm.Place({"A": 6, "B": 6, "C": 4, "D": 2}) == []
    assert pm.podNodeName == before and pm.lastScore == 18
    # scale A down, drop D, add E: released slots are reused and survivors stay put
    deleted = pm.Place({"A": 4, "B": 6, "C": 4, "E": 3})
    assert "D" not in pm.jobStates and pm.jobStates["E"].numWorkers == 3
    for pod in deleted:
        assert pod.split("-")[0] in ("A", "B", "C")
    for nd in pm.nodeStates.values():
        assert nd.freeSlots == nd.totalSlots - sum(nd.jobNumWorkers.values())
