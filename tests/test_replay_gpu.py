"""Trace replay (BASELINE configs[4] in miniature).  voda_replay (GPU kernels for every allocation
and placement decision, C++ bookkeeping around them) must take exactly the decisions of the CPU
replay built from the oracles (oracle/replay.py: oracle.allocate + oracle.place +
oracle/bookkeeping.py): same rounds, completions, migrations, actions and the same digest of every
round's allocations and bindings."""
import numpy as np
import pytest

from vodascheduler_b200 import replay as rp

pytestmark = pytest.mark.gpu

KEYS = ("done", "rounds", "sim_time", "jct_sum", "migrated_workers", "migrated_launchers", "placements_skipped",
        "actions", "history")


def test_replay_equals_cpu_replay(ctx, oracle):
    from oracle import replay as orp
    reps = rp.make_replicas(4, jobs=150, nodes=8, slots=8)
    stats = rp.replay(reps, ctx)
    assert stats["rounds"] == sum(s["rounds"] for s in stats["replicas"])
    for r, s in zip(reps, stats["replicas"]):
        ref = orp.replay_one(r.trace, r.nodes, r.slots, r.rate_limit)
        assert s["done"] == 150
        for k in KEYS:
            assert s[k] == ref[k], (r.idx, k, s[k], ref[k])
        assert s["rounds"] > 10 and s["mean_jct"] > 0 and s["migrated_workers"] > 0
        assert s["actions"]["starts"] >= 150  # every job starts at least once


def test_replay_is_replica_independent_and_thread_independent(ctx):
    """A replica replayed alone, or with another number of bookkeeping threads, gives the same history
    as inside a batch (rounds are independent across replicas)."""
    reps = rp.make_replicas(6, jobs=120, nodes=8, slots=8)
    stats = rp.replay(reps, ctx, host_threads=3)
    alone = rp.make_replicas(1, jobs=120, nodes=8, slots=8, first=4)
    s1 = rp.replay(alone, ctx, host_threads=1)["replicas"][0]
    assert s1 == stats["replicas"][4]
    again = rp.replay(reps, ctx, host_threads=1)
    assert again["replicas"] == stats["replicas"]


def test_replay_max_ticks_and_empty(ctx):
    reps = rp.make_replicas(2, jobs=80, nodes=4, slots=8)
    part = rp.replay(reps, ctx, max_ticks=50)
    assert part["ticks"] == 50 and all(s["done"] < 80 for s in part["replicas"])
    assert rp.replay([], ctx)["rounds"] == 0


def test_placement_manager_equals_bookkeeping_oracle(ctx, oracle):
    """The host mirror of PlacementManager (vodascheduler_b200/placement.py, GPU placement) against
    the bookkeeping restatement driven by oracle.place, over a random sequence of rounds: scale-outs,
    scale-downs, jobs leaving and arriving."""
    from oracle import bookkeeping as bk
    from vodascheduler_b200.batch import PlaceBatch
    from vodascheduler_b200.placement import PlacementManager
    rng = np.random.default_rng(11)
    nodes, slots = 6, 8
    pm = PlacementManager({f"n{i}": slots for i in range(nodes)}, ctx=ctx)
    ref = bk.Manager(nodes, slots)
    jobs = {}
    nxt = 0
    for rnd in range(60):
        for j in list(jobs):           # leave / resize
            u = rng.random()
            if u < 0.15:
                del jobs[j]
            elif u < 0.5:
                jobs[j] = int(rng.integers(0, 7))
        for _ in range(int(rng.integers(0, 3))):
            jobs[f"job{nxt}"] = int(rng.integers(1, 7))
            nxt += 1
        while sum(jobs.values()) > int(0.8 * nodes * slots):   # stay clear of the reference's endless loop
            jobs[max(jobs, key=jobs.get)] -= 1
        req = dict(jobs)
        deleting = pm.Place(req)
        # the same round through the oracles
        ref.release_slots(req)
        names = list(req)
        index = {n: q for q, n in enumerate(names)}
        prev = ref.previous_layout(index)
        off, pj, pw = [0], [], []
        for ent in prev:
            for q, w in ent:
                pj.append(q)
                pw.append(w)
            off.append(len(pj))
        out = oracle.place(PlaceBatch([0, nodes], [slots] * nodes, [0, len(names)], [req[n] for n in names], off,
                                      np.array(pj, np.int64), np.array(pw, np.int64)))
        assert out.status[0] == 0
        lay = [(int(out.lay_node[t]), names[int(out.lay_job[t])], int(out.lay_workers[t])) for t in range(int(out.lay_cnt[0]))]
        ref.bind(lay, [int(x) for x in out.free_slots[:nodes]], [int(x) for x in out.phys_of_logical[:nodes]])
        ref.update_job_states()
        dw, dl = ref.update_pod_node_name()
        assert (pm.deletedWorkers, pm.deletedLaunchers) == (dw, dl), rnd
        assert len(deleting) == dw + dl
        for c in range(nodes):
            nd = pm.nodeStates[f"n{c}"]
            assert dict(nd.jobNumWorkers) == ref.nodeStates[c].jobNumWorkers and nd.freeSlots == ref.nodeStates[c].freeSlots
        assert {j: [(int(x.node[1:]), x.numSlots) for x in js.nodeNumSlotsList] for j, js in pm.jobStates.items()} == \
               {j: [tuple(e) for e in js.nodeNumSlotsList] for j, js in ref.jobStates.items()}


def test_replay_parallel_contexts_equal_single(ctx):
    """Replicas dealt to several contexts (one host thread each) give the per-replica results of one
    context, in the same order."""
    from vodascheduler_b200 import Context
    reps = rp.make_replicas(7, jobs=100, nodes=8, slots=8)
    one = rp.replay(reps, ctx)
    extra = [Context(0) for _ in range(2)]
    try:
        par = rp.replay_parallel(reps, [ctx] + extra)
    finally:
        for c in extra:
            c.close()
    assert par["replicas"] == one["replicas"] and par["rounds"] == one["rounds"]
