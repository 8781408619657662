"""oracle/bookkeeping.py against hand-traced cases of the reference's releaseSlots
(placement_manager.go:337-411), updateJobStates (:548-566) and updatePodNodeName (:571-617).
The expected values were traced by hand from the Go text (not produced by running Go: no toolchain
here, and not by running this repository's code)."""
from oracle import bookkeeping as bk


def _state():
    """2 nodes x 4 slots.  A holds 4 on node 0 and 2 on node 1, B holds 2 on node 1 (cluster full)."""
    m = bk.Manager(2, 4)
    m.bind([(0, "A", 4), (1, "A", 2), (1, "B", 2)], [0, 0], [0, 1])
    m.update_job_states()
    assert m.update_pod_node_name() == (0, 0)  # no pod existed before: nothing migrates
    return m


def test_initial_tables():
    m = _state()
    assert m.jobStates["A"].nodeNumSlotsList == [[0, 4], [1, 2]] and m.jobStates["A"].numWorkers == 6
    assert m.jobStates["B"].nodeNumSlotsList == [[1, 2]]
    assert [m.podNodeName[("A", i)] for i in range(6)] == [0, 0, 0, 0, 1, 1]


def test_scale_down_across_two_nodes():
    """A: 6 -> 3.  toRelease = 3; the last node (1) holds 2 < 3: all of it goes (free 0 -> 2, A leaves
    node 1, the list entry is dropped), 1 left to release; node 0 holds 4 >= 1: 4 -> 3, free 0 -> 1."""
    m = _state()
    m.release_slots({"A": 3, "B": 2})
    assert m.nodeStates[0].jobNumWorkers == {"A": 3} and m.nodeStates[0].freeSlots == 1
    assert m.nodeStates[1].jobNumWorkers == {"B": 2} and m.nodeStates[1].freeSlots == 2
    assert m.jobStates["A"].nodeNumSlotsList == [[0, 3]]
    assert m.jobStates["B"].nodeNumSlotsList == [[1, 2]]  # untouched: 2 is not below 2


def test_partial_release_inside_the_last_node():
    """A: 6 -> 5: the last node holds 2 >= 1, so 2 -> 1 there and nothing else moves."""
    m = _state()
    m.release_slots({"A": 5, "B": 2})
    assert m.nodeStates[1].jobNumWorkers == {"A": 1, "B": 2} and m.nodeStates[1].freeSlots == 1
    assert m.jobStates["A"].nodeNumSlotsList == [[0, 4], [1, 1]]


def test_job_removed():
    """B is not in the request map any more (terminated): all its slots come back, it leaves its
    node; A is unchanged."""
    m = _state()
    m.release_slots({"A": 6})
    assert m.nodeStates[1].jobNumWorkers == {"A": 2} and m.nodeStates[1].freeSlots == 2
    assert m.jobStates["B"].nodeNumSlotsList == []
    assert m.jobStates["A"].nodeNumSlotsList == [[0, 4], [1, 2]]


def test_scale_to_zero_keeps_the_request_branch():
    """A job scheduled with 0 GPUs is IN the map (JobScheduleResult holds every ready job), so it
    takes the scale-down branch: 2 -> 0 on its only node."""
    m = _state()
    m.release_slots({"A": 6, "B": 0})
    assert m.nodeStates[1].jobNumWorkers == {"A": 2} and m.nodeStates[1].freeSlots == 2
    assert m.jobStates["B"].nodeNumSlotsList == []


def test_node_missing():
    """A node deleted between rounds (placement_manager.go:346-352, 389-392): its slots are not
    given back to anything, the job's list entry still goes."""
    m = _state()
    del m.nodeStates[1]
    m.release_slots({"A": 3})
    assert m.nodeStates[0].jobNumWorkers == {"A": 3} and m.nodeStates[0].freeSlots == 1
    assert m.jobStates["A"].nodeNumSlotsList == [[0, 3]]
    assert m.jobStates["B"].nodeNumSlotsList == []


def test_rename_and_migration_counts():
    """Next round: bestFit put A(3) on logical node 0 and B(2) on logical node 1, and the assignment
    bound logical 0 -> physical 1, logical 1 -> physical 0.  A's pods 0..2 were on node 0 and are now on
    node 1: 3 workers deleted = all of A's workers, so its launcher goes too.  B's pods 0..1 were on node
    1, now on node 0: 2 workers + its launcher."""
    m = _state()
    m.release_slots({"A": 3, "B": 2})
    m.bind([(0, "A", 3), (1, "B", 2)], [1, 2], [1, 0])
    assert m.nodeStates[1].jobNumWorkers == {"A": 3} and m.nodeStates[0].jobNumWorkers == {"B": 2}
    m.update_job_states()
    assert m.jobStates["A"].nodeNumSlotsList == [[1, 3]] and m.jobStates["B"].nodeNumSlotsList == [[0, 2]]
    assert m.update_pod_node_name() == (5, 2)
    assert sorted(m.podNodeName) == [("A", 0), ("A", 1), ("A", 2), ("B", 0), ("B", 1)]


def test_identity_binding_migrates_nothing_and_new_pods_are_free():
    """Same binding, A scaled out by one on node 1 (pod 6 is new): no pod changes node."""
    m = _state()
    m.release_slots({"A": 6, "B": 1})
    m.bind([(0, "A", 4), (1, "A", 3), (1, "B", 1)], [0, 0], [0, 1])
    m.update_job_states()
    assert m.update_pod_node_name() == (0, 0)
    assert m.podNodeName[("A", 6)] == 1


def test_partial_move_keeps_the_launcher():
    """A keeps node 0 for 4 workers and moves its other 2 from node 1 to a third place: only those two
    pods are deleted, deleted (2) != numWorkers (6), the launcher stays."""
    m = bk.Manager(3, 4)
    m.bind([(0, "A", 4), (1, "A", 2)], [0, 2, 4], [0, 1, 2])
    m.update_job_states()
    m.update_pod_node_name()
    m.release_slots({"A": 6})
    m.bind([(0, "A", 4), (1, "A", 2)], [0, 2, 4], [0, 2, 1])  # logical 1 now binds physical node 2
    m.update_job_states()
    assert m.jobStates["A"].nodeNumSlotsList == [[0, 4], [2, 2]]
    assert m.update_pod_node_name() == (2, 0)


def test_cpu_replay_runs_to_completion(oracle):
    """oracle/replay.py (the checker of voda_replay) on a small trace: every job completes, the
    cluster is never over-subscribed by a round that validateResult accepted, history is reproducible."""
    from oracle import replay as orp
    from vodascheduler_b200 import replay as rp
    r = rp.make_replicas(1, jobs=60, nodes=4, slots=8)[0]
    a = orp.replay_one(r.trace, r.nodes, r.slots, r.rate_limit)
    b = orp.replay_one(r.trace, r.nodes, r.slots, r.rate_limit, dense=False)
    assert a == b                      # dense and structure-faithful allocation ports agree round for round
    assert a["done"] == 60 and a["rounds"] > 5 and a["actions"]["starts"] >= 60


def _random_rounds(seed, nodes, slots, rounds):
    """A random history of request maps that stays clear of the reference's endless loop."""
    import numpy as np
    rng = np.random.default_rng(seed)
    jobs, nxt = {}, 0
    for _ in range(rounds):
        for j in list(jobs):
            u = rng.random()
            if u < 0.15:
                del jobs[j]
            elif u < 0.5:
                jobs[j] = int(rng.integers(0, 7))
        for _ in range(int(rng.integers(0, 3))):
            jobs[f"job{nxt}"] = int(rng.integers(1, 7))
            nxt += 1
        while sum(jobs.values()) > int(0.8 * nodes * slots):
            jobs[max(jobs, key=jobs.get)] -= 1
        yield dict(jobs)


def place_round(place_fn, ref, req, nodes, slots):
    """One Place() through `place_fn` (oracle.place or Context.place) with the pod table attached,
    applied to the bookkeeping restatement `ref`; returns (counts from the placement call, counts from
    updatePodNodeName of the restatement)."""
    import numpy as np
    from vodascheduler_b200.batch import PlaceBatch
    names = list(req)
    pods = bk.pod_table(ref, names)  # the table the round starts from
    ref.release_slots(req)
    index = {n: q for q, n in enumerate(names)}
    off, pj, pw = [0], [], []
    for ent in ref.previous_layout(index):
        for q, w in ent:
            pj.append(q)
            pw.append(w)
        off.append(len(pj))
    out = place_fn(PlaceBatch([0, nodes], [slots] * nodes, [0, len(names)], [req[n] for n in names], off,
                              np.array(pj, np.int64), np.array(pw, np.int64), pods_off=pods[0],
                              pods_node=np.array(pods[1], np.int64), pods_slots=np.array(pods[2], np.int64)))
    assert out.status[0] == 0
    lay = [(int(out.lay_node[t]), names[int(out.lay_job[t])], int(out.lay_workers[t])) for t in range(int(out.lay_cnt[0]))]
    ref.bind(lay, [int(x) for x in out.free_slots[:nodes]], [int(x) for x in out.phys_of_logical[:nodes]])
    ref.update_job_states()
    return (int(out.migrated_workers[0]), int(out.migrated_launchers[0])), ref.update_pod_node_name()


def test_oracle_place_migration_counts_equal_bookkeeping(oracle):
    """voda_oracle_place's migrated_workers / migrated_launchers (computed from the binding and the
    run-length pod table) against updatePodNodeName of the bookkeeping restatement, round after round."""
    nodes, slots = 6, 8
    ref = bk.Manager(nodes, slots)
    moved = 0
    for req in _random_rounds(5, nodes, slots, 80):
        got, want = place_round(oracle.place, ref, req, nodes, slots)
        assert got == want
        moved += want[0]
    assert moved > 0
