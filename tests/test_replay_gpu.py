"""Trace replay (BASELINE configs[4] in miniature): replicas stepped together, every job
completes, the cluster is never over-subscribed, and a replica replayed alone gives the same
history as inside a batch of replicas (rounds are independent across replicas)."""
import pytest

from vodascheduler_b200 import replay as rp

pytestmark = pytest.mark.gpu


def test_replay_completes_and_is_replica_independent(ctx):
    reps = rp.make_replicas(3, jobs=120, nodes=8, slots=8)
    stats = rp.replay(reps, ctx, max_ticks=20000)
    for r, s in zip(reps, stats["replicas"]):
        assert s["done"] == 120 and r.finished()
        assert s["rounds"] > 10 and s["mean_jct"] > 0
    alone = rp.make_replicas(1, jobs=120, nodes=8, slots=8, first=1)
    s1 = rp.replay(alone, ctx, max_ticks=20000)["replicas"][0]
    assert s1 == stats["replicas"][1]


def test_replay_never_oversubscribes(ctx):
    reps = rp.make_replicas(2, jobs=60, nodes=4, slots=8)
    seen = []

    def watch(ticks, replicas):
        for r in replicas:
            seen.append(sum(r.alloc.values()) <= r.total_gpu)

    rp.replay(reps, ctx, max_ticks=5000, progress=lambda t, rs: watch(t, rs))
    assert all(seen)
