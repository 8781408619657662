"""The N>1 host path on CPU: two gloo ranks shard a batch of rounds, each computes its range,
and the gathered result equals the single-process result byte for byte.
The compute function here is the CPU oracle standing in for Context.allocate, because no GPU
exists in this container; the sharding / gathering logic under test is the same."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, algo, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import oracle
    from vodascheduler_b200 import shard, synth
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    b = synth.quirk_batch(algo, rounds=37, jobs_per_round=20, total_gpu=40)
    g, s, sums = shard.run_sharded(b, lambda sub: oracle.allocate(sub, dense=True), rank, world, dist)
    q.put((rank, g, s, sums))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("algo", ["ElasticSRJF", "FfDLOptimizer"])
def test_two_ranks_equal_one(oracle, algo):
    from vodascheduler_b200 import shard, synth
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, algo, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=120) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    b = synth.quirk_batch(algo, rounds=37, jobs_per_round=20, total_gpu=40)
    g_ref, s_ref = oracle.allocate(b, dense=True)
    for rank, g, s, sums in got:
        assert np.array_equal(g, g_ref) and np.array_equal(s, s_ref)
        assert sums.shape == (2, 4)
        assert sums[:, 0].sum() == 37 and sums[:, 1].sum() == int(g_ref.sum())
        assert sums[:, 2].sum() == int((s_ref != 0).sum())
    assert np.array_equal(got[0][3], got[1][3])  # every rank sees the same summaries


def test_shard_bounds_cover_everything():
    from vodascheduler_b200.shard import shard_bounds
    for rounds in (0, 1, 7, 10000):
        for world in (1, 2, 3, 8):
            edges = [shard_bounds(rounds, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == rounds
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
