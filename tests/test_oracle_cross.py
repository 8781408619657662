"""Cross-checks inside the oracle (CPU only): the structure-faithful restatement against the
dense one on randomised and quirk-heavy rounds, the validateResult invariants, FfDL against
brute force, munkres against scipy's optimum."""
import itertools

import numpy as np
import pytest
from scipy.optimize import linear_sum_assignment

from vodascheduler_b200 import ALGO_NAMES, synth


@pytest.mark.parametrize("algo", ALGO_NAMES)
def test_faithful_equals_dense_random(oracle, algo):
    b = synth.jobs_batch(algo, rounds=60, jobs_per_round=40, total_gpu=96, ragged=True)
    g0, s0 = oracle.allocate(b)
    g1, s1 = oracle.allocate(b, dense=True, threads=3)
    assert np.array_equal(s0, s1)
    assert np.array_equal(g0, g1)
    assert (s0 == 0).mean() > 0.5


@pytest.mark.parametrize("algo", ALGO_NAMES)
def test_faithful_equals_dense_quirks(oracle, algo):
    b = synth.quirk_batch(algo, rounds=150, jobs_per_round=24, total_gpu=40)
    g0, s0 = oracle.allocate(b, threads=2)
    g1, s1 = oracle.allocate(b, dense=True)
    assert np.array_equal(s0, s1)
    assert np.array_equal(g0, g1)


@pytest.mark.parametrize("algo", ALGO_NAMES)
def test_faithful_equals_dense_min_zero(oracle, algo):
    """MinNumProc == 0: ElasticTiresias reaches the pick that changes nothing (the Go loop never
    ends there); both restatements cut it at the same moment with status 7."""
    b = synth.min_zero_batch(algo, rounds=120, jobs_per_round=24, total_gpu=40)
    g0, s0 = oracle.allocate(b, threads=2)
    g1, s1 = oracle.allocate(b, dense=True)
    assert np.array_equal(s0, s1)
    assert np.array_equal(g0, g1)
    if algo == "ElasticTiresias":
        assert (s0 == 7).any() and (s0 == 0).any()


@pytest.mark.parametrize("algo", ALGO_NAMES)
def test_validate_invariants(oracle, algo):
    """Rounds with status 0 satisfy what validateResult (utils.go:18-42) checks."""
    b = synth.quirk_batch(algo, rounds=100, jobs_per_round=30, total_gpu=64)
    g, st = oracle.allocate(b, dense=True)
    for r in range(b.rounds):
        if st[r] != 0:
            continue
        sl = slice(b.job_off[r], b.job_off[r + 1])
        x = g[sl]
        assert (x >= 0).all() and x.sum() <= b.total_gpu[r]
        assert ((x == 0) | (x >= b.min_proc[sl])).all() and (x <= b.max_proc[sl]).all()


def test_ffdl_optimal_vs_bruteforce(oracle):
    b = synth.jobs_batch("FfDLOptimizer", rounds=40, jobs_per_round=4, total_gpu=9)
    b.max_proc = np.minimum(b.max_proc, 5).astype(np.int32)
    g, st = oracle.allocate(b)
    S = b.speedup.reshape(-1, synth.W)
    for r in range(b.rounds):
        o = int(b.job_off[r])
        assert st[r] == 0
        best = max(
            sum(float(S[o + i, a]) for i, a in enumerate(alloc) if a > 0)
            for alloc in itertools.product(*[range(0, int(b.max_proc[o + i]) + 1) for i in range(4)])
            if sum(alloc) <= 9)
        got = sum(float(S[o + i, g[o + i]]) for i in range(4) if g[o + i] > 0)
        assert abs(got - best) < 1e-4


@pytest.mark.parametrize("n,kind,hi", [(5, "uniform", 3), (16, "uniform", 8), (32, "placement", 8),
                                       (64, "uniform", 999), (128, "placement", 8)])
def test_munkres_optimal(oracle, n, kind, hi):
    A = synth.matrices(6, n, kind=kind, hi=hi)
    cols, st = oracle.munkres_max(A, threads=2)
    for k in range(len(A)):
        assert sorted(cols[k].tolist()) == list(range(n))
        ri, ci = linear_sum_assignment(A[k], maximize=True)
        assert A[k][np.arange(n), cols[k]].sum() == A[k][ri, ci].sum()


def test_place_matches_parts(oracle):
    """voda_oracle_place == bestFit layout + score matrix + munkres done by hand."""
    pb = synth.place_batch(rounds=8, nodes=12, slots=8, jobs_per_round=20)
    out = oracle.place(pb)
    for r in range(pb.rounds):
        assert out.status[r] == 0
        n0, n = int(pb.node_off[r]), 12
        e0 = int(pb.req_off[r]) + n0
        lay = [dict() for _ in range(n)]
        for t in range(int(out.lay_cnt[r])):
            lay[out.lay_node[e0 + t]][int(out.lay_job[e0 + t])] = int(out.lay_workers[e0 + t])
        M = np.zeros((n, n), np.int64)
        for c in range(n):
            prev = {int(pb.prev_job[e]): int(pb.prev_workers[e])
                    for e in range(pb.prev_off[n0 + c], pb.prev_off[n0 + c + 1])}
            for i in range(n):
                M[i, c] = sum(min(w, prev.get(j, 0)) for j, w in lay[i].items())
        cols, _ = oracle.munkres_max(M[None])
        assert cols[0].tolist() == out.phys_of_logical[n0:n0 + n].tolist()
        assert int(M[np.arange(n), cols[0]].sum()) == int(out.score[r])
