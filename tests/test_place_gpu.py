"""GPU parity of the placement kernel (voda_place: bestFit + score + munkres + rename) against
the CPU restatement, through the C ABI."""
import numpy as np
import pytest

from helpers import load_kats, place_batch_from
from vodascheduler_b200 import synth

pytestmark = pytest.mark.gpu
KATS = load_kats()


def _same(pb, got, ref):
    assert np.array_equal(got.status[:pb.rounds], ref.status[:pb.rounds])
    for r in range(pb.rounds):
        if ref.status[r] != 0:
            continue
        n0, n1 = int(pb.node_off[r]), int(pb.node_off[r + 1])
        e0 = int(pb.req_off[r]) + n0
        c = int(ref.lay_cnt[r])
        assert got.lay_cnt[r] == c
        for name in ("lay_node", "lay_job", "lay_workers"):
            assert np.array_equal(getattr(got, name)[e0:e0 + c], getattr(ref, name)[e0:e0 + c]), (r, name)
        assert np.array_equal(got.free_slots[n0:n1], ref.free_slots[n0:n1]), r
        assert np.array_equal(got.phys_of_logical[n0:n1], ref.phys_of_logical[n0:n1]), r
        assert got.score[r] == ref.score[r] and got.cross[r] == ref.cross[r]


@pytest.mark.parametrize("kat", KATS["bestfit"], ids=lambda k: f"KAT-{k['kat']}")
def test_bestfit_kat(ctx, kat):
    pb = place_batch_from([dict(nodes=kat["nodes"], requests=kat["requests"])])
    out = ctx.place(pb)
    assert out.status[0] == kat["status"]
    if kat["layout"] is not None:
        c = int(out.lay_cnt[0])
        got = [[int(out.lay_node[t]), int(out.lay_job[t]), int(out.lay_workers[t])] for t in range(c)]
        assert got == kat["layout"]
        assert out.free_slots[:len(kat["nodes"])].tolist() == kat["free"]


@pytest.mark.parametrize("nodes,slots,jobs,load", [(4, 4, 6, 0.9), (16, 8, 30, 0.85), (37, 8, 70, 0.95),
                                                   (128, 8, 200, 0.85), (128, 8, 400, 1.2)])
def test_random_rounds(ctx, oracle, nodes, slots, jobs, load):
    pb = synth.place_batch(rounds=12, nodes=nodes, slots=slots, jobs_per_round=jobs, load=load)
    _same(pb, ctx.place(pb), oracle.place(pb, threads=8))


def test_oversubscribed_and_quirks(ctx, oracle):
    """Requests beyond the cluster: split jobs (Q11 over-assignment), early return when no slot
    is left, and the never-ending loop (Q12 -> VODA_ST_NONTERMINATING)."""
    rounds = [
        dict(nodes=[8, 8], requests=[6, 6, 4], prev=[[(0, 6)], [(1, 6), (2, 2)]]),
        dict(nodes=[8, 8], requests=[10, 3]),
        dict(nodes=[8, 8], requests=[16, 1]),
        dict(nodes=[4, 4, 4], requests=[5, 5, 2], prev=[[(0, 4)], [(0, 1), (1, 3)], [(2, 2)]]),
        dict(nodes=[2, 2, 2, 2], requests=[0, 3, 0, 1, 1, 1], prev=[[(1, 2)], [(3, 1)], [], [(5, 1)]]),
        dict(nodes=[], requests=[1, 2]),
        dict(nodes=[8], requests=[]),
        dict(nodes=[0, 0], requests=[1]),
    ]
    pb = place_batch_from(rounds)
    got, ref = ctx.place(pb), oracle.place(pb)
    _same(pb, got, ref)
    assert 7 in ref.status[:len(rounds)].tolist()


def test_ragged_batch(ctx, oracle):
    rng = np.random.default_rng(5)
    rounds = []
    for _ in range(60):
        n = int(rng.integers(1, 20))
        m = int(rng.integers(0, 30))
        nodes = [int(rng.integers(1, 9)) for _ in range(n)]
        req = [int(rng.integers(0, 7)) for _ in range(m)]
        prev = [[(int(rng.integers(0, max(1, m))), int(rng.integers(1, 4))) for _ in range(int(rng.integers(0, 3)))]
                if m else [] for _ in range(n)]
        prev = [list({j: (j, w) for j, w in p}.values()) for p in prev]  # one entry per job per node
        rounds.append(dict(nodes=nodes, requests=req, prev=prev))
    pb = place_batch_from(rounds)
    _same(pb, ctx.place(pb), oracle.place(pb))


def _pods_from_prev(pb):
    """A pod table consistent with the previous layout of a synthetic PlaceBatch: every job's worker
    pods sit, in pod-index order, on the nodes that hold it (by physical index)."""
    off, node, slots = [0], [], []
    for r in range(pb.rounds):
        n0, n1 = int(pb.node_off[r]), int(pb.node_off[r + 1])
        per_job = {}
        for c in range(n0, n1):
            for e in range(int(pb.prev_off[c]), int(pb.prev_off[c + 1])):
                per_job.setdefault(int(pb.prev_job[e]), []).append((c - n0, int(pb.prev_workers[e])))
        for q in range(int(pb.req_off[r + 1]) - int(pb.req_off[r])):
            for c, w in per_job.get(q, []):
                node.append(c)
                slots.append(w)
            off.append(len(node))
    return np.array(off, np.int32), np.array(node, np.int32), np.array(slots, np.int32)


def test_migration_counts_match_oracle(ctx, oracle):
    """migrated_workers / migrated_launchers of the kernel (updateJobStates + updatePodNodeName as
    counts, placement_manager.go:548-617) against the CPU restatement, over batched rounds."""
    pb = synth.place_batch(rounds=300, nodes=24, slots=8, jobs_per_round=40)
    pb.pods_off, pb.pods_node, pb.pods_slots = _pods_from_prev(pb)
    ref = oracle.place(pb, threads=4)
    out = ctx.place(pb)
    assert np.array_equal(out.status, ref.status)
    ok = ref.status == 0
    assert np.array_equal(out.migrated_workers[ok], ref.migrated_workers[ok])
    assert np.array_equal(out.migrated_launchers[ok], ref.migrated_launchers[ok])
    assert ref.migrated_workers[ok].sum() > 0 and ref.migrated_launchers[ok].sum() > 0
    # without a pod table nothing is counted and nothing else changes
    pb.pods_off = pb.pods_node = pb.pods_slots = None
    plain = ctx.place(pb)
    assert np.array_equal(plain.phys_of_logical, out.phys_of_logical) and not plain.migrated_workers.any()


def test_migration_counts_match_bookkeeping_history(ctx):
    """A history of rounds: the kernel's counts equal updatePodNodeName of the bookkeeping
    restatement fed with the kernel's own bindings."""
    from oracle import bookkeeping as bk
    from test_bookkeeping_cpu import _random_rounds, place_round
    nodes, slots = 6, 8
    ref = bk.Manager(nodes, slots)
    for req in _random_rounds(9, nodes, slots, 80):
        got, want = place_round(ctx.place, ref, req, nodes, slots)
        assert got == want


@pytest.mark.parametrize("scale", [1, 40, 300, 20000], ids=["u8", "u16-low", "u16", "i32"])
@pytest.mark.parametrize("nodes", [16, 37, 128])
def test_score_ranges_pick_the_cell_width(ctx, oracle, scale, nodes):
    """The munkres step inside the placement kernel stores rowmax - score in 8-, 16- or 32-bit cells
    depending on the score range of the round (scores are worker counts): every class, at node
    counts that are and are not multiples of 4, against the oracle."""
    pb = synth.place_batch(rounds=10, nodes=nodes, slots=8, jobs_per_round=nodes + 20, load=0.9)
    pb.node_slots = (pb.node_slots * scale).astype(pb.node_slots.dtype)
    pb.req_workers = (pb.req_workers * scale).astype(pb.req_workers.dtype)
    pb.prev_workers = (pb.prev_workers * scale).astype(pb.prev_workers.dtype)
    ref = oracle.place(pb, threads=8)
    assert int(ref.score.max()) >= 8 * scale  # the class is really reached
    _same(pb, ctx.place(pb), ref)


@pytest.mark.parametrize("split", ["0", "1"], ids=["fused", "split"])
def test_split_and_fused_forms_agree_with_the_oracle(ctx, oracle, monkeypatch, split):
    """Large batches without a pod table run as two kernels (score matrix, then a warp per round for the
    assignment); VODA_PLACE_SPLIT forces either form.  Shapes: ragged node counts, every cell width,
    refused and non-terminating rounds among solvable ones."""
    monkeypatch.setenv("VODA_PLACE_SPLIT", split)
    for nodes, slots, jobs, load, scale in ((16, 8, 30, 0.85, 1), (37, 8, 70, 0.95, 40), (128, 8, 200, 0.85, 1),
                                            (128, 8, 160, 0.9, 300), (61, 4, 90, 1.2, 20000)):
        pb = synth.place_batch(rounds=24, nodes=nodes, slots=slots, jobs_per_round=jobs, load=load)
        pb.node_slots = (pb.node_slots * scale).astype(pb.node_slots.dtype)
        pb.req_workers = (pb.req_workers * scale).astype(pb.req_workers.dtype)
        pb.prev_workers = (pb.prev_workers * scale).astype(pb.prev_workers.dtype)
        _same(pb, ctx.place(pb), oracle.place(pb, threads=8))
    rounds = [
        dict(nodes=[8, 8], requests=[6, 6, 4], prev=[[(0, 6)], [(1, 6), (2, 2)]]),
        dict(nodes=[8, 8], requests=[16, 1]),
        dict(nodes=[], requests=[1, 2]),
        dict(nodes=[4, 4, 4], requests=[5, 5, 2], prev=[[(0, 4)], [(0, 1), (1, 3)], [(2, 2)]]),
        dict(nodes=[8], requests=[]),
        dict(nodes=[0, 0], requests=[1]),
    ]
    pb = place_batch_from(rounds)
    _same(pb, ctx.place(pb), oracle.place(pb))


def test_split_form_large_batch(ctx, oracle):
    """A batch large enough to take the split form by itself (and more than one 1 GiB sub-batch of
    slabs would need only at 16k+ rounds of 128 nodes: 6 000 rounds stay in one)."""
    pb = synth.place_batch(rounds=6000, nodes=32, slots=8, jobs_per_round=50, load=0.9)
    _same(pb, ctx.place(pb), oracle.place(pb, threads=16))


def test_chunked_host_path_equals_single_shot(ctx, oracle):
    """voda_place on a batch large enough for the two-stream chunked path (at least eight waves of
    rounds) gives what the oracle computes, ragged round sizes included."""
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    rounds = 2 * max(2 * 10 * sms, 1) * 2 + 37  # >= 2 chunks of >= 2 * 10 * sms rounds, not a multiple
    pb = synth.place_batch(rounds=rounds, nodes=12, slots=8, jobs_per_round=20, load=0.9)
    _same(pb, ctx.place(pb), oracle.place(pb, threads=16))


@pytest.mark.parametrize("split", ["0", "1"], ids=["fused", "split"])
@pytest.mark.parametrize("nodes", [3, 4, 16, 37, 64])
def test_negative_previous_counts_follow_the_reference_wrap(ctx, oracle, monkeypatch, split, nodes):
    """A previous layout with negative worker counts puts negative scores next to non-negative ones in a
    row of the score matrix; the reference's MaxInt64 - score then wraps (test_munkres_wrap_cpu.py) and
    the binding it finds is what the kernel has to find too.  Negative node capacities ride along."""
    monkeypatch.setenv("VODA_PLACE_SPLIT", split)
    rng = np.random.default_rng(40 + nodes)
    for it in range(4):
        pb = synth.place_batch(rounds=20, nodes=nodes, slots=int(rng.choice([2, 4, 8])), jobs_per_round=nodes + 4,
                               load=float(rng.choice([0.6, 0.9, 1.1])), seed=900 + 7 * nodes + it)
        m = rng.random(pb.prev_workers.shape[0]) < 0.15
        pb.prev_workers = np.where(m, -pb.prev_workers - int(rng.integers(0, 3)), pb.prev_workers).astype(pb.prev_workers.dtype)
        if it % 2:
            m = rng.random(pb.node_slots.shape[0]) < 0.1
            pb.node_slots = np.where(m, -int(rng.integers(0, 3)), pb.node_slots).astype(pb.node_slots.dtype)
        _same(pb, ctx.place(pb), oracle.place(pb, threads=8))
