"""Synthetic inputs for the hot path (the reference ships none).

Counter-based PRNG so any (seed, round, job, field) value can be regenerated anywhere:
    x = splitmix64(seed * 0x9E3779B97F4A7C15 + ((round << 20 | job) * 16 + field))
    u = top 24 bits of x / 2**24
Workload shapes follow BASELINE.json `configs`.
"""
from __future__ import annotations

import numpy as np

from .batch import AllocBatch, PlaceBatch

SEED = 20261106
W = 34  # dense speedup-table width: indices 0..33 (maxNumGpu+1, trainingjob.go:13,173)
MAXTIME_SEC = (1 << 63) - 62135596801  # types.MaxTime, pkg/common/types/types.go:65
MAXTIME_NSEC = 999999999

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M
    z = x
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M
    return z ^ (z >> np.uint64(31))


def _u(seed, rnd, job, field):
    """Uniform [0,1) with 24 bits, vectorised over rnd/job arrays."""
    with np.errstate(over="ignore"):
        ctr = ((rnd.astype(np.uint64) << np.uint64(20)) | job.astype(np.uint64)) * np.uint64(16) \
            + np.uint64(field)
        x = splitmix64(np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + ctr)
    return (x >> np.uint64(40)).astype(np.float64) / float(1 << 24)


def jobs_batch(algo, rounds, jobs_per_round, total_gpu, seed=SEED, *, force_min1=None,
               ragged=False, round_base=0):
    """Appendix D.1 jobs for `rounds` rounds.

    min in {1:60%, 2:25%, 4:15%}; max = min(32, min*m), m in {2,4,8} (so max > min);
    NumProc uniform in [min, max]; Priority {0:70%, 1:30%}; SubmitTime = base + U{0..N-1} s
    (unsorted, with ties); FirstStartTime 40% the never-started sentinel, else uniform in an
    hour; remaining = f32(exp(7.5 + 1.5 z)); S[g] = f32(g / (1 + a (g-1))), a ~ U(0, 0.35).
    FfDL and AFS-L ignore MinNumProc, so their rounds use min = 1 with the
    max drawn as above.
    """
    if isinstance(algo, int):
        from ._abi import ALGO_NAMES
        algo = ALGO_NAMES[algo]
    if force_min1 is None:
        force_min1 = algo in ("FfDLOptimizer", "AFS-L")
    if ragged:
        rr = np.arange(rounds, dtype=np.int64) + round_base
        cnt = 1 + np.floor(_u(seed, rr, np.zeros(rounds, np.int64), 15) * jobs_per_round).astype(np.int64)
    else:
        cnt = np.full(rounds, jobs_per_round, np.int64)
    job_off = np.zeros(rounds + 1, np.int64)
    np.cumsum(cnt, out=job_off[1:])
    n = int(job_off[-1])
    rnd = np.repeat(np.arange(rounds, dtype=np.int64), cnt) + round_base
    job = np.arange(n, dtype=np.int64) - np.repeat(job_off[:-1], cnt)

    u = _u(seed, rnd, job, 0)
    mn = np.where(u < 0.60, 1, np.where(u < 0.85, 2, 4)).astype(np.int64)
    m = np.array([2, 4, 8])[np.minimum(2, np.floor(_u(seed, rnd, job, 1) * 3).astype(np.int64))]
    mx = np.minimum(32, mn * m)
    npc = mn + np.floor(_u(seed, rnd, job, 2) * (mx - mn + 1)).astype(np.int64)
    prio = (_u(seed, rnd, job, 3) >= 0.70).astype(np.int64)
    base = 1_790_000_000
    submit = base + np.floor(_u(seed, rnd, job, 4) * np.repeat(cnt, cnt)).astype(np.int64)
    never = _u(seed, rnd, job, 5) < 0.40
    first_sec = np.where(never, MAXTIME_SEC, base + np.floor(_u(seed, rnd, job, 6) * 3600).astype(np.int64))
    first_nsec = np.where(never, MAXTIME_NSEC, np.floor(_u(seed, rnd, job, 7) * 1e9).astype(np.int64))
    u1 = np.maximum(_u(seed, rnd, job, 8), 2.0 ** -24)
    z = np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * _u(seed, rnd, job, 9))
    rem = np.exp(7.5 + 1.5 * z).astype(np.float32)
    alpha = 0.35 * _u(seed, rnd, job, 10)
    g = np.arange(W, dtype=np.float64)[None, :]
    S = (g / (1.0 + alpha[:, None] * (g - 1.0))).astype(np.float32)
    S[:, 0] = 0.0
    if force_min1:
        mn = np.ones_like(mn)
        npc = np.maximum(1, npc)
    tot = np.full(rounds, total_gpu, np.int64) if np.isscalar(total_gpu) else np.asarray(total_gpu)
    return AllocBatch(
        algo, job_off.astype(np.int32), tot.astype(np.int32),
        submit_sec=submit, submit_nsec=np.zeros(n, np.int32),
        first_sec=first_sec, first_nsec=first_nsec.astype(np.int32),
        priority=prio.astype(np.int32), num_proc=npc.astype(np.int32),
        min_proc=mn.astype(np.int32), max_proc=mx.astype(np.int32),
        remaining_sec=rem, speedup=S.reshape(-1), speedup_stride=W)


def quirk_batch(algo, rounds, jobs_per_round, total_gpu, seed=SEED + 1):
    """D.1 plus the reference's sharp edges: min == max, min > max, Priority 2, zero and
    negative totals, negative minimum, tie-heavy keys, flat curves (zero / negative gains),
    ragged job counts."""
    b = jobs_batch(algo, rounds, jobs_per_round, total_gpu, seed, ragged=True, force_min1=False)
    n = b.jobs
    rnd = np.repeat(np.arange(rounds, dtype=np.int64), np.diff(b.job_off))
    job = np.arange(n, dtype=np.int64) - np.repeat(b.job_off[:-1].astype(np.int64), np.diff(b.job_off))
    u = _u(seed, rnd, job, 11)
    b.max_proc = np.where(u < 0.10, b.min_proc, b.max_proc).astype(np.int32)          # min == max
    b.max_proc = np.where((u >= 0.10) & (u < 0.13), np.maximum(1, b.min_proc - 1), b.max_proc).astype(np.int32)  # min > max
    b.priority = np.where(_u(seed, rnd, job, 12) < 0.05, 2, b.priority).astype(np.int32)
    flat = _u(seed, rnd, job, 13) < 0.15
    S = b.speedup.reshape(n, W).copy()
    cap = 1 + np.floor(_u(seed, rnd, job, 14) * 6).astype(np.int64)
    for_idx = np.nonzero(flat)[0]
    for i in for_idx:  # plateau (zero gain) then decline (negative gain)
        c = int(cap[i])
        S[i, c:] = S[i, c]
        S[i, min(W - 1, c + 3):] = S[i, c] * np.float32(0.75)
    b.speedup = np.ascontiguousarray(S.reshape(-1))
    tie = _u(seed, rnd, job, 15) < 0.5
    b.remaining_sec = np.where(tie, np.float32(1000.0), b.remaining_sec).astype(np.float32)
    b.submit_sec = np.where(tie, b.submit_sec[0], b.submit_sec).astype(np.int64)
    ru = _u(seed, np.arange(rounds, dtype=np.int64), np.zeros(rounds, np.int64), 14)
    tg = np.asarray(b.total_gpu).copy()
    tg[ru < 0.05] = 0
    tg[(ru >= 0.05) & (ru < 0.15)] = 1 + (tg[(ru >= 0.05) & (ru < 0.15)] // 7)
    tg[(ru >= 0.15) & (ru < 0.18)] = -1                      # negative total -> "exceeded total GPUs"
    b.total_gpu = tg.astype(np.int32)
    neg = _u(seed, rnd, job, 9) < 0.004                      # negative minimum -> "can't be negative"
    b.min_proc = np.where(neg, -1, b.min_proc).astype(np.int32)
    return b


def min_zero_batch(algo, rounds, jobs_per_round, total_gpu, seed=SEED + 2, frac=0.08):
    """D.1 rounds in which a fraction of the jobs has MinNumProc == 0 — the input for which the
    reference's ElasticTiresias greedy loop never ends (elastic_tiresias.go:136-143; reported as
    VODA_ST_NONTERMINATING).  Those jobs get S[0] = S[1]/4 so that their first gain S[0]/0 is +Inf
    rather than NaN (NaN sort keys are outside the reference's own domain)."""
    b = jobs_batch(algo, rounds, jobs_per_round, total_gpu, seed, ragged=True, force_min1=False)
    n = b.jobs
    cnt = np.diff(b.job_off)
    rnd = np.repeat(np.arange(rounds, dtype=np.int64), cnt)
    job = np.arange(n, dtype=np.int64) - np.repeat(b.job_off[:-1].astype(np.int64), cnt)
    z = _u(seed, rnd, job, 11) < frac
    b.min_proc = np.where(z, 0, b.min_proc).astype(np.int32)
    # half of them ask for more than any total: never started by the basic portion, they meet the
    # greedy loop at 0 workers
    big = z & (_u(seed, rnd, job, 12) < 0.5)
    b.num_proc = np.where(big, 1 << 20, b.num_proc).astype(np.int32)
    S = b.speedup.reshape(n, W).copy()
    S[z, 0] = S[z, 1] * np.float32(0.25)
    b.speedup = np.ascontiguousarray(S.reshape(-1))
    return b


def matrices(batch, n, kind="placement", seed=SEED, hi=8):
    """Appendix D.2 score matrices, int64 [batch, n, n].

    kind="uniform": iid integers in [0, hi].  kind="placement": overlap scores of two random
    layouts of the same jobs on n nodes x 8 slots — sparse, small, tie-heavy, like the
    matrices bindNodes builds (placement_manager.go:499-502)."""
    b = np.repeat(np.arange(batch, dtype=np.int64), n * n)
    c = np.tile(np.arange(n * n, dtype=np.int64), batch)
    u = _u(seed + 7, b, c, 0).reshape(batch, n, n)
    if kind == "uniform":
        return np.floor(u * (hi + 1)).astype(np.int64)
    # sparse overlaps: ~3 candidate columns per row share jobs with it
    dense = np.floor(_u(seed + 7, b, c, 1).reshape(batch, n, n) * (hi + 1)).astype(np.int64)
    return np.where(u < 3.0 / max(3, n), np.maximum(1, dense), 0).astype(np.int64)


def place_batch(rounds, nodes, slots, jobs_per_round, seed=SEED, load=0.85):
    """Placement rounds: `nodes` nodes x `slots` slots, requests drawn as D.1 sizes until
    `load` of the cluster is asked for; the previous layout puts each job on random nodes."""
    rng = np.random.Generator(np.random.Philox(key=seed + 11))
    node_off = np.arange(rounds + 1, dtype=np.int64) * nodes
    node_slots = np.full(rounds * nodes, slots, np.int64)
    req_off = [0]
    req_workers, prev_off, prev_job, prev_workers = [], [0], [], []
    for _ in range(rounds):
        sizes = rng.choice([1, 2, 3, 4, 6, 8, 12, 16], size=jobs_per_round,
                           p=[.3, .2, .1, .15, .08, .1, .04, .03])
        budget = int(load * nodes * slots)
        keep = np.cumsum(sizes) <= budget
        sizes = sizes[keep]
        m = len(sizes)
        req_workers.extend(int(s) for s in sizes)
        req_off.append(req_off[-1] + m)
        # previous layout: a different random packing of (mostly) the same jobs
        free = np.full(nodes, slots)
        per_node = [[] for _ in range(nodes)]
        for q in rng.permutation(m):
            if rng.random() < 0.15:
                continue  # job is new this round
            need = max(1, int(sizes[q]) + int(rng.integers(-1, 2)))
            for c in rng.permutation(nodes):
                if need == 0:
                    break
                take = min(need, int(free[c]))
                if take > 0:
                    per_node[c].append((int(q), take))
                    free[c] -= take
                    need -= take
        for c in range(nodes):
            for q, w in per_node[c]:
                prev_job.append(q)
                prev_workers.append(w)
            prev_off.append(len(prev_job))
    return PlaceBatch(node_off, node_slots, np.array(req_off), np.array(req_workers or [0])[:req_off[-1]],
                      np.array(prev_off), np.array(prev_job, dtype=np.int64), np.array(prev_workers, dtype=np.int64))


# ---- generators that run on the device (plumbing: torch integer ops), so that the large
#      throughput batches never cross PCIe.  Same counter-based PRNG, same values as the numpy
#      generators above.
def _i64(c):
    return c - (1 << 64) if c >= (1 << 63) else c


def _lsr(x, k):
    return (x >> k) & ((1 << (64 - k)) - 1)


def _splitmix64_t(x):
    x = x + _i64(0x9E3779B97F4A7C15)
    z = x
    z = (z ^ _lsr(z, 30)) * _i64(0xBF58476D1CE4E5B9)
    z = (z ^ _lsr(z, 27)) * _i64(0x94D049BB133111EB)
    return z ^ _lsr(z, 31)


def _u_t(seed, rnd, job, field):
    """torch version of _u: int64 tensors in, float64 uniform [0,1) with 24 bits out."""
    import torch
    ctr = ((rnd << 20) | job) * 16 + field
    base = _i64((seed * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF)
    x = _splitmix64_t(ctr + base)
    return _lsr(x, 40).to(torch.float64) / float(1 << 24)


def matrices_device(batch, n, device, kind="placement", seed=SEED, hi=8, first=0, chunk=8192):
    """matrices() generated on `device` (int64 [batch, n, n]); matrix b equals matrices()[first + b]."""
    import torch
    out = torch.empty((batch, n, n), dtype=torch.int64, device=device)
    cells = torch.arange(n * n, dtype=torch.int64, device=device)[None, :]
    for b0 in range(0, batch, chunk):
        b1 = min(batch, b0 + chunk)
        b = torch.arange(first + b0, first + b1, dtype=torch.int64, device=device)[:, None]
        u = _u_t(seed + 7, b, cells, 0)
        if kind == "uniform":
            m = torch.floor(u * (hi + 1)).to(torch.int64)
        else:
            dense = torch.floor(_u_t(seed + 7, b, cells, 1) * (hi + 1)).to(torch.int64)
            m = torch.where(u < 3.0 / max(3, n), torch.clamp(dense, min=1), torch.zeros_like(dense))
        out[b0:b1] = m.view(b1 - b0, n, n)
    return out


def chain_batch(rounds, nodes, slots, jobs_per_round, seed=SEED, round_base=0, load=0.85):
    """One scheduling round the way resched() runs it (scheduler.go:326-364): an Elastic-SRJF
    allocation over `jobs_per_round` ready jobs with total = load * nodes * slots GPUs (a cluster
    asked for every last slot sends bestFit into the loop the reference never leaves, Q12), whose result is the
    request list of a placement round on `nodes` nodes x `slots` slots.  Returns (AllocBatch,
    PlaceBatch); the PlaceBatch's req_workers is a placeholder to be replaced by the allocation
    output (same offsets: req_off == job_off).  The previous layout packs a random earlier
    allocation of the same jobs (1..8 workers, ~15 % of the jobs new) into the nodes in a random
    job order, slot after slot, so jobs straddle node boundaries like a real history does."""
    ab = jobs_batch("ElasticSRJF", rounds, jobs_per_round, int(load * nodes * slots), seed, round_base=round_base)
    J = jobs_per_round
    rnd = np.repeat(np.arange(rounds, dtype=np.int64) + round_base, J)
    job = np.tile(np.arange(J, dtype=np.int64), rounds)
    wprev = 1 + np.floor(_u(seed + 3, rnd, job, 0) * 8).astype(np.int64)
    wprev[_u(seed + 3, rnd, job, 1) < 0.15] = 0                       # not running before
    order = np.argsort(_u(seed + 3, rnd, job, 2).reshape(rounds, J), axis=1, kind="stable")
    w = np.take_along_axis(wprev.reshape(rounds, J), order, axis=1)
    end = np.cumsum(w, axis=1)
    start = end - w
    cap = nodes * slots
    w = np.where(end <= cap, w, 0)                                     # jobs that did not fit were not running
    end = start + w
    # a job occupying slots [start, end) touches nodes start//slots .. (end-1)//slots (at most
    # ceil(8/slots)+1 of them); emit one (node, job, workers) entry per touched node
    span = int(np.ceil(8 / slots)) + 1
    n0 = start // slots
    ent_node, ent_job, ent_w, ent_round = [], [], [], []
    rr = np.repeat(np.arange(rounds, dtype=np.int64)[:, None], J, axis=1)
    for t in range(span):
        nd = n0 + t
        lo = np.maximum(start, nd * slots)
        hi_ = np.minimum(end, (nd + 1) * slots)
        ww = hi_ - lo
        m = (ww > 0) & (w > 0)
        ent_node.append(nd[m]); ent_job.append(order[m]); ent_w.append(ww[m]); ent_round.append(rr[m])
    ent_node = np.concatenate(ent_node); ent_job = np.concatenate(ent_job)
    ent_w = np.concatenate(ent_w); ent_round = np.concatenate(ent_round)
    gnode = ent_round * nodes + ent_node
    o = np.argsort(gnode, kind="stable")
    gnode, ent_job, ent_w = gnode[o], ent_job[o], ent_w[o]
    prev_off = np.zeros(rounds * nodes + 1, np.int64)
    np.cumsum(np.bincount(gnode, minlength=rounds * nodes), out=prev_off[1:])
    pb = PlaceBatch(np.arange(rounds + 1, dtype=np.int64) * nodes, np.full(rounds * nodes, slots, np.int64),
                    ab.job_off.astype(np.int64), np.zeros(max(1, ab.jobs), np.int64)[:ab.jobs], prev_off, ent_job, ent_w,
                    max_nodes=nodes, max_requests=J)
    return ab, pb


def share_curves(b: AllocBatch, want_dense: bool = True):
    """The what-if form of a batch: every round is the SAME set of jobs (the jobs of round 0: their
    speedup curves, sizes and times) under the round's own total and remaining times, with the curves
    stored ONCE as a pool that all rounds point into (voda_alloc_batch.speedup_off).  Needs rounds of
    equal size.  Returns (pooled batch, the equivalent dense batch)."""
    cnt = np.diff(b.job_off)
    assert (cnt == cnt[0]).all(), "share_curves needs rounds of equal size"
    J, R, Wd = int(cnt[0]), b.rounds, b.speedup_stride
    tile = lambda v: None if v is None else np.ascontiguousarray(np.tile(v[:J], R))
    pool = np.ascontiguousarray(b.speedup[:J * Wd])
    common = dict(submit_sec=tile(b.submit_sec), submit_nsec=tile(b.submit_nsec), first_sec=tile(b.first_sec),
                  first_nsec=tile(b.first_nsec), priority=tile(b.priority), num_proc=tile(b.num_proc),
                  min_proc=tile(b.min_proc), max_proc=tile(b.max_proc), remaining_sec=b.remaining_sec,
                  speedup_stride=Wd, queue_num=b.queue_num, compaction_threshold=b.compaction_threshold)
    pooled = AllocBatch(b.algo, b.job_off, b.total_gpu, speedup=pool,
                        speedup_off=np.tile(np.arange(J, dtype=np.int32) * Wd, R), **common)
    if not want_dense:
        return pooled, None
    dense = AllocBatch(b.algo, b.job_off, b.total_gpu, speedup=np.ascontiguousarray(np.tile(pool, R)), **common)
    return pooled, dense
