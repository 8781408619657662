"""Trace replay (BASELINE configs[4]): a synthetic Philly-like trace driven through
Elastic-SRJF allocation + placement, for many what-if REPLICAS at once.

One replica is sequential in simulated time (each round depends on the previous one), so the
parallelism is across replicas ("replicas only"): at every rate-limit tick the
rounds of all replicas that need one go to the GPU as ONE allocation batch and ONE placement
batch.  Replicas differ in seed, cluster size and load.

Model:
  * jobs arrive as a Poisson process tuned to `load` x cluster capacity; a job carries D.1
    min/max sizes, a speedup curve S and a demand of serial work (seconds at one GPU);
    holding w GPUs it progresses at S[w] work-seconds per second;
  * the scheduler reschedules at most once per `rate_limit` seconds (30 s in the reference,
    pkg/scheduler/scheduler/scheduler.go:212) and only when something changed (arrival or
    completion; scheduler.go:648,887);
  * a round = ElasticSRJF.Schedule over the live jobs (EstimatedRemainningTimeSec = remaining
    work) followed by PlacementManager.Place of the resulting allocation.
Host side: plain Python bookkeeping (the reference's is map manipulation too); all policy and
placement arithmetic runs on the GPU through the C ABI.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List

import numpy as np

from . import synth
from .algorithm import SchedulePanic
from .batch import AllocBatch, PlaceBatch, PlaceOut
from .placement import PlacementManager

W = synth.W


@dataclasses.dataclass
class Trace:
    arrival: np.ndarray   # [J] seconds, ascending
    work: np.ndarray      # [J] serial work, seconds at one GPU
    min_proc: np.ndarray
    max_proc: np.ndarray
    speedup: np.ndarray   # [J, W] float32


def make_trace(jobs: int, gpus: int, load: float = 0.85, seed: int = synth.SEED, max_size: int = 32) -> Trace:
    rng = np.random.Generator(np.random.Philox(key=seed))
    u = rng.random(jobs)
    mn = np.where(u < 0.60, 1, np.where(u < 0.85, 2, 4))
    # max > min; capped at one node's slots: a job wider than a node always
    # triggers the reference's split bug (Q11) and soon after the endless loop (Q12)
    mx = np.minimum(max(max_size, 5), mn * rng.choice([2, 4, 8], size=jobs))
    mx = np.maximum(mx, mn + 1)
    alpha = 0.35 * rng.random(jobs)
    g = np.arange(W, dtype=np.float64)[None, :]
    S = (g / (1.0 + alpha[:, None] * (g - 1.0))).astype(np.float32)
    S[:, 0] = 0.0
    work = np.exp(7.5 + 1.5 * rng.standard_normal(jobs))            # log-normal service demand
    # GPU-seconds a job costs when it runs at its minimum size ~ work / S[min] * min
    cost = work / S[np.arange(jobs), mn] * mn
    rate = load * gpus / cost.mean()                                 # arrivals per second
    arrival = np.cumsum(rng.exponential(1.0 / rate, size=jobs))
    return Trace(arrival, work.astype(np.float64), mn.astype(np.int32), mx.astype(np.int32), S)


class Replica:
    """One what-if variant: its own trace, cluster and state."""

    def __init__(self, idx: int, trace: Trace, nodes: int, slots: int, rate_limit: float = 30.0):
        self.idx, self.trace = idx, trace
        self.total_gpu = nodes * slots
        self.rate_limit = rate_limit
        self.pm = PlacementManager({f"gpu{i}": slots for i in range(nodes)}, ctx=None)
        self.now = 0.0
        self.next_arrival = 0
        self.live: List[int] = []                 # job ids, in arrival order
        self.remaining: Dict[int, float] = {}
        self.alloc: Dict[int, int] = {}
        self.dirty = False
        self.done = 0
        self.rounds = 0
        self.migrated_workers = 0
        self.placements_skipped = 0   # rounds where the reference's bestFit would never return
        self.jct_sum = 0.0

    def finished(self) -> bool:
        return self.done == len(self.trace.arrival)

    def advance(self) -> None:
        """Move simulated time one rate-limit period forward: progress, completions, arrivals."""
        t1 = self.now + self.rate_limit
        tr = self.trace
        if not self.live and self.next_arrival < len(tr.arrival) and tr.arrival[self.next_arrival] > t1:
            t1 = float(tr.arrival[self.next_arrival])  # idle cluster: jump to the next arrival
        dt = t1 - self.now
        still = []
        for j in self.live:
            w = self.alloc.get(j, 0)
            if w > 0:
                self.remaining[j] -= dt * float(tr.speedup[j, min(w, W - 1)])
            if self.remaining[j] <= 0.0:
                self.done += 1
                self.jct_sum += t1 - float(tr.arrival[j])
                self.alloc.pop(j, None)
                del self.remaining[j]
                self.dirty = True   # job finished (scheduler.go:648)
            else:
                still.append(j)
        self.live = still
        while self.next_arrival < len(tr.arrival) and tr.arrival[self.next_arrival] <= t1:
            j = self.next_arrival
            self.live.append(j)
            self.remaining[j] = float(tr.work[j])
            self.next_arrival += 1
            self.dirty = True       # job created (scheduler.go:887)
        self.now = t1

    def alloc_round(self):
        """(job ids, per-job arrays) of the ElasticSRJF round this replica wants now."""
        ids = np.array(self.live, np.int64)
        rem = np.array([self.remaining[j] for j in self.live], np.float32)
        return ids, rem


def replay(replicas: List[Replica], ctx, max_ticks: int = 10 ** 9, progress=None):
    """Run all replicas to completion (or max_ticks); returns per-replica stats and totals."""
    for r in replicas:
        r.pm._ctx = ctx
    ticks = alloc_rounds = 0
    while ticks < max_ticks and not all(r.finished() for r in replicas):
        ticks += 1
        todo = []
        for r in replicas:
            if r.finished():
                continue
            r.advance()
            if r.dirty and r.live:
                todo.append(r)
            r.dirty = False
        if not todo:
            continue
        # ---- one allocation batch: a round per replica -------------------------------------
        off, tot, parts = [0], [], []
        for r in todo:
            ids, rem = r.alloc_round()
            parts.append((ids, rem))
            off.append(off[-1] + len(ids))
            tot.append(r.total_gpu)
        ids_all = np.concatenate([p[0] for p in parts])
        tr_of = np.concatenate([np.full(len(p[0]), k) for k, p in enumerate(parts)])
        mn = np.concatenate([todo[k].trace.min_proc[p[0]] for k, p in enumerate(parts)])
        mx = np.concatenate([todo[k].trace.max_proc[p[0]] for k, p in enumerate(parts)])
        b = AllocBatch("ElasticSRJF", np.array(off, np.int32), np.array(tot, np.int32),
                       min_proc=mn, max_proc=mx, remaining_sec=np.concatenate([p[1] for p in parts]))
        gpus, status = ctx.allocate(b)
        alloc_rounds += len(todo)
        # ---- one placement batch ------------------------------------------------------------
        reqs, pbs = [], []
        for k, r in enumerate(todo):
            if status[k] != 0:
                from ._lib import status_str
                raise SchedulePanic(int(status[k]), status_str(status[k]))
            sl = slice(off[k], off[k + 1])
            r.alloc = {int(j): int(g) for j, g in zip(parts[k][0], gpus[sl])}
            req = {f"job{j}": g for j, g in r.alloc.items() if g > 0}
            reqs.append(req)
            pbs.append(r.pm.begin_place(req))
        pb = _concat_place(pbs)
        out = ctx.place(pb)
        for k, r in enumerate(todo):
            r.rounds += 1
            if out.status[k] == 7:
                # VODA_ST_NONTERMINATING: the reference's bestFit loops for ever on this round
                # (a split job over-books a node, then a later request meets a
                # cluster without usable slots).  A what-if replay cannot hang, so the round's
                # placement is skipped: bindings stay as they were, the allocation stands.
                r.placements_skipped += 1
                r.pm.updateJobStates()  # keep jobStates consistent with the released nodeStates
                continue
            r.pm.finish_place(reqs[k], out, k, int(pb.node_off[k]), int(pb.node_off[k]) + int(pb.req_off[k]))
            r.migrated_workers += r.pm.deletedWorkers
        if progress and ticks % 200 == 0:
            progress(ticks, replicas)
    return dict(ticks=ticks, rounds=alloc_rounds,
                replicas=[dict(idx=r.idx, done=r.done, rounds=r.rounds, sim_time=r.now,
                               mean_jct=r.jct_sum / max(1, r.done), migrated_workers=r.migrated_workers,
                               placements_skipped=r.placements_skipped)
                          for r in replicas])


def _concat_place(pbs: List[PlaceBatch]) -> PlaceBatch:
    node_off, req_off, prev_off = [0], [0], [0]
    for p in pbs:
        node_off.append(node_off[-1] + p.nodes)
        req_off.append(req_off[-1] + p.requests)
        prev_off.extend((p.prev_off[1:] + prev_off[-1]).tolist())
    cat = lambda name: np.concatenate([getattr(p, name) for p in pbs]) if pbs else np.zeros(0, np.int32)
    return PlaceBatch(node_off, cat("node_slots"), req_off, cat("req_workers"), prev_off,
                      cat("prev_job"), cat("prev_workers"))


def make_replicas(count: int, jobs: int, nodes: int = 128, slots: int = 8, seed: int = synth.SEED,
                  first: int = 0) -> List[Replica]:
    """What-if variants: seed, load (0.70-0.95) and cluster size (100% .. 75% of `nodes`) vary."""
    out = []
    for k in range(first, first + count):
        load = 0.70 + 0.25 * ((k * 7) % 11) / 10.0
        n_k = max(1, nodes - (k % 5) * max(1, nodes // 16))
        out.append(Replica(k, make_trace(jobs, n_k * slots, load, seed + 1000 * k, max_size=slots), n_k, slots))
    return out
