"""Trace replay (BASELINE configs[4]): a synthetic Philly-like trace driven through
Elastic-SRJF allocation + placement, for many what-if REPLICAS at once.

One replica is sequential in simulated time (each round depends on the previous one), so the
parallelism is across replicas ("replicas only"): at every rate-limit tick the
rounds of all replicas that need one go to the GPU as ONE allocation batch and ONE placement
batch.  Replicas differ in seed, cluster size and load.

The event loop and the placement bookkeeping live in the library (voda_replay, csrc/replay.cu:
C++ host code around the kernels); this module builds the synthetic traces (SURVEY.md Appendix
D.3) and calls it.  Model:
  * jobs arrive as a Poisson process tuned to `load` x cluster capacity; a job carries D.1
    min/max sizes, a speedup curve S and a demand of serial work (seconds at one GPU);
    holding w GPUs it progresses at S[w] work-seconds per second;
  * the scheduler reschedules at most once per `rate_limit` seconds (30 s in the reference,
    pkg/scheduler/scheduler/scheduler.go:212) and only when something changed (arrival or
    completion; scheduler.go:648,887);
  * a round = ElasticSRJF.Schedule over the live jobs (EstimatedRemainningTimeSec = remaining
    work) followed by PlacementManager.Place of the resulting allocation.
"""
from __future__ import annotations

import ctypes as C
import dataclasses
from typing import List

import numpy as np

from . import _abi, synth

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


@dataclasses.dataclass
class Replica:
    """One what-if variant: its own trace, cluster and rate limit."""
    idx: int
    trace: Trace
    nodes: int
    slots: int
    rate_limit: float = 30.0


def make_replicas(count: int, jobs: int, nodes: int = 128, slots: int = 8, seed: int = synth.SEED,
                  first: int = 0) -> List[Replica]:
    """What-if variants: seed, load (0.70-0.95) and cluster size (100% .. 75% of `nodes`) vary."""
    out = []
    for k in range(first, first + count):
        load = 0.70 + 0.25 * ((k * 7) % 11) / 10.0
        n_k = max(1, nodes - (k % 5) * max(1, nodes // 16))
        out.append(Replica(k, make_trace(jobs, n_k * slots, load, seed + 1000 * k, max_size=slots), n_k, slots))
    return out


STAT_FIELDS = ("rounds", "done", "sim_time", "jct_sum", "migrated_workers", "migrated_launchers",
               "placements_skipped", "actions", "history")


def replay(replicas: List[Replica], ctx, max_ticks: int = 0, host_threads: int = 0):
    """Run all replicas to completion (or max_ticks) through voda_replay; returns totals and
    per-replica stats (`history` is the digest of every round's allocations and bindings)."""
    R = len(replicas)
    off = np.zeros(R + 1, np.int32)
    np.cumsum([len(r.trace.arrival) for r in replicas], out=off[1:])
    cat = lambda f, dt: np.ascontiguousarray(np.concatenate([f(r) for r in replicas]).astype(dt)) if R else np.zeros(0, dt)
    arrival = cat(lambda r: r.trace.arrival, np.float64)
    work = cat(lambda r: r.trace.work, np.float64)
    mn = cat(lambda r: r.trace.min_proc, np.int32)
    mx = cat(lambda r: r.trace.max_proc, np.int32)
    sp = cat(lambda r: r.trace.speedup.reshape(-1), np.float32)
    nodes = np.array([r.nodes for r in replicas], np.int32)
    slots = np.array([r.slots for r in replicas], np.int32)
    rate = np.array([r.rate_limit for r in replicas], np.float64)
    b = _abi.ReplayBatchStruct(R, off.ctypes.data, arrival.ctypes.data, work.ctypes.data, mn.ctypes.data,
                               mx.ctypes.data, sp.ctypes.data, W, nodes.ctypes.data, slots.ctypes.data,
                               rate.ctypes.data, int(max_ticks), int(host_threads))
    n = max(1, R)
    o = {k: np.zeros(n, np.int64) for k in ("rounds", "done", "migrated_workers", "migrated_launchers", "placements_skipped")}
    o["sim_time"], o["jct_sum"] = np.zeros(n, np.float64), np.zeros(n, np.float64)
    o["actions"] = np.zeros(4 * n, np.int64)
    o["history"] = np.zeros(n, np.uint64)
    ticks = np.zeros(1, np.int64)
    s = _abi.ReplayOutStruct(o["rounds"].ctypes.data, o["done"].ctypes.data, o["sim_time"].ctypes.data,
                             o["jct_sum"].ctypes.data, o["migrated_workers"].ctypes.data,
                             o["migrated_launchers"].ctypes.data, o["placements_skipped"].ctypes.data,
                             o["actions"].ctypes.data, o["history"].ctypes.data, ticks.ctypes.data)
    ctx._check(ctx._lib.voda_replay(ctx._h, C.byref(b), C.byref(s)), "voda_replay")
    reps = [dict(idx=r.idx, done=int(o["done"][k]), rounds=int(o["rounds"][k]), sim_time=float(o["sim_time"][k]),
                 mean_jct=float(o["jct_sum"][k]) / max(1, int(o["done"][k])), jct_sum=float(o["jct_sum"][k]),
                 migrated_workers=int(o["migrated_workers"][k]), migrated_launchers=int(o["migrated_launchers"][k]),
                 placements_skipped=int(o["placements_skipped"][k]),
                 actions=dict(zip(("halts", "scale_ins", "scale_outs", "starts"), (int(x) for x in o["actions"][4 * k:4 * k + 4]))),
                 history=int(o["history"][k]))
            for k, r in enumerate(replicas)]
    return dict(ticks=int(ticks[0]), rounds=int(o["rounds"][:R].sum()), replicas=reps)


def replay_parallel(replicas: List[Replica], contexts, max_ticks: int = 0, host_threads: int = 0):
    """The same replay with the replicas dealt round-robin to several contexts of one GPU, one host
    thread per context (ctypes releases the GIL during voda_replay): while one group waits for its
    kernels another does its bookkeeping.  Replicas are independent, so the per-replica results are
    those of replay(); they come back in the order of `replicas`."""
    import threading
    K = max(1, min(len(contexts), len(replicas)))
    out: list = [None] * K
    err: list = [None] * K

    def work(i):
        try:
            out[i] = replay(replicas[i::K], contexts[i], max_ticks, host_threads)
        except Exception as e:  # surfaced below
            err[i] = e

    threads = [threading.Thread(target=work, args=(i,)) for i in range(K)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for e in err:
        if e is not None:
            raise e
    reps = [None] * len(replicas)
    for i in range(K):
        for j, r in enumerate(out[i]["replicas"]):
            reps[i + j * K] = r
    return dict(ticks=max(o["ticks"] for o in out), rounds=sum(o["rounds"] for o in out), replicas=reps)
