"""CPU replay of a trace — TEST INFRASTRUCTURE, not product.

The same event model as voda_replay (include/voda_accel.h, "trace replay"), one replica at a time, with
every decision taken by the CPU restatements: oracle.allocate (ElasticSRJF), oracle.place (bestFit,
score, munkres) and oracle/bookkeeping.py (releaseSlots, updateJobStates, updatePodNodeName).  It
produces the per-replica statistics and the history digest voda_replay produces, so a test can demand
that the GPU-driven replay took exactly the same decisions round after round.  Slow (Python loop):
meant for traces of a few hundred jobs.
"""
from __future__ import annotations

import numpy as np

from vodascheduler_b200.batch import AllocBatch, PlaceBatch

from . import bookkeeping, oracle

MASK = (1 << 64) - 1


def _fnv(h, v):
    return ((h ^ (int(v) & MASK)) * 0x100000001B3) & MASK


def _actions(old, new):
    """Scheduler.compareResults (scheduler.go:448-480) for jobs that are neither completed nor failed:
    0 halts, 1 scale-ins, 2 scale-outs, 3 starts."""
    if old > new:
        return 0 if new == 0 else 1
    if old < new:
        return 3 if old == 0 else 2
    return None


def replay_one(trace, nodes, slots, rate_limit=30.0, max_ticks=0, dense=True):
    """trace: arrays arrival, work (float64), min_proc, max_proc (int32), speedup [J, W] float32."""
    J = len(trace.arrival)
    W = trace.speedup.shape[1]
    pm = bookkeeping.Manager(nodes, slots)
    now, nxt = 0.0, 0
    live, remaining, alloc = [], {}, {}
    done = rounds = mig_w = mig_l = skipped = ticks = 0
    act = [0, 0, 0, 0]
    jct = 0.0
    hist = 0xCBF29CE484222325
    while done < J and (max_ticks <= 0 or ticks < max_ticks):
        ticks += 1
        # ---- advance one rate-limit period --------------------------------------------------
        t1 = now + rate_limit
        if not live and nxt < J and trace.arrival[nxt] > t1:
            t1 = float(trace.arrival[nxt])
        dt = t1 - now
        dirty = False
        still = []
        for j in live:
            w = alloc.get(j, 0)
            if w > 0:
                remaining[j] -= dt * float(trace.speedup[j, min(w, W - 1)])
            if remaining[j] <= 0.0:
                done += 1
                jct += t1 - float(trace.arrival[j])
                alloc[j] = 0
                dirty = True
            else:
                still.append(j)
        live = still
        while nxt < J and trace.arrival[nxt] <= t1:
            live.append(nxt)
            remaining[nxt] = float(trace.work[nxt])
            nxt += 1
            dirty = True
        now = t1
        if not (dirty and live):
            continue
        # ---- the round: ElasticSRJF over the live jobs ------------------------------------------
        ids = np.array(live, np.int64)
        b = AllocBatch("ElasticSRJF", np.array([0, len(live)], np.int32), np.array([nodes * slots], np.int32),
                       min_proc=trace.min_proc[ids], max_proc=trace.max_proc[ids],
                       remaining_sec=np.array([remaining[j] for j in live], np.float32))
        g, st = oracle.allocate(b, dense=dense)
        if st[0] != 0:
            raise RuntimeError(f"allocation status {st[0]} at tick {ticks}")
        for j, new in zip(live, g):
            a = _actions(alloc.get(j, 0), int(new))
            if a is not None:
                act[a] += 1
            alloc[j] = int(new)
        # ---- Place: releaseSlots, bestFit + score + munkres, rename, job states, pod table ----------
        requests = {j: alloc[j] for j in live}
        pm.release_slots(requests)
        index = {j: q for q, j in enumerate(live)}
        prev = pm.previous_layout(index)
        prev_off, pj, pw = [0], [], []
        for ent in prev:
            for q, w in ent:
                pj.append(q)
                pw.append(w)
            prev_off.append(len(pj))
        pb = PlaceBatch([0, nodes], [slots] * nodes, [0, len(live)], [alloc[j] for j in live], prev_off,
                        np.array(pj, np.int64), np.array(pw, np.int64))
        out = oracle.place(pb)
        rounds += 1
        hist = _fnv(hist, ticks)
        hist = _fnv(hist, len(live))
        for j in live:
            hist = _fnv(hist, alloc[j])
        if out.status[0] == 7:
            skipped += 1
            pm.update_job_states()
            hist = _fnv(hist, -1)
            continue
        if out.status[0] != 0:
            raise RuntimeError(f"placement status {out.status[0]} at tick {ticks}")
        e0 = 0
        layout = [(int(out.lay_node[e0 + t]), live[int(out.lay_job[e0 + t])], int(out.lay_workers[e0 + t]))
                  for t in range(int(out.lay_cnt[0]))]
        pm.bind(layout, [int(x) for x in out.free_slots[:nodes]], [int(x) for x in out.phys_of_logical[:nodes]])
        pm.update_job_states()
        dw, dl = pm.update_pod_node_name()
        mig_w += dw
        mig_l += dl
        for c in range(nodes):
            hist = _fnv(hist, int(out.phys_of_logical[c]))
        hist = _fnv(hist, int(out.score[0]))
        hist = _fnv(hist, int(out.cross[0]))
        hist = _fnv(hist, dw)
        hist = _fnv(hist, dl)
    return dict(done=done, rounds=rounds, sim_time=now, jct_sum=jct, migrated_workers=mig_w, migrated_launchers=mig_l,
                placements_skipped=skipped, actions=dict(halts=act[0], scale_ins=act[1], scale_outs=act[2], starts=act[3]),
                history=hist, ticks=ticks)
