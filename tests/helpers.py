"""Shared helpers for the tests: KAT loading and batch construction."""
import json
import os

import numpy as np

from vodascheduler_b200.batch import AllocBatch, PlaceBatch
from vodascheduler_b200.synth import W

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_kats():
    with open(os.path.join(GOLDEN, "kats.json")) as f:
        return json.load(f)


def batch_from_jobs(algo, rounds, width=W):
    """rounds: list of (jobs:list[dict], total:int) -> AllocBatch (input order preserved)."""
    off, tot = [0], []
    cols = {k: [] for k in ("submit", "first", "priority", "num_proc", "min", "max", "remaining")}
    S = []
    for jobs, total in rounds:
        tot.append(total)
        off.append(off[-1] + len(jobs))
        for j in jobs:
            for k in cols:
                cols[k].append(j[k])
            row = np.zeros(width, np.float32)
            c = np.asarray(j["speedup"], np.float32)[:width]
            row[:len(c)] = c
            S.append(row)
    n = off[-1]
    S = np.concatenate(S) if S else np.zeros(0, np.float32)
    return AllocBatch(
        algo, np.array(off, np.int32), np.array(tot, np.int32),
        submit_sec=np.array(cols["submit"], np.int64), submit_nsec=np.zeros(n, np.int32),
        first_sec=np.array(cols["first"], np.int64), first_nsec=np.zeros(n, np.int32),
        priority=np.array(cols["priority"], np.int32), num_proc=np.array(cols["num_proc"], np.int32),
        min_proc=np.array(cols["min"], np.int32), max_proc=np.array(cols["max"], np.int32),
        remaining_sec=np.array(cols["remaining"], np.float32), speedup=S, speedup_stride=width)


def place_batch_from(rounds):
    """rounds: list of dict(nodes=[slots...], requests=[n...], prev=[[(job,w),...] per node])"""
    node_off, req_off, prev_off = [0], [0], [0]
    slots, req, pj, pw = [], [], [], []
    for r in rounds:
        n = len(r["nodes"])
        slots += list(r["nodes"])
        req += list(r["requests"])
        node_off.append(node_off[-1] + n)
        req_off.append(req_off[-1] + len(r["requests"]))
        prev = r.get("prev") or [[] for _ in range(n)]
        for c in range(n):
            for (j, w) in prev[c]:
                pj.append(j)
                pw.append(w)
            prev_off.append(len(pj))
    return PlaceBatch(np.array(node_off), np.array(slots, np.int64), np.array(req_off),
                      np.array(req, np.int64), np.array(prev_off), np.array(pj, np.int64),
                      np.array(pw, np.int64))
