"""Randomised parity run for the assignment and placement kernels (the counterpart of tests/soak.py
for the allocation policies): voda_munkres_max and voda_place against the CPU oracle over matrix sizes
1..128, value ranges from 1 to 2^40 (every cell width), batches shifted to hold negative scores (the
reference's wrapping MaxInt64 - score arithmetic), and placement rounds with ragged clusters, every
score width, negative previous counts and node capacities, zero and negative requests — in the fused
and the split form of the placement kernels.  Not collected by pytest: run it on a GPU box,

    python tests/soak_place.py --out profiles/r2_soak_place.json [--seeds 6]

and commit the record.  Exit status 1 on the first mismatch."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import oracle  # noqa: E402
from vodascheduler_b200 import Context, synth  # noqa: E402


def place_same(pb, got, ref):
    if not np.array_equal(got.status[:pb.rounds], ref.status[:pb.rounds]):
        return False
    for r in range(pb.rounds):
        if ref.status[r] != 0:
            continue
        n0, n1 = int(pb.node_off[r]), int(pb.node_off[r + 1])
        e0, c = int(pb.req_off[r]) + n0, int(ref.lay_cnt[r])
        if not (got.lay_cnt[r] == c and got.score[r] == ref.score[r] and got.cross[r] == ref.cross[r] and
                np.array_equal(got.phys_of_logical[n0:n1], ref.phys_of_logical[n0:n1]) and
                np.array_equal(got.free_slots[n0:n1], ref.free_slots[n0:n1]) and
                all(np.array_equal(getattr(got, k)[e0:e0 + c], getattr(ref, k)[e0:e0 + c])
                    for k in ("lay_node", "lay_job", "lay_workers"))):
            return False
    return True


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--seeds", type=int, default=6)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    args = ap.parse_args()
    oracle.build()
    ctx = Context(0)
    t0 = time.time()
    rec = {"seeds": args.seeds, "threads": args.threads, "munkres": {"matrices": 0, "with_negative_scores": 0},
           "placement": {"fused": 0, "split": 0, "odd_input_rounds": 0}}
    for seed in range(args.seeds):
        rng = np.random.default_rng(1000 + seed)
        for _ in range(60):
            n = int(rng.choice([1, 2, 3, 4, 5, 7, 8, 12, 16, 31, 32, 33, 48, 63, 64, 65, 96, 100, 127, 128]))
            kind = str(rng.choice(["placement", "uniform"]))
            hi = int(rng.choice([1, 2, 8, 255, 256, 999, 65535, 70000, 2 ** 31 - 1, 2 ** 40]))
            B = int(rng.integers(8, 120))
            A = synth.matrices(B, n, kind=kind, hi=hi, seed=int(rng.integers(1, 1 << 30)))
            if rng.random() < 0.3:
                A = A - int(rng.integers(0, hi + 1))
                rec["munkres"]["with_negative_scores"] += B
            got, st = ctx.munkres_max(A)
            ref, _ = oracle.munkres_max(A, threads=args.threads)
            if not (np.array_equal(got, ref) and (st == 0).all()):
                print(f"MUNKRES MISMATCH seed {seed} n {n} kind {kind} hi {hi} batch {B}")
                return 1
            rec["munkres"]["matrices"] += B
        for split in ("0", "1"):
            os.environ["VODA_PLACE_SPLIT"] = split
            for _ in range(30):
                nodes = int(rng.choice([1, 2, 3, 4, 8, 13, 16, 32, 37, 64, 100, 128]))
                pb = synth.place_batch(rounds=int(rng.integers(4, 32)), nodes=nodes, slots=int(rng.choice([1, 2, 4, 8, 16])),
                                       jobs_per_round=int(rng.integers(1, 2 * nodes + 8)), load=float(rng.choice([0.3, 0.7, 0.9, 1.0, 1.3])),
                                       seed=int(rng.integers(1, 1 << 30)))
                scale = int(rng.choice([1, 1, 1, 50, 400, 30000]))
                pb.node_slots = (pb.node_slots * scale).astype(pb.node_slots.dtype)
                pb.req_workers = (pb.req_workers * scale).astype(pb.req_workers.dtype)
                pb.prev_workers = (pb.prev_workers * scale).astype(pb.prev_workers.dtype)
                odd = rng.random()
                if odd < 0.25:    # a previous layout with negative counts: mixed-sign score rows
                    m = rng.random(pb.prev_workers.shape[0]) < 0.15
                    pb.prev_workers = np.where(m, -pb.prev_workers - int(rng.integers(0, 3)), pb.prev_workers).astype(pb.prev_workers.dtype)
                elif odd < 0.4:   # nodes without (or with negative) capacity
                    m = rng.random(pb.node_slots.shape[0]) < 0.15
                    pb.node_slots = np.where(m, -int(rng.integers(0, 3)), pb.node_slots).astype(pb.node_slots.dtype)
                elif odd < 0.5:   # zero and negative requests
                    m = rng.random(pb.req_workers.shape[0]) < 0.12
                    pb.req_workers = np.where(m, -rng.integers(0, 4, size=m.shape[0]), pb.req_workers).astype(pb.req_workers.dtype)
                if odd < 0.5:
                    rec["placement"]["odd_input_rounds"] += pb.rounds
                if not place_same(pb, ctx.place(pb), oracle.place(pb, threads=args.threads)):
                    print(f"PLACEMENT MISMATCH seed {seed} split {split} nodes {nodes} scale {scale} odd {odd:.2f}")
                    return 1
                rec["placement"]["split" if split == "1" else "fused"] += pb.rounds
    os.environ.pop("VODA_PLACE_SPLIT", None)
    rec["seconds"] = round(time.time() - t0, 1)
    rec["mismatches"] = 0
    print("soak ok:", json.dumps(rec))
    if args.out:
        with open(args.out, "w") as f:
            json.dump(rec, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
