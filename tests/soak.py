"""Long randomised parity run: for every policy, at least 10^5
randomised rounds GPU vs the dense oracle and at least 10^4 rounds dense vs structure-faithful
oracle, over the shape / quirk / table-width / parameter space of tests/test_fuzz_gpu.py.
Not collected by pytest (no test_ prefix): run it on a GPU box,

    python tests/soak.py --out profiles/r2_soak.json [--gpu-rounds 100000] [--cpu-rounds 10000]

and commit the record it writes.  Exit status 1 on the first mismatch (the failing case is
printed with everything needed to reproduce it)."""
import argparse
import json
import os
import sys
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import oracle  # noqa: E402
from vodascheduler_b200 import ALGO_NAMES, Context, synth  # noqa: E402


def reshape_table(b, width):
    S = b.speedup.reshape(-1, b.speedup_stride)
    T = np.zeros((S.shape[0], width), np.float32)
    w = min(width, S.shape[1])
    T[:, :w] = S[:, :w]
    b.speedup = np.ascontiguousarray(T.reshape(-1))
    b.speedup_stride = width
    return b


def draw(rng, algo, small):
    jobs = int(rng.choice([1, 2, 5, 17, 33, 64, 65] if small else [1, 2, 5, 17, 33, 64, 65, 97, 130, 260, 520, 1024]))
    total = int(rng.choice([0, 1, 3, jobs, 4 * jobs, 40 * jobs + 7]))
    rounds = int(rng.integers(20, 200)) if jobs <= 130 else int(rng.integers(4, 24))
    kind = int(rng.integers(0, 3))
    seed = int(rng.integers(1, 1 << 30))
    gen = (synth.jobs_batch, synth.quirk_batch, synth.min_zero_batch)[kind]
    b = gen(algo, rounds, jobs, total, seed)
    width = int(rng.choice([2, 5, 9, 34, 40]))
    b = reshape_table(b, width)
    b.queue_num = int(rng.choice([0, 1, 2, 3]))
    b.compaction_threshold = int(rng.choice([-1, 0, 3, 10, 1000]))
    if rng.integers(0, 3) == 0:
        b.max_proc = (b.max_proc.astype(np.int64) * int(rng.choice([1, 3, 20]))).astype(np.int32)
    tag = dict(algo=algo, jobs=jobs, total=total, rounds=rounds, kind=kind, seed=seed, width=width,
               queue_num=b.queue_num, compaction=b.compaction_threshold)
    return b, tag


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpu-rounds", type=int, default=100000)
    ap.add_argument("--cpu-rounds", type=int, default=10000)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_soak.json"))
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args()
    oracle.build()
    ctx = Context(0)
    threads = os.cpu_count() or 1
    record = {"seed": args.seed, "threads": threads, "policies": {}}
    t_all = time.time()
    for algo in ALGO_NAMES:
        rng = np.random.default_rng(zlib.crc32(algo.encode()) + args.seed)
        done = cases = nonzero = 0
        t0 = time.time()
        while done < args.gpu_rounds:
            b, tag = draw(rng, algo, small=False)
            g_ref, s_ref = oracle.allocate(b, dense=True, threads=threads)
            g, s = ctx.allocate(b)
            if not (np.array_equal(g, g_ref) and np.array_equal(s, s_ref)):
                print("MISMATCH gpu vs dense:", json.dumps(tag))
                sys.exit(1)
            done += b.rounds
            cases += 1
            nonzero += int((s != 0).sum())
        cdone = ccases = 0
        while cdone < args.cpu_rounds:
            b, tag = draw(rng, algo, small=True)
            g0, s0 = oracle.allocate(b, dense=False, threads=threads)
            g1, s1 = oracle.allocate(b, dense=True, threads=threads)
            if not (np.array_equal(g0, g1) and np.array_equal(s0, s1)):
                print("MISMATCH dense vs faithful:", json.dumps(tag))
                sys.exit(1)
            cdone += b.rounds
            ccases += 1
        record["policies"][algo] = {"gpu_vs_dense_rounds": done, "gpu_vs_dense_cases": cases,
                                    "rounds_with_nonzero_status": nonzero, "dense_vs_faithful_rounds": cdone,
                                    "dense_vs_faithful_cases": ccases, "seconds": round(time.time() - t0, 1), "mismatches": 0}
        print(algo, record["policies"][algo], flush=True)
    record["seconds"] = round(time.time() - t_all, 1)
    try:
        import subprocess
        record["git"] = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
    except OSError:
        pass
    with open(args.out, "w") as f:
        json.dump(record, f, indent=1)
    ctx.close()
    print("soak ok ->", args.out)


if __name__ == "__main__":
    main()
