"""Randomised shape / parameter sweep of the allocation kernels against the oracle: table
widths other than 34 (narrow tables make g run past the table: the Go map-miss path), queue
counts, compaction thresholds, totals from 0 to far above the demand, rounds from 1 to a few
hundred jobs.  VODA_FUZZ_CASES raises the number of cases for a long soak run."""
import os
import zlib

import numpy as np
import pytest

from vodascheduler_b200 import ALGO_NAMES, synth
from vodascheduler_b200.batch import AllocBatch

pytestmark = pytest.mark.gpu
CASES = int(os.environ.get("VODA_FUZZ_CASES", "6"))


def _reshape_table(b: AllocBatch, width: int) -> AllocBatch:
    """Same jobs with a speedup table of `width` entries (truncated or zero-padded)."""
    S = b.speedup.reshape(-1, b.speedup_stride)
    T = np.zeros((S.shape[0], width), np.float32)
    w = min(width, S.shape[1])
    T[:, :w] = S[:, :w]
    b.speedup = np.ascontiguousarray(T.reshape(-1))
    b.speedup_stride = width
    return b


@pytest.mark.parametrize("algo", ALGO_NAMES)
def test_fuzz(ctx, oracle, algo):
    rng = np.random.default_rng(zlib.crc32(algo.encode()) + int(os.environ.get("VODA_FUZZ_SEED", "0")))
    for case in range(CASES):
        jobs = int(rng.choice([1, 2, 5, 17, 33, 64, 65, 97, 130, 260]))
        total = int(rng.choice([0, 1, 3, jobs, 4 * jobs, 40 * jobs + 7]))
        rounds = int(rng.integers(3, 40))
        quirky = bool(rng.integers(0, 2))
        seed = int(rng.integers(1, 1 << 30))
        b = (synth.quirk_batch if quirky else synth.jobs_batch)(algo, rounds, jobs, total, seed)
        width = int(rng.choice([2, 5, 9, 34, 40]))
        b = _reshape_table(b, width)
        b.queue_num = int(rng.choice([0, 1, 2, 3]))
        b.compaction_threshold = int(rng.choice([-1, 0, 3, 10, 1000]))
        if rng.integers(0, 3) == 0:
            b.max_proc = (b.max_proc.astype(np.int64) * int(rng.choice([1, 3, 20]))).astype(np.int32)
        g_ref, s_ref = oracle.allocate(b, dense=True, threads=8)
        g, s = ctx.allocate(b)
        tag = f"{algo} case {case}: jobs={jobs} total={total} rounds={rounds} W={width} Q={b.queue_num} " \
              f"thr={b.compaction_threshold} quirky={quirky} seed={seed}"
        assert np.array_equal(s, s_ref), tag
        assert np.array_equal(g, g_ref), tag
