"""voda_tick / voda_compare_results through the C ABI against the restatement of
scheduler.go:769-803 and :448-480 on random job states (integers: bit-exact)."""
import math

import numpy as np
import pytest

from oracle import tick as T

pytestmark = pytest.mark.gpu
S = 1_000_000_000
FIELDS = ("last_update_ns", "running_ns", "gpu_ns", "total_ns", "waiting_ns", "last_running_ns", "last_gpu_ns",
          "last_waiting_ns")


def _random_state(n, seed):
    rng = np.random.default_rng(seed)
    st = {"status": rng.integers(0, 5, size=n).astype(np.int32), "num_gpu": rng.integers(0, 33, size=n).astype(np.int32),
          "priority": rng.choice([0, 0, 1, 1, 2, -1, 7], size=n).astype(np.int32)}
    for k in FIELDS:
        st[k] = (rng.integers(0, 5000, size=n) * S + rng.integers(0, S, size=n)).astype(np.int64)
    # thresholds hit exactly, and a few values near the int64 edge (Go wraps)
    st["last_gpu_ns"][::17] = 3600 * S
    st["gpu_ns"][::29] = (1 << 63) - 1 - rng.integers(0, 50, size=st["gpu_ns"][::29].size)
    st["last_running_ns"][::13] = 0
    return st


@pytest.mark.parametrize("tiresias", [False, True])
@pytest.mark.parametrize("n", [1, 33, 5000])
def test_tick_matches_the_restatement(ctx, tiresias, n):
    st = _random_state(n, 11 + n)
    ref = {k: [int(x) for x in v] for k, v in st.items()}
    now = 6000 * S + 123
    want = T.tick(ref, now, tiresias)
    got = ctx.tick(st, now, tiresias)
    assert got.tolist() == want
    for k in ("priority",) + FIELDS:
        assert st[k].tolist() == ref[k], k


def test_tick_queue_tables(ctx):
    st = _random_state(400, 5)
    ref = {k: [int(x) for x in v] for k, v in st.items()}
    thr = (100.0, 2000.0, math.inf)
    want = T.tick(ref, 7000 * S, True, thr, promote_knob=3)
    got = ctx.tick(st, 7000 * S, True, thr, promote_knob=3)
    assert got.tolist() == want and st["priority"].tolist() == ref["priority"]


def test_compare_results(ctx):
    rng = np.random.default_rng(3)
    old = rng.integers(0, 6, size=3000).astype(np.int32)
    new = rng.integers(0, 6, size=3000).astype(np.int32)
    status = rng.integers(0, 5, size=3000).astype(np.int32)
    assert ctx.compare_results(old, new, status).tolist() == T.compare_results(old.tolist(), new.tolist(), status.tolist())
    assert ctx.compare_results(old[:0], new[:0], status[:0]).size == 0
