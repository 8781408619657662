"""CPU restatement of the scheduler's between-round bookkeeping — TEST INFRASTRUCTURE, never used
by the product path.  Plain Python loops in the reference's statement order (small cases only):
  tick()             pkg/scheduler/scheduler/scheduler.go:769-803 (updateTimeMetrics loop body) with
                     pkg/algorithm/tiresias.go:27-36 (thresholds, knob) and :109-119 (demote / promote)
  compare_results()  scheduler.go:448-480
Durations are Go time.Duration: int64 nanoseconds with wrapping arithmetic; Duration.Seconds()
as in time/time.go.  Parity unpinned (no Go toolchain here): pinned only by the hand-written
cases of tests/test_tick_cpu.py."""
import math

RUNNING, WAITING, COMPLETED, FAILED = 2, 1, 3, 4
M64 = (1 << 64) - 1


def _i64(x):  # wrap to int64 like Go
    x &= M64
    return x - (1 << 64) if x >> 63 else x


def _trunc_divmod(a, b):  # Go's / and % truncate toward zero
    q = abs(a) // abs(b)
    q = q if (a >= 0) == (b >= 0) else -q
    return q, a - q * b


def duration_seconds(d):
    sec, nsec = _trunc_divmod(d, 1_000_000_000)
    return float(sec) + float(nsec) / 1e9


def tick(state, now_ns, tiresias, thresholds_sec=(3600.0, math.inf), promote_knob=8):
    """state: dict of lists (same fields as voda_tick_batch); updated in place; returns `changed`."""
    n = len(state["status"])
    qn = len(thresholds_sec)
    changed = [0] * n
    for i in range(n):
        status = state["status"][i]
        elapsed = _i64(now_ns - state["last_update_ns"][i])
        g = state["num_gpu"][i]
        if status == RUNNING:
            state["running_ns"][i] = _i64(state["running_ns"][i] + elapsed)
            state["gpu_ns"][i] = _i64(state["gpu_ns"][i] + _i64(elapsed * g))
            state["total_ns"][i] = _i64(state["total_ns"][i] + elapsed)
            state["last_running_ns"][i] = _i64(state["last_running_ns"][i] + elapsed)
            state["last_gpu_ns"][i] = _i64(state["last_gpu_ns"][i] + _i64(elapsed * g))
        elif status == WAITING:
            state["waiting_ns"][i] = _i64(state["waiting_ns"][i] + elapsed)
            state["total_ns"][i] = _i64(state["total_ns"][i] + elapsed)
            state["last_waiting_ns"][i] = _i64(state["last_waiting_ns"][i] + elapsed)
        state["last_update_ns"][i] = now_ns
        if tiresias and status in (RUNNING, WAITING):
            pr = state["priority"][i]
            thr = thresholds_sec[pr] if 0 <= pr < qn else 0.0  # Go map miss
            if duration_seconds(state["last_gpu_ns"][i]) > thr:
                state["priority"][i] = pr + 1 if pr < qn - 1 else pr
                changed[i] = 1
            elif state["last_waiting_ns"][i] >= _i64(state["last_running_ns"][i] * promote_knob) and pr > 0:
                state["priority"][i] = 0
                changed[i] = 1
    return changed


def compare_results(old_gpu, new_gpu, status):
    out = []
    for n, m, st in zip(old_gpu, new_gpu, status):
        if n > m:
            out.append((1 if st not in (COMPLETED, FAILED) else 0) if m == 0 else 2)
        elif n < m:
            out.append(4 if n == 0 else 3)
        else:
            out.append(0)
    return out
