"""Hand-traced cases for the bookkeeping oracle (scheduler.go:769-803, :448-480; tiresias.go:27-36,
109-119).  The reference has no test for these functions, so these traces are the pin."""
import math

from oracle import tick as T

S = 1_000_000_000


def _state(**kw):
    n = len(kw["status"])
    base = {k: [0] * n for k in ("num_gpu", "priority", "last_update_ns", "running_ns", "gpu_ns", "total_ns",
                                 "waiting_ns", "last_running_ns", "last_gpu_ns", "last_waiting_ns")}
    base.update(kw)
    return base


def test_duration_accounting_by_status():
    st = _state(status=[2, 1, 3, 0], num_gpu=[4, 0, 2, 1], last_update_ns=[10 * S] * 4)
    ch = T.tick(st, 40 * S, tiresias=False)
    assert ch == [0, 0, 0, 0]
    assert st["running_ns"] == [30 * S, 0, 0, 0] and st["gpu_ns"] == [120 * S, 0, 0, 0]
    assert st["waiting_ns"] == [0, 30 * S, 0, 0] and st["total_ns"] == [30 * S, 30 * S, 0, 0]
    assert st["last_gpu_ns"] == [120 * S, 0, 0, 0] and st["last_update_ns"] == [40 * S] * 4


def test_tiresias_demote_and_promote():
    # job 0: 901 s on 4 GPUs -> LastGpuDuration 3604 s > 3600 -> demoted 0 -> 1
    # job 1: exactly 3600 s of GPU time is not > the threshold: stays
    # job 2: priority 1 (threshold +Inf), waited 80 s, ran 10 s, 80 >= 10*8 -> promoted to 0
    # job 3: priority 1, waited 79 s -> stays;  job 4: priority 7 names no queue: threshold 0 -> "demoted" (stays 7, flag set)
    st = _state(status=[2, 2, 1, 1, 2], num_gpu=[4, 1, 0, 0, 1], priority=[0, 0, 1, 1, 7],
                last_update_ns=[0] * 5, last_running_ns=[0, 0, 10 * S, 10 * S, 0], last_waiting_ns=[0, 0, 0, 0, 0])
    st["last_update_ns"] = [0, 0, 0, 1 * S, 0]
    ch = T.tick(st, 80 * S, tiresias=True)  # elapsed 80 s (job 3: 79 s)
    assert st["last_gpu_ns"][0] == 320 * S
    st2 = _state(status=[2, 2], num_gpu=[4, 1], priority=[0, 0], last_update_ns=[0, 0])
    ch2 = T.tick(st2, 901 * S, tiresias=True)
    assert ch2 == [1, 0] and st2["priority"] == [1, 0]
    st3 = _state(status=[2], num_gpu=[1], priority=[0], last_update_ns=[0])
    assert T.tick(st3, 3600 * S, tiresias=True) == [0] and T.tick(st3, 3600 * S + 1, tiresias=True) == [1]
    assert ch[2] == 1 and st["priority"][2] == 0
    assert ch[3] == 0 and st["priority"][3] == 1
    assert ch[4] == 1 and st["priority"][4] == 7


def test_duration_seconds_and_wrap():
    assert T.duration_seconds(1_500_000_000) == 1.5 and T.duration_seconds(-1_500_000_000) == -1.5
    st = _state(status=[2], num_gpu=[3], last_update_ns=[0], gpu_ns=[(1 << 63) - 5])
    T.tick(st, 10, tiresias=False)
    assert st["gpu_ns"] == [-(1 << 63) + 25]  # int64 wrap, as in Go


def test_compare_results():
    old = [4, 4, 4, 0, 2, 3, 3]
    new = [4, 0, 2, 2, 5, 0, 0]
    status = [2, 2, 2, 1, 2, 3, 4]
    assert T.compare_results(old, new, status) == [0, 1, 2, 4, 3, 0, 0]
