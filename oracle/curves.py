"""CPU restatement of the speedup-curve builder — TEST INFRASTRUCTURE, never used by the product
path.  Follows python/metrics_collector/metrics_collector.py:131-167:
  _update_time_metric :139-141  epoch_time[str(w)] = mean of the epoch times recorded with w workers
  _update_speedup     :150-154  speedup[k] = epoch_time['1'] / epoch_time[k]   for every key k != '0'
  _update_efficiency  :163-167  efficiency[k] = speedup[k] / float(k)
over the dictionaries CreateBaseJobInfo (pkg/common/mongo/mongo.go:76-100) starts a job with
(time 1.0, speedup k, efficiency 1.0 for k = 1..33, and 0 at k = 0).  Everything is float64 here;
the mean is pandas' (sum / count) and the sum is numpy's pairwise summation, restated below so
that the result is bit-identical.  PINNED: tests/test_curves_cpu.py checks this file against
tests/golden/curve_kats.json, produced by running the reference's own three methods
(tests/golden/make_curve_kats.py).
"""
import numpy as np

PW_BLOCK = 128  # numpy's PW_BLOCKSIZE


def pairwise_sum(a):
    """numpy/core/src/umath/loops_utils.h.src  @TYPE@_pairwise_sum for a contiguous float64 run."""
    n = len(a)
    if n < 8:
        res = 0.0  # (numpy starts from -0.0 only for n == 0 results; adding 0.0 first is the same for n >= 1)
        for x in a:
            res += float(x)
        return res if n else 0.0
    if n <= PW_BLOCK:
        r = [float(a[i]) for i in range(8)]
        i = 8
        while i < n - (n % 8):
            for u in range(8):
                r[u] += float(a[i + u])
            i += 8
        res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]))
        while i < n:
            res += float(a[i])
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return pairwise_sum(a[:n2]) + pairwise_sum(a[n2:])


def build_curves(workers, epoch_time_sec, width=34):
    """One job: records (workers[i], epoch_time_sec[i]) -> (epoch_time, speedup, efficiency), dense
    float64 arrays indexed by worker count; `width` grows to hold the largest count seen."""
    workers = np.asarray(workers, np.int64)
    t = np.asarray(epoch_time_sec, np.float64)
    width = max(width, int(workers.max()) + 1 if workers.size else 0)
    et = np.ones(width)
    et[0] = 0.0
    for w in np.unique(workers):
        if w < 0:
            continue
        vals = t[workers == w]
        et[w] = pairwise_sum(vals) / float(len(vals))
    sp = np.zeros(width)
    ef = np.zeros(width)
    with np.errstate(divide="ignore", invalid="ignore"):
        sp[1:] = et[1] / et[1:]
        ef[1:] = sp[1:] / np.arange(1, width, dtype=np.float64)
    return et, sp, ef
