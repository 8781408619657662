"""Measurement of the speedup-curve builder: jobs/s with the records resident in
HBM (CUDA events on the context's stream), end to end from host arrays, algorithmic bytes against
the measured HBM peak, and the CPU restatement timed on one core for a bounded sample.
    python profiles/bench_curves.py [--jobs 200000] [--records 120]  ->  one JSON line
The CPU arm imports oracle/ (checker and timed baseline only)."""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from vodascheduler_b200 import Context, _abi  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--jobs", type=int, default=200000)
ap.add_argument("--records", type=int, default=120)
ap.add_argument("--steps", type=int, default=20)
args = ap.parse_args()

W = 34
rng = np.random.default_rng(7)
counts = rng.integers(args.records // 2, args.records * 3 // 2 + 1, size=args.jobs)
off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
workers = rng.choice([1, 1, 2, 2, 4, 8, 16, 32], size=off[-1]).astype(np.int32)
t = np.exp(rng.normal(3.0, 1.0, size=off[-1])) / workers

ctx = Context(0)
dev = "cuda:0"
stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
d_off, d_w, d_t = (torch.from_numpy(x).to(dev) for x in (off, workers, t))
d_out = [torch.empty(args.jobs * W, dtype=torch.float64, device=dev) for _ in range(3)]
d_st = torch.empty(args.jobs, dtype=torch.int32, device=dev)
s = _abi.CurveBatchStruct(args.jobs, W, d_off.data_ptr(), d_w.data_ptr(), d_t.data_ptr())


def step():
    rc = ctx._lib.voda_build_curves_dev(ctx._h, C.byref(s), d_out[0].data_ptr(), d_out[1].data_ptr(),
                                        d_out[2].data_ptr(), d_st.data_ptr())
    assert rc == 0


torch.cuda.synchronize()
for _ in range(3):
    step()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(stream):
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
ctx.sync()
ms = ev0.elapsed_time(ev1) / args.steps
t0 = time.perf_counter()
for _ in range(3):
    out = ctx.build_curves(off, workers, t, W)
e2e_ms = (time.perf_counter() - t0) / 3 * 1e3
# CPU restatement on a bounded sample, one core
from oracle import curves  # noqa: E402
n_cpu = min(args.jobs, 300)
t0 = time.perf_counter()
for j in range(n_cpu):
    ref = curves.build_curves(workers[off[j]:off[j + 1]], t[off[j]:off[j + 1]], W)
cpu_s = time.perf_counter() - t0
ok = all(out[i][n_cpu - 1].tobytes() == ref[i].tobytes() for i in range(3))
alg_bytes = 12 * int(off[-1]) + args.jobs * (4 + 3 * 8 * W + 4)
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 7700.0
print(json.dumps({"metric": "speedup curves built per second", "value": args.jobs / (ms / 1e3), "unit": "jobs/s",
                  "ms_per_step": ms, "config": {"workload": f"{args.jobs} jobs x ~{args.records} epoch records, width {W}"},
                  "e2e": {"value": args.jobs / (e2e_ms / 1e3), "unit": "jobs/s", "h2d_bytes_per_step": 12 * int(off[-1]) + 4 * args.jobs,
                          "d2h_bytes_per_step": args.jobs * (3 * 8 * W + 4)},
                  "roofline": {"bound": "hbm", "achieved": alg_bytes / (ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                               "frac": alg_bytes / (ms / 1e3) / 1e9 / peak, "traffic": None},
                  "cpu_baseline": {"value": n_cpu / cpu_s, "unit": "jobs/s", "cores": 1, "kind": "port",
                                   "sample": f"first {n_cpu} jobs, pure-Python restatement of the pandas/numpy arithmetic"},
                  "verified_vs_oracle": bool(ok)}))
