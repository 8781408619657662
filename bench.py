#!/usr/bin/env python
"""bench.py — throughput of the allocation/placement hot path on B200.

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic rounds (SURVEY.md Appendix D
generators).  The default workload is BASELINE.json configs[1]: FfDL Optimizer DP allocation,
10k batched rounds of 256 jobs x 1024 GPUs.  Rounds are independent, so N GPUs each take their
own 10k rounds (weak scaling, no data-path collective).

One JSON line on stdout (rank 0):
  value     rounds/s, inputs already resident in HBM, CUDA events on the launching stream
  e2e       rounds/s through the C ABI with pinned HOST buffers (H2D + kernels + D2H timed)
  roofline  algorithmic bytes (compulsory input+output bytes per round) per
            launch / measured launch time, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the CPU oracle (a C++ port of the Go reference; Go is not installable here)
            timed on this box's host cores on a bounded sample of the same workload
`--impl reference` times that CPU port alone, with every host thread, on the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from vodascheduler_b200 import synth  # noqa: E402

WORKLOADS = {
    # BASELINE.json configs[1] — the headline
    "ffdl_c2": dict(kind="alloc", algo="FfDLOptimizer", rounds=10000, jobs=256, total=1024,
                    desc="FfDL Optimizer DP allocation, 10k batched rounds of 256 jobs x 1024 GPUs (BASELINE configs[1])"),
    # configs[2]
    # 2368 rounds = 4 full waves of the kernel's 592 resident CTAs on a B200 (a round is one long
    # dependent chain, so a partial last wave costs a whole wave: 2000 rounds measure 13 % lower)
    "afsl_c3": dict(kind="alloc", algo="AFS-L", rounds=2368, jobs=1024, total=4096,
                    desc="AFS-L greedy allocation, batched rounds of 1024 jobs x 4096 GPUs (BASELINE configs[2], 2368-round batch)"),
    # configs[0] batched
    "tiresias_c1": dict(kind="alloc", algo="ElasticTiresias", rounds=200000, jobs=64, total=16,
                        desc="Elastic-Tiresias, batched rounds of 64 jobs on 16 GPUs (BASELINE configs[0], 200k-round batch)"),
    "esrjf": dict(kind="alloc", algo="ElasticSRJF", rounds=20000, jobs=512, total=1024,
                  desc="Elastic-SRJF, 20k batched rounds of 512 jobs x 1024 GPUs (the policy of BASELINE configs[4])"),
    "efifo": dict(kind="alloc", algo="ElasticFIFO", rounds=20000, jobs=512, total=1024,
                  desc="Elastic-FIFO, 20k batched rounds of 512 jobs x 1024 GPUs"),
    # configs[3]
    "munkres_c4": dict(kind="munkres", n=128, rounds=20000,
                       desc="Placement munkres Hungarian, batched 128x128 worker-to-node score matrices (BASELINE configs[3], 20k-matrix batch)"),
    "place": dict(kind="place", nodes=128, slots=8, jobs=160, rounds=4000,
                  desc="bestFit + score + munkres placement rounds, 128 nodes x 8 slots"),
    # configs[4], scaled down: the full 100k-job trace is `--rounds 100000`
    "replay_c5": dict(kind="replay", replicas=8, rounds=1500, nodes=128, slots=8,
                      desc="Synthetic Philly-like trace replay, Elastic-SRJF + placement, 8 what-if replicas per GPU x "
                           "1500 jobs on 128 nodes x 8 GPUs (BASELINE configs[4] scaled down; --rounds sets jobs per replica)"),
}

W = synth.W
# compulsory bytes per job at the reference's dtypes and 8 bytes per round
BYTES_PER_JOB = {"ElasticFIFO": 24, "FIFO": 24, "ElasticSRJF": 16, "SRJF": 16,
                 "ElasticTiresias": 32 + 4 * W, "Tiresias": 32, "FfDLOptimizer": 24 + 4 * W,
                 "AFS-L": 28 + 4 * W}


def alg_bytes(wl):
    if wl["kind"] == "alloc":
        return wl["rounds"] * (wl["jobs"] * BYTES_PER_JOB[wl["algo"]] + 8)
    if wl["kind"] == "munkres":
        n = wl["n"]
        return wl["rounds"] * (8 * n * n + 4 * n + 4)
    return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        rows = [r for (t, r) in self.rows if t0 - 0.1 <= t <= t1 + 0.3] or [r for (_, r) in self.rows]
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except (ValueError, IndexError):
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_inputs(wl, rank):
    """Synthetic inputs for one rank (its own rounds: round_base keeps ranks distinct)."""
    if wl["kind"] == "alloc":
        return synth.jobs_batch(wl["algo"], wl["rounds"], wl["jobs"], wl["total"],
                                round_base=rank * wl["rounds"])
    if wl["kind"] == "munkres":
        return synth.matrices(wl["rounds"], wl["n"], kind="placement", seed=synth.SEED + rank)
    return synth.place_batch(wl["rounds"], wl["nodes"], wl["slots"], wl["jobs"], seed=synth.SEED + rank)


def cpu_run(wl, inputs, rounds, threads, dense):
    """Time the CPU oracle on the first `rounds` rounds; returns rounds/s."""
    from oracle import oracle
    if wl["kind"] == "alloc":
        sub = inputs.slice_rounds(0, rounds)
        t = time.perf_counter()
        oracle.allocate(sub, dense=dense, threads=threads)
        return rounds / (time.perf_counter() - t)
    if wl["kind"] == "munkres":
        t = time.perf_counter()
        oracle.munkres_max(inputs[:rounds], threads=threads)
        return rounds / (time.perf_counter() - t)
    raise NotImplementedError


def cpu_sample_size(wl, inputs, threads, dense, target_s):
    probe = max(1, min(wl["rounds"], threads))
    rate = cpu_run(wl, inputs, probe, threads, dense)
    return int(max(probe, min(wl["rounds"], rate * target_s)))


def run_reference(args, wl, rank):
    """--impl reference: the CPU port of the reference, every host thread, bounded sample."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample_wl = dict(wl)
    inputs = make_inputs(dict(wl, rounds=min(wl["rounds"], 4096)), 0)
    sample_wl["rounds"] = min(wl["rounds"], 4096)
    per_step = cpu_sample_size(sample_wl, inputs, threads, False, target_s=max(1.0, 60.0 / (args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_run(sample_wl, inputs, per_step, threads, False)
    t = time.perf_counter()
    for _ in range(args.steps):
        cpu_run(sample_wl, inputs, per_step, threads, False)
    dt = time.perf_counter() - t
    val = per_step * args.steps / dt
    sample = f"{per_step} rounds per step of the same workload, structure-faithful C++ port of the Go code"
    print(json.dumps({
        "impl": "reference", "metric": "alloc+placement rounds/sec", "value": val, "unit": "rounds/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "name": args.workload},
        "cpu_baseline": {"value": val, "unit": "rounds/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "rounds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference Go timing unavailable: no Go toolchain on this image; this is oracle/ (C++ restatement)",
    }))


def run_replay(args, wl, rank, world, ctx, dev, dist):
    """configs[4]: every rank replays its own replicas (replicas only: a trace is sequential in
    simulated time).  One 'step' = the whole replay; value = allocation+placement rounds / s of
    wall clock, host event loop included (so value == e2e here)."""
    import torch
    from vodascheduler_b200 import replay as rp
    t_all, rounds_all = 0.0, 0
    stats = None
    for it in range(max(1, min(args.steps, 3))):
        reps = rp.make_replicas(wl["replicas"], wl["rounds"], wl["nodes"], wl["slots"], first=rank * wl["replicas"])
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        l0 = ctx.launches
        t0 = time.perf_counter()
        stats = rp.replay(reps, ctx)
        dt = time.perf_counter() - t0
        t_all += dt
        rounds_all += stats["rounds"]
    t = torch.tensor([t_all, float(rounds_all)], dtype=torch.float64, device=dev)
    if world > 1:
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ts = t.clone()
        dist.all_reduce(ts, op=dist.ReduceOp.SUM)
        t_all, rounds_all = float(tm[0]), float(ts[1])
    if rank == 0:
        val = rounds_all / t_all
        print(json.dumps({
            "metric": "alloc+placement rounds/sec", "value": val, "unit": "rounds/s", "n_gpus": world,
            "steps": max(1, min(args.steps, 3)), "warmup": 0, "ms_per_step": 1e3 * t_all / max(1, min(args.steps, 3)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "name": args.workload, "replicas_per_gpu": wl["replicas"],
                       "jobs_per_replica": wl["rounds"], "parallelism": "replicas only (a trace is sequential in time)"},
            "e2e": {"value": val, "unit": "rounds/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None},
            "gpu_launches": int(ctx.launches - l0), "roofline": None, "cpu_baseline": None,
            "replay": {"ticks": stats["ticks"], "replica0": stats["replicas"][0]},
        }))
    if world > 1:
        dist.destroy_process_group()
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="ffdl_c2", choices=sorted(WORKLOADS))
    ap.add_argument("--rounds", type=int, default=0, help="override rounds per GPU")
    ap.add_argument("--e2e-steps", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = dict(WORKLOADS[args.workload])
    if args.rounds:
        wl["rounds"] = args.rounds
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, wl, rank)
        return

    import torch
    import torch.distributed as dist
    from vodascheduler_b200 import Context

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    ctx = Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    if wl["kind"] == "replay":
        return run_replay(args, wl, rank, world, ctx, dev, dist if world > 1 else None)
    inputs = make_inputs(wl, rank)
    R = wl["rounds"]

    # ---- device-resident arm ------------------------------------------------------------
    if wl["kind"] == "alloc":
        dbatch = inputs.to_device(dev)
        d_out = torch.zeros(inputs.jobs, dtype=torch.int32, device=dev)
        d_st = torch.full((R,), -1, dtype=torch.int32, device=dev)
        in_bytes = inputs.nbytes()

        def step_dev():
            ctx.allocate_dev(dbatch, d_out, d_st)

        hbatch = ctx.pin_batch(inputs)
        h_out = ctx.pinned((inputs.jobs,), np.int32)
        h_st = ctx.pinned((R,), np.int32)

        def step_e2e():
            ctx.allocate(hbatch, h_out, h_st)

        h2d = in_bytes
        d2h = 4 * inputs.jobs + 4 * R
    elif wl["kind"] == "munkres":
        n = wl["n"]
        d_A = torch.from_numpy(inputs).to(dev)
        d_out = torch.zeros(R * n, dtype=torch.int32, device=dev)
        d_st = torch.full((R,), -1, dtype=torch.int32, device=dev)
        in_bytes = inputs.nbytes

        def step_dev():
            ctx.munkres_max_dev(n, R, d_A, d_out, d_st)

        h_A = ctx.pinned(inputs.shape, np.int64)
        h_A[...] = inputs
        lib, hnd = ctx._lib, ctx._h
        h_out = ctx.pinned((R * n,), np.int32)
        h_st = ctx.pinned((R,), np.int32)

        def step_e2e():
            rc = lib.voda_munkres_max(hnd, n, R, h_A.ctypes.data, h_out.ctypes.data, h_st.ctypes.data)
            if rc:
                raise SystemExit(f"voda_munkres_max rc={rc}")

        h2d, d2h = in_bytes, 4 * R * n + 4 * R
    else:
        from vodascheduler_b200.batch import PlaceOut
        pout = PlaceOut(R, inputs.nodes, inputs.requests)
        in_bytes = sum(getattr(inputs, k).nbytes for k in ("node_off", "node_slots", "req_off", "req_workers", "prev_off", "prev_job", "prev_workers"))
        step_dev = None

        def step_e2e():
            ctx.place(inputs, pout)

        h2d = in_bytes
        d2h = 4 * (2 * inputs.nodes + 3 * (inputs.nodes + inputs.requests) + 3 * R) + 8 * R

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity spot check before any timing is accepted ---------------------------------
    verified = None
    if rank == 0 and not args.no_cpu and wl["kind"] == "alloc":
        from oracle import oracle
        nchk = min(R, 16)
        sub = inputs.slice_rounds(0, nchk)
        g_ref, s_ref = oracle.allocate(sub, dense=True, threads=os.cpu_count() or 1)
        g, s = ctx.allocate(sub)
        verified = bool(np.array_equal(g, g_ref) and np.array_equal(s, s_ref))
        if not verified:
            raise SystemExit("GPU result differs from the oracle on the verification sample; no timing reported")

    sampler = ClockSampler(local)
    if step_dev is None:
        step_dev = step_e2e
    for _ in range(args.warmup):
        step_dev()
    ctx.sync()
    barrier()
    l0 = ctx.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step_dev()
    ev1.record(stream)
    ctx.sync()
    barrier()
    t_wall1 = time.time()
    launches = ctx.launches - l0
    ms = ev0.elapsed_time(ev1) if wl["kind"] != "place" else 1e3 * (t_wall1 - t_wall0)

    # ---- end-to-end arm (host buffers through the C ABI) ---------------------------------
    e2e_steps = args.e2e_steps or max(3, min(args.steps, 10))
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    t_wall2 = time.time()
    clocks = sampler.stop(t_wall0, t_wall2)

    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])

    if rank == 0:
        total_rounds = R * world
        value = total_rounds * args.steps / (ms / 1e3)
        e2e_val = total_rounds * e2e_steps / (e2e_ms / 1e3)
        peak, peak_src = peaks()
        ab = alg_bytes(wl)
        roof = None
        if ab is not None:
            # the step is the dominant kernel plus (FfDL only) its small backtrack launch, ~10 % of the
            # step in the ncu launch list (profiles/r1_final_launches.csv): the step time is used as
            # the launch time, which can only understate the achieved figure
            kernel_ms = ms / args.steps
            ach = ab / (kernel_ms / 1e3) / 1e9
            traffic = None
            tpath = os.path.join(ROOT, "profiles", "r1_traffic.json")
            if os.path.exists(tpath):  # DRAM bytes of the dominant kernel from the committed ncu capture
                with open(tpath) as f:
                    t = json.load(f).get(args.workload)
                if t:
                    traffic = t["bytes_per_round"] * R
            roof = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": traffic, "peak_source": peak_src, "kernel": f"{wl.get('algo', wl['kind'])} batch kernel",
                    "algorithmic_bytes_per_launch": ab,
                    "note": "compute/latency-bound path: HBM fraction is low by nature (SURVEY 8d); see profiles/"}
        cpu = None
        if world == 1 and not args.no_cpu and wl["kind"] in ("alloc", "munkres"):
            threads = os.cpu_count() or 1
            n_f = cpu_sample_size(wl, inputs, threads, False, 10.0)
            v_f = cpu_run(wl, inputs, n_f, threads, False)
            cpu = {"value": v_f, "unit": "rounds/s", "cores": threads, "kind": "port",
                   "sample": f"first {n_f} rounds of the same workload; structure-faithful C++ port of the Go code, {threads} threads"}
            if wl["kind"] == "alloc":
                n_d = cpu_sample_size(wl, inputs, threads, True, 5.0)
                cpu["dense_port_value"] = cpu_run(wl, inputs, n_d, threads, True)
                n_1 = cpu_sample_size(wl, inputs, 1, True, 3.0)
                cpu["dense_port_1thread"] = cpu_run(wl, inputs, n_1, 1, True)
        out = {
            "metric": "alloc+placement rounds/sec", "value": value, "unit": "rounds/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl["desc"], "name": args.workload, "rounds_per_gpu": R,
                       "l2": f"inputs {in_bytes / 1e6:.0f} MB per GPU vs 126 MB L2: every step re-reads from HBM" if in_bytes > 130e6
                       else f"inputs {in_bytes / 1e6:.0f} MB per GPU (fits L2)",
                       "parallelism": f"rounds sharded over {world} GPU(s), no collective on the data path"},
            "e2e": {"value": e2e_val, "unit": "rounds/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps},
            "gpu_launches": int(launches),
            "roofline": roof, "cpu_baseline": cpu, "clocks": clocks, "verified_vs_oracle": verified,
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
