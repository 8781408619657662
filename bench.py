#!/usr/bin/env python
"""bench.py — throughput of the allocation + placement hot path on B200.

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--impl reference] [--only-headline]

A "step" is one pass of the hot path over one batch of synthetic rounds (SURVEY.md Appendix D
generators).  The HEADLINE is BASELINE.json configs[2], the configuration the north star is quoted
on: AFS-L greedy allocation, 10 000 batched rounds of 1024 jobs x 4096 GPUs per GPU.  Rounds are
independent, so N GPUs each take their own 10 000 rounds (weak scaling, no data-path collective).

One JSON line on stdout (rank 0):
  value      rounds/s, inputs already resident in HBM, CUDA events on the launching stream
  e2e        rounds/s through the C ABI with pinned HOST buffers (H2D + kernels + D2H timed); the
             byte counts are what the call actually copied (voda_ctx_copied_bytes)
  roofline   algorithmic bytes (compulsory input+output bytes per round) per launch /
             measured launch time, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the CPU oracle (a C++ port of the Go reference; Go is not installable here) timed on
             this box's host cores on a bounded sample: structure-faithful port (all threads) and
             the dense port (all threads, one thread)
  verified_vs_oracle  every rank compares >= 256 rounds of BOTH arms with the oracle before any
             number is printed
  per_config the other BASELINE configs on the same GPU(s), each with the same keys:
             tiresias_c1 (configs[0]: batched, and the latency of ONE round through the C ABI),
             ffdl_c2 (configs[1]), munkres_c4 (configs[3], 100 000 matrices generated on the
             device), alloc_place (Elastic-SRJF allocation feeding placement on the device, the
             order resched() runs them, scheduler.go:326-364), replay_c5 (configs[4])
`--impl reference` times the CPU port of the headline alone, with every host thread.
"""
from __future__ import annotations

import argparse
import ctypes as C
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

HEADLINE = "afsl_c3"
PER_CONFIG = ("tiresias_c1", "ffdl_c2", "munkres_c4", "alloc_place", "replay_c5")

WORKLOADS = {
    # BASELINE.json configs[2] — the headline
    "afsl_c3": dict(kind="alloc", algo="AFS-L", rounds=10000, jobs=1024, total=4096, steps=12, e2e_steps=4,
                    desc="AFS-L greedy allocation, 10k batched rounds of 1024 jobs x 4096 GPUs (BASELINE configs[2])"),
    # configs[1]
    "ffdl_c2": dict(kind="alloc", algo="FfDLOptimizer", rounds=10000, jobs=256, total=1024, steps=40, e2e_steps=10,
                    desc="FfDL Optimizer DP allocation, 10k batched rounds of 256 jobs x 1024 GPUs (BASELINE configs[1])"),
    # configs[0] batched
    "tiresias_c1": dict(kind="alloc", algo="ElasticTiresias", rounds=200000, jobs=64, total=16, steps=40, e2e_steps=6,
                        desc="Elastic-Tiresias, 64 jobs on a 16-GPU cluster (BASELINE configs[0]): 200k-round batch, "
                             "and one round at a time"),
    "esrjf": dict(kind="alloc", algo="ElasticSRJF", rounds=20000, jobs=512, total=1024, steps=40, e2e_steps=10,
                  desc="Elastic-SRJF, 20k batched rounds of 512 jobs x 1024 GPUs (the policy of BASELINE configs[4])"),
    "efifo": dict(kind="alloc", algo="ElasticFIFO", rounds=20000, jobs=512, total=1024, steps=40, e2e_steps=10,
                  desc="Elastic-FIFO, 20k batched rounds of 512 jobs x 1024 GPUs"),
    # configs[3]
    "munkres_c4": dict(kind="munkres", n=128, rounds=100000, e2e_rounds=20000, steps=10, e2e_steps=3,
                       desc="Placement munkres Hungarian, 100k batched 128x128 worker-to-node score matrices "
                            "(BASELINE configs[3]); generated on the device; the end-to-end arm streams 20k of them "
                            "per step from a 2.6 GB pinned host buffer"),
    "place": dict(kind="place", nodes=128, slots=8, jobs=160, rounds=4000, steps=10, e2e_steps=5,
                  desc="bestFit + score + munkres placement rounds, 128 nodes x 8 slots"),
    "alloc_place": dict(kind="chain", nodes=128, slots=8, jobs=160, rounds=20000, steps=10, e2e_steps=4,
                        desc="allocation + placement per round, on the device with no host hop: Elastic-SRJF over 160 "
                             "ready jobs (870 GPUs) -> bestFit + score + munkres on 128 nodes x 8 slots "
                             "(resched(), scheduler.go:326-364)"),
    # configs[4]
    "replay_c5": dict(kind="replay", replicas=64, rounds=1562, nodes=128, slots=8, steps=1,
                      desc="Synthetic Philly-like trace replay, Elastic-SRJF + placement on 128 nodes x 8 GPUs: 64 what-if "
                           "replicas x 1562 jobs = 100k jobs per GPU, replicas sharded over the GPUs (BASELINE configs[4]; "
                           "--rounds sets jobs per replica)"),
}

W = synth.W
# compulsory bytes per job at the reference's dtypes and 8 bytes per round
BYTES_PER_JOB = {"ElasticFIFO": 24, "FIFO": 24, "ElasticSRJF": 16, "SRJF": 16,
                 "ElasticTiresias": 32 + 4 * W, "Tiresias": 32, "FfDLOptimizer": 24 + 4 * W,
                 "AFS-L": 28 + 4 * W}


def alg_bytes(wl, rounds=None):
    """Algorithmic bytes of one launch over `rounds` rounds."""
    R = wl["rounds"] if rounds is None else rounds
    if wl["kind"] == "alloc":
        return R * (wl["jobs"] * BYTES_PER_JOB[wl["algo"]] + 8)
    if wl["kind"] == "munkres":
        n = wl["n"]
        return R * (8 * n * n + 4 * n + 4)
    if wl["kind"] == "chain":
        # allocation (Elastic-SRJF: 16 B per job + 8) + placement: node slots and offsets (8 B per node),
        # the previous layout (8 B per entry, ~1.1 per running job), the request list is the
        # allocation's own output (not read again from the host), outputs: binding + free slots
        # (8 B per node), layout (12 B per entry), score/cross/status/count (20 B)
        return R * (wl["jobs"] * 16 + 8 + wl["nodes"] * 16 + wl["jobs"] * 8 + wl["jobs"] * 12 + 20)
    return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def traffic_per_unit(name):
    """DRAM bytes per round of the dominant kernel from the committed ncu capture of this round."""
    for fn in ("r2_traffic.json", "r1_traffic.json"):
        p = os.path.join(ROOT, "profiles", fn)
        if os.path.exists(p):
            with open(p) as f:
                t = json.load(f).get(name)
            if t:
                return t["bytes_per_round"], fn
    return None, None


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


def concat_batches(parts):
    """AllocBatch of the rounds of `parts`, in order."""
    from vodascheduler_b200.batch import AllocBatch, _ALLOC_ARRAYS
    first = parts[0]
    kw = {}
    for name, _ in _ALLOC_ARRAYS[2:]:
        vs = [getattr(p, name) for p in parts]
        kw[name] = None if vs[0] is None else np.concatenate(vs)
    offs, base = [np.zeros(1, np.int64)], 0
    for p in parts:
        offs.append(p.job_off[1:].astype(np.int64) + base)
        base += int(p.job_off[-1])
    return AllocBatch(first.algo, np.concatenate(offs).astype(np.int32), np.concatenate([p.total_gpu for p in parts]),
                      speedup_stride=first.speedup_stride, **kw)


def make_alloc_inputs(wl, rank, rounds=None):
    """The rank's own rounds (round_base keeps ranks distinct), generated in pieces of <= 1M jobs."""
    R = wl["rounds"] if rounds is None else rounds
    per = max(1, min(R, (1 << 20) // max(1, wl["jobs"])))
    parts = [synth.jobs_batch(wl["algo"], min(per, R - r0), wl["jobs"], wl["total"], round_base=rank * wl["rounds"] + r0)
             for r0 in range(0, R, per)]
    return parts[0] if len(parts) == 1 else concat_batches(parts)


def oracle_alloc_rate(batch, rounds, threads, dense):
    from oracle import oracle
    sub = batch.slice_rounds(0, rounds)
    t = time.perf_counter()
    oracle.allocate(sub, dense=dense, threads=threads)
    return rounds / (time.perf_counter() - t)


def sized_rate(fn, total, threads, target_s):
    """fn(count) -> units/s on the first `count` units; a probe sizes the sample to about target_s."""
    probe = max(1, min(total, threads))
    rate = fn(probe)
    count = int(max(probe, min(total, rate * target_s)))
    return fn(count), count


# ---------------------------------------------------------------------------------------------
class Env:
    def __init__(self, rank, world, local, dev, dist):
        self.rank, self.world, self.local, self.dev, self.dist = rank, world, local, dev, dist
        self.threads = os.cpu_count() or 1

    def barrier(self):
        import torch
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, *vals):
        import torch
        t = torch.tensor(list(vals), dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def all_ok(self, ok):
        import torch
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return bool(int(t[0]))

    def gather_summaries(self, mine):
        """all-gather of the 4-number per-rank summary over NCCL (the only collective besides timing)."""
        import torch
        t = torch.from_numpy(np.asarray(mine, np.int64)).to(self.dev)
        if self.world == 1:
            return [t.cpu().numpy().tolist()]
        out = [torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [o.cpu().numpy().tolist() for o in out]


def time_steps(ctx, env, step, steps, warmup):
    """`warmup` untimed steps, then exactly `steps` steps between CUDA events on the context's
    stream, bracketed by barrier + synchronize; returns (ms total, launches)."""
    import torch
    stream = torch.cuda.ExternalStream(ctx.stream, device=env.dev)
    for _ in range(warmup):
        step()
    ctx.sync()
    env.barrier()
    l0 = ctx.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(steps):
        step()
    ev1.record(stream)
    ctx.sync()
    env.barrier()
    return ev0.elapsed_time(ev1), ctx.launches - l0


def time_host_steps(env, step, steps, warmup=2):
    for _ in range(warmup):
        step()
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    env.barrier()
    return 1e3 * (time.perf_counter() - t0)


def roofline_entry(name, wl, ms_per_launch, rounds, kernel):
    peak, peak_src = peaks()
    ab = alg_bytes(wl, rounds)
    if ab is None:
        return None
    ach = ab / (ms_per_launch / 1e3) / 1e9
    tpr, tsrc = traffic_per_unit(name)
    return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": None if tpr is None else tpr * rounds, "traffic_source": tsrc,
            "peak_source": peak_src, "kernel": kernel, "algorithmic_bytes_per_launch": ab,
            "note": "issue/latency-bound path: the HBM fraction is low by nature (SURVEY 8d); see profiles/"}


# ---------------------------------------------------------------------------------------------
def run_alloc(name, wl, ctx, env, steps, warmup, e2e_steps, with_cpu):
    import torch
    from oracle import oracle
    R = wl["rounds"]
    inputs = make_alloc_inputs(wl, env.rank)
    dbatch = inputs.to_device(env.dev)
    d_out = torch.zeros(inputs.jobs, dtype=torch.int32, device=env.dev)
    d_st = torch.full((R,), -1, dtype=torch.int32, device=env.dev)
    ms, launches = time_steps(ctx, env, lambda: ctx.allocate_dev(dbatch, d_out, d_st), steps, warmup)

    hbatch = ctx.pin_batch(inputs)
    h_out = ctx.pinned((inputs.jobs,), np.int32)
    h_st = ctx.pinned((R,), np.int32)
    e2e_ms = time_host_steps(env, lambda: ctx.allocate(hbatch, h_out, h_st), e2e_steps)
    h2d, d2h = ctx.copied_bytes

    # parity on EVERY rank, both arms, before any number is accepted
    nchk = min(R, max(256, 2 * env.threads))
    sub = inputs.slice_rounds(0, nchk)
    g_ref, s_ref = oracle.allocate(sub, dense=True, threads=max(1, env.threads // env.world))
    nj = int(inputs.job_off[nchk])
    g_dev, s_dev = d_out[:nj].cpu().numpy(), d_st[:nchk].cpu().numpy()
    ok = (np.array_equal(g_dev, g_ref) and np.array_equal(s_dev, s_ref) and
          np.array_equal(h_out[:nj], g_ref) and np.array_equal(h_st[:nchk], s_ref))
    g_all, s_all = d_out.cpu().numpy(), d_st.cpu().numpy()
    ok = ok and np.array_equal(g_all, h_out) and np.array_equal(s_all, h_st)  # device arm == host arm, whole batch
    from vodascheduler_b200 import shard
    sums = env.gather_summaries(shard.summary(g_all, s_all))
    verified = env.all_ok(ok)
    if not verified:
        raise SystemExit(f"{name}: GPU result differs from the oracle on the verification sample; no timing reported")
    ms, e2e_ms = env.max_over_ranks(ms, e2e_ms)

    entry = None
    if env.rank == 0:
        total = R * env.world
        entry = {
            "workload": wl["desc"], "value": total * steps / (ms / 1e3), "unit": "rounds/s", "ms_per_step": ms / steps,
            "steps": steps, "rounds_per_gpu": R,
            "e2e": {"value": total * e2e_steps / (e2e_ms / 1e3), "unit": "rounds/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
                    "h2d_gbs_per_gpu": h2d / (e2e_ms / e2e_steps / 1e3) / 1e9},
            "gpu_launches": int(launches),
            "roofline": roofline_entry(name, wl, ms / steps, R, f"{wl['algo']} batch kernel(s), one launch set per step"),
            "verified_vs_oracle": {"rounds_per_rank": nchk, "ranks": env.world, "arms": ["device", "host"], "ok": True},
            "rank_summaries": {"fields": ["rounds", "gpus_handed_out", "nonzero_status", "digest64"], "per_rank": sums},
            "l2": f"inputs {inputs.nbytes() / 1e6:.0f} MB per GPU vs 126 MB L2: every step re-reads them from HBM"
                  if inputs.nbytes() > 130e6 else f"inputs {inputs.nbytes() / 1e6:.0f} MB per GPU (fit L2; outputs are rewritten every step)",
        }
        if with_cpu:
            th = env.threads
            v_f, n_f = sized_rate(lambda c: oracle_alloc_rate(inputs, c, th, False), R, th, 8.0)
            v_d, n_d = sized_rate(lambda c: oracle_alloc_rate(inputs, c, th, True), R, th, 4.0)
            v_1, n_1 = sized_rate(lambda c: oracle_alloc_rate(inputs, c, 1, True), R, 1, 3.0)
            entry["cpu_baseline"] = {
                "value": v_f, "unit": "rounds/s", "cores": th, "kind": "port",
                "sample": f"first {n_f} rounds of the same workload; structure-faithful C++ port of the Go code "
                          f"(string-keyed maps, a stable sort per pick), {th} threads",
                "dense_port_value": v_d, "dense_port_sample": f"first {n_d} rounds, array-based port, {th} threads",
                "dense_port_1thread": v_1, "dense_port_1thread_sample": f"first {n_1} rounds"}
        else:
            entry["cpu_baseline"] = None
        if name == "tiresias_c1":
            entry["single_round"] = single_round_latency(wl, ctx, inputs, with_cpu)
    if inputs.speedup is not None and wl["algo"] in ("ElasticTiresias", "FfDLOptimizer", "AFS-L"):
        pooled = shared_curves_e2e(name, wl, ctx, env, inputs, e2e_steps)
        if entry is not None:
            entry["e2e_shared_curves"] = pooled
    del dbatch, d_out, d_st
    return entry


def shared_curves_e2e(name, wl, ctx, env, inputs, e2e_steps):
    """A SEPARATE, labelled end-to-end line (it never replaces `e2e`): the same batch shape as a what-if
    study over one job set — every round is the jobs of round 0 under its own total and remaining
    times — with the speedup curves passed once as a pool (voda_alloc_batch.speedup_off) instead of a
    dense table per round."""
    from oracle import oracle
    pooled, _ = synth.share_curves(inputs, want_dense=False)
    hb = ctx.pin_batch(pooled)
    R = pooled.rounds
    h_out = ctx.pinned((pooled.jobs,), np.int32)
    h_st = ctx.pinned((R,), np.int32)
    ms = time_host_steps(env, lambda: ctx.allocate(hb, h_out, h_st), e2e_steps)
    h2d, d2h = ctx.copied_bytes
    nchk = min(R, 256)
    g_ref, s_ref = oracle.allocate(pooled.slice_rounds(0, nchk), dense=True, threads=max(1, env.threads // env.world))
    ok = np.array_equal(h_out[:int(pooled.job_off[nchk])], g_ref) and np.array_equal(h_st[:nchk], s_ref)
    if not env.all_ok(ok):
        raise SystemExit(f"{name}: shared-curve result differs from the oracle; no timing reported")
    (ms,) = env.max_over_ranks(ms)
    return {"value": R * env.world * e2e_steps / (ms / 1e3), "unit": "rounds/s", "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": ms / e2e_steps,
            "workload": "what-if form of the same shape: all rounds over the jobs of round 0, curves passed once as a pool "
                        "(speedup_off); NOT comparable with `e2e` or the reference arm, which take a dense table per round",
            "verified_vs_oracle": {"rounds_per_rank": nchk, "ok": True}}


def single_round_latency(wl, ctx, inputs, with_cpu):
    """configs[0] as the reference runs it: ONE round per call, host buffers, through the C ABI."""
    from oracle import oracle
    one = inputs.slice_rounds(0, 1)
    pb = ctx.pin_batch(one)
    o = ctx.pinned((one.jobs,), np.int32)
    s = ctx.pinned((1,), np.int32)
    st = pb.struct()  # built once: the loop times the C ABI call, not Python's struct marshalling
    call, h, po, ps = ctx._lib.voda_allocate, ctx._h, o.ctypes.data, s.ctypes.data
    for _ in range(50):
        call(h, C.byref(st), po, ps)
    n = 2000
    t = time.perf_counter()
    for _ in range(n):
        call(h, C.byref(st), po, ps)
    us = 1e6 * (time.perf_counter() - t) / n
    g_ref, s_ref = oracle.allocate(one, dense=True)
    assert np.array_equal(o, g_ref) and np.array_equal(s, s_ref)
    out = {"us_per_round_gpu": us, "calls": n,
           "note": "voda_allocate on a 1-round batch (host buffers): gathered into one page-locked block that the kernel reads and writes directly (one launch, one wait)"}
    if with_cpu:
        m = 2000
        t = time.perf_counter()
        for _ in range(m):
            oracle.allocate(one, dense=False, threads=1)
        out["us_per_round_cpu_faithful"] = 1e6 * (time.perf_counter() - t) / m
        t = time.perf_counter()
        for _ in range(m):
            oracle.allocate(one, dense=True, threads=1)
        out["us_per_round_cpu_dense"] = 1e6 * (time.perf_counter() - t) / m
        out["cpu_note"] = ("one round per call through the ctypes binding of the CPU ports (same call overhead as the GPU "
                           "line); a single small round is launch-latency bound on the GPU: the tuned dense port is faster "
                           "there, batches are where the GPU pays off")
    return out


def run_munkres(name, wl, ctx, env, steps, warmup, e2e_steps, with_cpu):
    import torch
    from oracle import oracle
    n, R, Re = wl["n"], wl["rounds"], min(wl["rounds"], wl["e2e_rounds"])
    d_A = synth.matrices_device(R, n, env.dev, kind="placement", first=env.rank * R)
    d_out = torch.zeros(R * n, dtype=torch.int32, device=env.dev)
    d_st = torch.full((R,), -1, dtype=torch.int32, device=env.dev)
    ms, launches = time_steps(ctx, env, lambda: ctx.munkres_max_dev(n, R, d_A, d_out, d_st), steps, warmup)

    h_A = ctx.pinned((Re, n, n), np.int64)
    h_A[...] = d_A[:Re].cpu().numpy()
    h_out = ctx.pinned((Re * n,), np.int32)
    h_st = ctx.pinned((Re,), np.int32)
    lib, hnd = ctx._lib, ctx._h

    def step_e2e():
        rc = lib.voda_munkres_max(hnd, n, Re, h_A.ctypes.data, h_out.ctypes.data, h_st.ctypes.data)
        if rc:
            raise SystemExit(f"voda_munkres_max rc={rc}")

    e2e_ms = time_host_steps(env, step_e2e, e2e_steps, warmup=1)
    h2d, d2h = ctx.copied_bytes
    nchk = min(R, 256)
    c_ref, s_ref = oracle.munkres_max(np.ascontiguousarray(h_A[:nchk]), threads=max(1, env.threads // env.world))
    ok = (np.array_equal(d_out[:nchk * n].cpu().numpy().reshape(nchk, n), c_ref) and
          np.array_equal(h_out[:nchk * n].reshape(nchk, n), c_ref) and bool((d_st == 0).all()) and
          np.array_equal(d_out[:Re * n].cpu().numpy(), h_out))
    verified = env.all_ok(ok)
    if not verified:
        raise SystemExit(f"{name}: GPU result differs from the oracle; no timing reported")
    ms, e2e_ms = env.max_over_ranks(ms, e2e_ms)
    # iid uniform matrices, small and wide value ranges (8-bit and 16-bit cells)
    variants = {}
    for vname, hi in (("uniform_0_8", 8), ("uniform_0_999", 999)):
        Rv = min(R, 20000)
        dv = synth.matrices_device(Rv, n, env.dev, kind="uniform", hi=hi, first=env.rank * Rv)
        vms, _ = time_steps(ctx, env, lambda: ctx.munkres_max_dev(n, Rv, dv, d_out, d_st), 3, 2)
        c_ref, _ = oracle.munkres_max(np.ascontiguousarray(dv[:64].cpu().numpy()), threads=max(1, env.threads // env.world))
        if not env.all_ok(np.array_equal(d_out[:64 * n].cpu().numpy().reshape(64, n), c_ref)):
            raise SystemExit(f"{name}/{vname}: GPU result differs from the oracle; no timing reported")
        (vms,) = env.max_over_ranks(vms)
        variants[vname] = {"value": Rv * env.world * 3 / (vms / 1e3), "unit": "matrices/s", "matrices_per_gpu": Rv,
                           "verified_matrices_per_rank": 64}
        del dv
    entry = None
    if env.rank == 0:
        entry = {
            "workload": wl["desc"], "value": R * env.world * steps / (ms / 1e3), "unit": "matrices/s",
            "variants": variants,
            "ms_per_step": ms / steps, "steps": steps, "matrices_per_gpu": R,
            "e2e": {"value": Re * env.world * e2e_steps / (e2e_ms / 1e3), "unit": "matrices/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                    "ms_per_step": e2e_ms / e2e_steps, "matrices_per_step": Re,
                    "h2d_gbs_per_gpu": h2d / (e2e_ms / e2e_steps / 1e3) / 1e9},
            "gpu_launches": int(launches),
            "roofline": roofline_entry(name, wl, ms / steps, R, "munkres_kernel (cell-width ladder), one launch set per step"),
            "verified_vs_oracle": {"matrices_per_rank": nchk, "ranks": env.world, "arms": ["device", "host"], "ok": True},
        }
        if with_cpu:
            th = env.threads
            A = np.ascontiguousarray(h_A[:min(Re, 4096)])

            def rate(c):
                t = time.perf_counter()
                oracle.munkres_max(A[:c], threads=th)
                return c / (time.perf_counter() - t)

            v, cnt = sized_rate(rate, len(A), th, 6.0)
            entry["cpu_baseline"] = {"value": v, "unit": "matrices/s", "cores": th, "kind": "port",
                                     "sample": f"first {cnt} matrices; six-step Munkres restatement (the Go module is not in the reference tree), {th} threads"}
        else:
            entry["cpu_baseline"] = None
    del d_A, d_out, d_st
    return entry


def run_chain(name, wl, ctx, env, steps, warmup, e2e_steps, with_cpu):
    """allocation -> placement for every round, device-resident: voda_allocate_dev writes the per-job
    GPU counts, voda_place_dev reads them as its request list (same offsets), no host hop."""
    import torch
    from oracle import oracle
    from vodascheduler_b200 import _abi
    from vodascheduler_b200.batch import PlaceOut
    R = wl["rounds"]
    ab, pb = synth.chain_batch(R, wl["nodes"], wl["slots"], wl["jobs"], round_base=env.rank * R)
    dab = ab.to_device(env.dev)
    dpb = pb.to_device(env.dev)
    d_g = torch.zeros(ab.jobs, dtype=torch.int32, device=env.dev)
    d_s = torch.full((R,), -1, dtype=torch.int32, device=env.dev)
    dpb.req_workers = d_g  # the allocation's output IS the placement's request list
    cap = pb.nodes + pb.requests
    outs = {"phys_of_logical": torch.zeros(pb.nodes, dtype=torch.int32, device=env.dev),
            "free_slots": torch.zeros(pb.nodes, dtype=torch.int32, device=env.dev),
            "lay_cnt": torch.zeros(R, dtype=torch.int32, device=env.dev),
            "lay_node": torch.zeros(cap, dtype=torch.int32, device=env.dev),
            "lay_job": torch.zeros(cap, dtype=torch.int32, device=env.dev),
            "lay_workers": torch.zeros(cap, dtype=torch.int32, device=env.dev),
            "score": torch.zeros(R, dtype=torch.int64, device=env.dev),
            "cross": torch.zeros(R, dtype=torch.int32, device=env.dev),
            "status": torch.full((R,), -1, dtype=torch.int32, device=env.dev)}
    po = _abi.PlaceOutStruct()
    for k, _ in _abi.PlaceOutStruct._fields_:
        setattr(po, k, outs[k].data_ptr() if k in outs else None)

    def step_dev():
        ctx.allocate_dev(dab, d_g, d_s)
        ctx.place_dev(dpb, po)

    ms, launches = time_steps(ctx, env, step_dev, steps, warmup)

    # end to end: host buffers; the allocation result comes back to the host and goes out again as
    # the request list (two C-ABI calls, the way a caller without device pointers chains them)
    hab = ctx.pin_batch(ab)
    h_g = ctx.pinned((ab.jobs,), np.int32)
    h_s = ctx.pinned((R,), np.int32)
    hpb = ctx.pin_place_batch(pb)
    hout = ctx.pinned_place_out(R, pb.nodes, pb.requests)
    moved = [0, 0]

    def step_e2e():
        ctx.allocate(hab, h_g, h_s)
        a, b = ctx.copied_bytes
        hpb.req_workers = h_g  # the allocation result (page-locked) is the request list
        ctx.place(hpb, hout)
        moved[0], moved[1] = a, b

    e2e_ms = time_host_steps(env, step_e2e, e2e_steps, warmup=1)
    place_in = sum(getattr(pb, k).nbytes for k in ("node_off", "node_slots", "req_off", "req_workers", "prev_off", "prev_job", "prev_workers"))
    place_out = 4 * (2 * pb.nodes + 3 * cap + 3 * R) + 8 * R
    h2d, d2h = moved[0] + place_in, moved[1] + place_out

    nchk = min(R, 256)
    sub = ab.slice_rounds(0, nchk)
    g_ref, s_ref = oracle.allocate(sub, dense=True, threads=max(1, env.threads // env.world))
    nj, nn = int(ab.job_off[nchk]), nchk * wl["nodes"]
    from vodascheduler_b200.batch import PlaceBatch
    spb = PlaceBatch(pb.node_off[:nchk + 1], pb.node_slots[:nn], pb.req_off[:nchk + 1], g_ref,
                     pb.prev_off[:nn + 1], pb.prev_job[:int(pb.prev_off[nn])], pb.prev_workers[:int(pb.prev_off[nn])])
    ref = oracle.place(spb)
    okr = ref.status[:nchk] == 0
    dev_phys = outs["phys_of_logical"][:nn].cpu().numpy().reshape(nchk, -1)
    ok = (np.array_equal(d_g[:nj].cpu().numpy(), g_ref) and np.array_equal(d_s[:nchk].cpu().numpy(), s_ref) and
          np.array_equal(outs["status"][:nchk].cpu().numpy(), ref.status[:nchk]) and
          np.array_equal(dev_phys[okr], ref.phys_of_logical[:nn].reshape(nchk, -1)[okr]) and
          np.array_equal(outs["score"][:nchk].cpu().numpy()[okr], ref.score[:nchk][okr]) and
          np.array_equal(outs["cross"][:nchk].cpu().numpy()[okr], ref.cross[:nchk][okr]) and
          np.array_equal(hout.status, outs["status"].cpu().numpy()) and
          np.array_equal(hout.score[hout.status == 0], outs["score"].cpu().numpy()[hout.status == 0]))
    verified = env.all_ok(ok)
    if not verified:
        raise SystemExit(f"{name}: GPU result differs from the oracle; no timing reported")
    ms, e2e_ms = env.max_over_ranks(ms, e2e_ms)
    entry = None
    if env.rank == 0:
        entry = {
            "workload": wl["desc"], "value": R * env.world * steps / (ms / 1e3), "unit": "rounds/s",
            "ms_per_step": ms / steps, "steps": steps, "rounds_per_gpu": R,
            "e2e": {"value": R * env.world * e2e_steps / (e2e_ms / 1e3), "unit": "rounds/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                    "ms_per_step": e2e_ms / e2e_steps},
            "gpu_launches": int(launches),
            "roofline": roofline_entry(name, wl, ms / steps, R, "fill_kernel + place_kernel, one launch each per step"),
            "verified_vs_oracle": {"rounds_per_rank": nchk, "ranks": env.world, "arms": ["device", "host"], "ok": True,
                                   "nonterminating_rounds_in_sample": int((~okr).sum())},
        }
        if with_cpu:
            th = env.threads

            def rate(c):
                s_ = ab.slice_rounds(0, c)
                cn = c * wl["nodes"]
                t = time.perf_counter()
                g, _ = oracle.allocate(s_, dense=False, threads=th)
                p = PlaceBatch(pb.node_off[:c + 1], pb.node_slots[:cn], pb.req_off[:c + 1], g, pb.prev_off[:cn + 1],
                               pb.prev_job[:int(pb.prev_off[cn])], pb.prev_workers[:int(pb.prev_off[cn])])
                oracle.place(p, threads=th)
                return c / (time.perf_counter() - t)

            v, cnt = sized_rate(rate, R, th, 6.0)
            entry["cpu_baseline"] = {"value": v, "unit": "rounds/s", "cores": th, "kind": "port",
                                     "sample": f"first {cnt} rounds: structure-faithful allocation port + placement port (bestFit, score, munkres), {th} threads"}
        else:
            entry["cpu_baseline"] = None
    return entry


def run_place(name, wl, ctx, env, steps, warmup, e2e_steps, with_cpu):
    from vodascheduler_b200.batch import PlaceOut
    R = wl["rounds"]
    inputs = synth.place_batch(R, wl["nodes"], wl["slots"], wl["jobs"], seed=synth.SEED + env.rank)
    pout = ctx.pinned_place_out(R, inputs.nodes, inputs.requests)
    hin = ctx.pin_place_batch(inputs)  # page-locked, like every other end-to-end input
    e2e_ms = time_host_steps(env, lambda: ctx.place(hin, pout), e2e_steps)
    (e2e_ms,) = env.max_over_ranks(e2e_ms)
    if env.rank != 0:
        return None
    v = R * env.world * e2e_steps / (e2e_ms / 1e3)
    return {"workload": wl["desc"], "value": v, "unit": "rounds/s", "ms_per_step": e2e_ms / e2e_steps,
            "e2e": {"value": v, "unit": "rounds/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None},
            "roofline": None, "cpu_baseline": None}


def run_replay(name, wl, ctx, env, steps, warmup, e2e_steps, with_cpu):
    """configs[4]: every rank replays its own replicas (replicas only: a trace is sequential in
    simulated time) through voda_replay.  value = allocation+placement rounds / s of wall clock, host
    bookkeeping and every H2D/D2H hop included (so value == e2e here)."""
    import torch
    from vodascheduler_b200 import replay as rp
    reps = rp.make_replicas(wl["replicas"], wl["rounds"], wl["nodes"], wl["slots"], first=env.rank * wl["replicas"])
    from vodascheduler_b200 import Context
    K = wl.get("contexts", 4)  # replica groups, each on its own context (stream) and host thread
    ctxs = [ctx] + [Context(env.local) for _ in range(K - 1)]
    for c in ctxs:
        rp.replay(rp.make_replicas(2, 200, 8, 8), c)  # warm-up: kernels loaded, staging buffers grown
    env.barrier()
    l0 = sum(c.launches for c in ctxs)
    t0 = time.perf_counter()
    stats = rp.replay_parallel(reps, ctxs, host_threads=max(1, env.threads // (K * env.world)))
    dt = time.perf_counter() - t0
    launches = sum(c.launches for c in ctxs) - l0
    for c in ctxs[1:]:
        c.close()
    t = torch.tensor([dt, float(stats["rounds"])], dtype=torch.float64, device=env.dev)
    if env.world > 1:
        tm = t.clone()
        env.dist.all_reduce(tm, op=env.dist.ReduceOp.MAX)
        ts = t.clone()
        env.dist.all_reduce(ts, op=env.dist.ReduceOp.SUM)
        dt, rounds = float(tm[0]), float(ts[1])
    else:
        rounds = float(stats["rounds"])
    # parity: replica 0 of every rank against the CPU replay built from the oracles, on its first jobs
    from oracle import replay as orp
    nchk = min(wl["rounds"], 300)
    small = rp.make_replicas(1, nchk, wl["nodes"], wl["slots"], first=env.rank * wl["replicas"])
    got = rp.replay(small, ctx)["replicas"][0]
    tc = time.perf_counter()
    ref = orp.replay_one(small[0].trace, small[0].nodes, small[0].slots, small[0].rate_limit)
    cpu_dt = time.perf_counter() - tc
    ok = all(got[k] == ref[k] for k in ("done", "rounds", "migrated_workers", "migrated_launchers", "history", "actions"))
    if not env.all_ok(ok):
        raise SystemExit(f"{name}: GPU replay history differs from the CPU replay; no timing reported")
    if env.rank != 0:
        return None
    val = rounds / dt
    return {"workload": wl["desc"], "value": val, "unit": "rounds/s", "ms_per_step": 1e3 * dt, "steps": 1,
            "replicas_per_gpu": wl["replicas"], "jobs_per_replica": wl["rounds"], "contexts_per_gpu": K,
            "parallelism": "replicas only (a trace is sequential in time); the replicas of a GPU are dealt to "
                           f"{K} contexts, one host thread each",
            "e2e": {"value": val, "unit": "rounds/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
                    "note": "every tick copies its rounds to the device and the results back; host bookkeeping included"},
            "gpu_launches": int(launches), "roofline": None,
            "cpu_baseline": {"value": ref["rounds"] / cpu_dt, "unit": "rounds/s", "cores": 1, "kind": "port",
                             "sample": f"one replica of {nchk} jobs through oracle/replay.py (dense allocation port + placement "
                                       f"port + Python bookkeeping restatement), 1 thread"} if with_cpu else None,
            "verified_vs_oracle": {"replica_jobs": nchk, "ranks": env.world, "ok": True},
            "replay": {"ticks": stats["ticks"], "rounds_per_gpu": stats["rounds"], "replica0": stats["replicas"][0]}}


RUNNERS = {"alloc": run_alloc, "munkres": run_munkres, "chain": run_chain, "place": run_place, "replay": run_replay}


def run_reference(args, name, wl, rank):
    """--impl reference: the CPU port of the reference on the headline config, every host thread,
    a bounded sample per step."""
    if rank != 0:
        return
    if wl["kind"] != "alloc":
        raise SystemExit("--impl reference follows an allocation workload")
    threads = os.cpu_count() or 1
    inputs = make_alloc_inputs(wl, 0, rounds=min(wl["rounds"], max(64, 4 * threads)))
    R = inputs.rounds
    budget = max(1.0, 90.0 / (args.steps + args.warmup))
    rate, per_step = sized_rate(lambda c: oracle_alloc_rate(inputs, c, threads, False), R, threads, budget)
    for _ in range(args.warmup):
        oracle_alloc_rate(inputs, per_step, threads, False)
    t = time.perf_counter()
    for _ in range(args.steps):
        oracle_alloc_rate(inputs, per_step, threads, False)
    dt = time.perf_counter() - t
    val = per_step * args.steps / dt
    sample = (f"{per_step} rounds per step of the same workload, structure-faithful C++ port of the Go code "
              f"(string-keyed maps, a stable sort per pick), {threads} threads")
    print(json.dumps({
        "impl": "reference", "metric": "alloc+placement rounds/sec", "value": val, "unit": "rounds/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_of(name, wl, max(1, args.gpus)),
        "cpu_baseline": {"value": val, "unit": "rounds/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "rounds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference Go timing unavailable: no Go toolchain on this image; this is oracle/ (C++ restatement)",
    }))


def release(ctx):
    """Between workloads: page-locked host buffers and cached device memory go back (eight ranks on one
    box would otherwise hold ~7 GB of pinned memory each by the end of the run)."""
    import gc
    import torch
    ctx.sync()
    ctx.free_pinned()
    gc.collect()
    torch.cuda.empty_cache()


def config_of(name, wl, world):
    """The same dictionary in both arms (ours and --impl reference) for the same command line."""
    cfg = {"workload": wl["desc"], "name": name, "rounds_per_gpu": wl["rounds"],
           "parallelism": f"rounds sharded over {world} GPU(s), no collective on the data path"}
    if wl["kind"] == "alloc":
        # every array of the synthetic batch: 44 bytes of scalars + the W-entry float32 curve per job
        mb = (wl["rounds"] * wl["jobs"] * (44 + 4 * W) + 8 * wl["rounds"]) / 1e6
        cfg["l2"] = (f"inputs {mb:.0f} MB per GPU vs 126 MB L2: every step re-reads them from HBM" if mb > 130
                     else f"inputs {mb:.0f} MB per GPU (fit L2; outputs are rewritten every step)")
    return cfg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps of the headline (default: >= 1 s of timed region)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=HEADLINE, choices=sorted(WORKLOADS))
    ap.add_argument("--rounds", type=int, default=0, help="override rounds per GPU")
    ap.add_argument("--e2e-steps", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--only-headline", action="store_true", help="skip per_config")
    args = ap.parse_args()
    name = args.workload
    wl = dict(WORKLOADS[name])
    if args.rounds:
        wl["rounds"] = args.rounds
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        args.steps = args.steps or 5
        run_reference(args, name, wl, rank)
        return
    args.warmup = max(args.warmup, 3)
    steps = args.steps or wl.get("steps", 10)

    import torch
    import torch.distributed as dist
    from vodascheduler_b200 import Context

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    env = Env(rank, world, local, dev, dist if world > 1 else None)
    ctx = Context(local)
    with_cpu = (world == 1) and not args.no_cpu

    sampler = ClockSampler(local)
    t_wall0 = time.time()
    head = RUNNERS[wl["kind"]](name, wl, ctx, env, steps, args.warmup, args.e2e_steps or wl.get("e2e_steps", 4), with_cpu)
    t_wall1 = time.time()
    release(ctx)
    clocks = sampler.stop(t_wall0, t_wall1)

    per = {}
    if name == HEADLINE and not args.only_headline:
        for pn in PER_CONFIG:
            pw = dict(WORKLOADS[pn])
            try:
                per[pn] = RUNNERS[pw["kind"]](pn, pw, ctx, env, pw.get("steps", 10), 3, pw.get("e2e_steps", 4), with_cpu)
                release(ctx)
            except SystemExit:
                raise
            except Exception as e:  # a per-config failure must not hide the headline
                per[pn] = {"error": f"{type(e).__name__}: {e}"}

    if rank == 0:
        out = {
            "metric": "alloc+placement rounds/sec", "value": head["value"], "unit": head["unit"], "n_gpus": world,
            "steps": steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(name, wl, world),
            "e2e": head["e2e"], "gpu_launches": head.get("gpu_launches"),
            "roofline": head.get("roofline"), "cpu_baseline": head.get("cpu_baseline"), "clocks": clocks,
            "verified_vs_oracle": head.get("verified_vs_oracle"), "rank_summaries": head.get("rank_summaries"),
        }
        for k in ("single_round", "replay", "e2e_shared_curves", "variants"):
            if k in head:
                out[k] = head[k]
        if per:
            out["per_config"] = per
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
