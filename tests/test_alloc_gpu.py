"""GPU parity of the allocation kernels against the CPU oracle, through the C ABI
(voda_allocate: host buffers in/out).  Integer outputs: the bar is bit-exact equality of every
per-job GPU count and every per-round status."""
import numpy as np
import pytest

from helpers import batch_from_jobs, load_kats
from vodascheduler_b200 import ALGO_NAMES, synth

pytestmark = pytest.mark.gpu

KATS = load_kats()
# policies whose kernels exist; extended as kernels land
from vodascheduler_b200.supported import SUPPORTED_ALGOS  # noqa: E402

ALGOS = [a for a in ALGO_NAMES if a in SUPPORTED_ALGOS]


def _same(ctx, oracle, b, dense=True, threads=8):
    g_ref, s_ref = oracle.allocate(b, dense=dense, threads=threads)
    g, s = ctx.allocate(b)
    bad = np.nonzero(s != s_ref)[0]
    assert bad.size == 0, f"status differs in rounds {bad[:5]}: gpu {s[bad[:5]]} ref {s_ref[bad[:5]]}"
    badj = np.nonzero(g != g_ref)[0]
    assert badj.size == 0, f"{badj.size} job allocations differ, first at {badj[:5]}: gpu {g[badj[:5]]} ref {g_ref[badj[:5]]}"
    return g, s


@pytest.mark.parametrize("kat", [k for k in KATS["alloc"] if k["algo"] in SUPPORTED_ALGOS],
                         ids=lambda k: f"KAT-{k['kat']}-{k['algo']}")
def test_kat(ctx, kat):
    b = batch_from_jobs(kat["algo"], [(kat["jobs"], kat["total"])])
    g, st = ctx.allocate(b)
    assert st[0] == kat["status"]
    if kat["expect"] is not None:
        assert {j["name"]: int(x) for j, x in zip(kat["jobs"], g)} == kat["expect"]


@pytest.mark.parametrize("algo", ALGOS)
def test_random_vs_faithful_oracle(ctx, oracle, algo):
    """Small enough for the structure-faithful (string-map) restatement."""
    b = synth.jobs_batch(algo, rounds=64, jobs_per_round=48, total_gpu=120, ragged=True)
    _same(ctx, oracle, b, dense=False, threads=4)


@pytest.mark.parametrize("algo", ALGOS)
def test_quirks(ctx, oracle, algo):
    """min == max, min > max, Priority 2, zero totals, flat curves, heavy ties: statuses too."""
    b = synth.quirk_batch(algo, rounds=600, jobs_per_round=40, total_gpu=64)
    g, s = _same(ctx, oracle, b)
    assert (s != 0).any() and (s == 0).any()


@pytest.mark.parametrize("algo", ALGOS)
def test_min_zero(ctx, oracle, algo):
    """MinNumProc == 0 (what a JSON request without min_num_proc decodes to): the reference's
    ElasticTiresias loop never ends on such a job; the kernel must return, with status 7."""
    b = synth.min_zero_batch(algo, rounds=400, jobs_per_round=40, total_gpu=64)
    g, s = _same(ctx, oracle, b)
    if algo == "ElasticTiresias":
        assert (s == 7).any()


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("jobs,total", [(1, 1), (3, 0), (7, 3), (33, 31), (64, 16), (129, 257), (300, 1000)])
def test_shapes(ctx, oracle, algo, jobs, total):
    b = synth.jobs_batch(algo, rounds=24, jobs_per_round=jobs, total_gpu=total)
    _same(ctx, oracle, b)


@pytest.mark.parametrize("algo", ALGOS)
def test_empty_and_ragged(ctx, oracle, algo):
    b = synth.jobs_batch(algo, rounds=40, jobs_per_round=20, total_gpu=50, ragged=True)
    # splice in empty rounds
    off = b.job_off.copy()
    off[5] = off[4]
    off = np.maximum.accumulate(off)
    cnt = np.diff(off)
    keep = np.concatenate([np.arange(b.job_off[r], b.job_off[r] + cnt[r]) for r in range(b.rounds)]).astype(np.int64)
    from vodascheduler_b200.batch import _ALLOC_ARRAYS
    for name, _ in _ALLOC_ARRAYS[2:]:
        v = getattr(b, name)
        if name == "speedup":
            v = v.reshape(-1, b.speedup_stride)[keep].reshape(-1)
        else:
            v = v[keep]
        setattr(b, name, np.ascontiguousarray(v))
    b.job_off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    _same(ctx, oracle, b)
    # zero rounds
    z = synth.jobs_batch(algo, rounds=0, jobs_per_round=4, total_gpu=8)
    g, s = ctx.allocate(z)
    assert g.size == 0 and s.size == 0


def test_ffdl_full_size_rounds(ctx, oracle):
    """BASELINE config 2 shape (256 jobs x 1024 GPUs), a sample checked against the oracle."""
    b = synth.jobs_batch("FfDLOptimizer", rounds=48, jobs_per_round=256, total_gpu=1024)
    g, s = _same(ctx, oracle, b)
    assert (s == 0).all()
    per_round = g.reshape(48, 256).sum(1)
    assert (per_round <= 1024).all() and (per_round > 900).all()


def test_afsl_full_size_rounds(ctx, oracle):
    """BASELINE config 3 shape (1024 jobs x 4096 GPUs): a sample against the oracle, plus the
    size-independent properties (every GPU handed out, nobody above max)."""
    b = synth.jobs_batch("AFS-L", rounds=24, jobs_per_round=1024, total_gpu=4096)
    g, s = _same(ctx, oracle, b)
    assert (s == 0).all()
    per_round = g.reshape(24, 1024).sum(1)
    assert (per_round == 4096).all()
    assert (g <= b.max_proc).all() and (g >= 0).all()


def test_tiresias_c1_shape(ctx, oracle):
    """BASELINE config 0 shape: 64 jobs on 16 GPUs, many rounds, against the string-map oracle."""
    b = synth.jobs_batch("ElasticTiresias", rounds=400, jobs_per_round=64, total_gpu=16)
    g, s = _same(ctx, oracle, b, dense=False, threads=8)
    assert (g.reshape(400, 64).sum(1) <= 16).all()


@pytest.mark.parametrize("algo", ["Tiresias", "ElasticTiresias"])
def test_tiresias_times_outside_the_packed_key(ctx, oracle, algo):
    """The queue sort packs (queue, FirstStartTime) into 64 bits when every key of the round fits
    (seconds in [0, 2^32-1), or the MaxTime sentinel); rounds with anything else — times before
    1970, beyond 2106, the sentinel's seconds with other nanoseconds — take the general compare.
    Both kinds of round in one batch, ties included."""
    b = synth.jobs_batch(algo, rounds=96, jobs_per_round=64, total_gpu=16)
    sec = b.first_sec.copy()
    ns = b.first_nsec.copy()
    rnd = np.repeat(np.arange(b.rounds), 64)
    j = np.arange(b.jobs) % 64
    sec[(rnd % 4 == 1) & (j % 7 == 0)] = -5                       # before the epoch
    sec[(rnd % 4 == 2) & (j % 5 == 0)] = (1 << 32) + 17           # beyond the 32-bit field
    ns[(rnd % 4 == 3) & (j % 3 == 0) & (sec == synth.MAXTIME_SEC)] = 5  # sentinel seconds, other nanoseconds
    sec[(rnd % 8 == 0) & (j % 2 == 0)] = 1_790_000_000            # packable rounds with heavy ties
    ns[(rnd % 8 == 0) & (j % 2 == 0)] = 7
    b.first_sec, b.first_nsec = sec, ns.astype(b.first_nsec.dtype)
    _same(ctx, oracle, b)


@pytest.mark.parametrize("algo", ["FIFO", "ElasticFIFO", "FfDLOptimizer", "AFS-L"])
def test_submit_times_outside_the_packed_key(ctx, oracle, algo):
    """SubmitTime sorts run on 64-bit packed keys when the round's times lie in [0, 2^32) s and on
    the general (seconds, nanoseconds, index) compare otherwise; both kinds of round, with ties."""
    b = synth.jobs_batch(algo, rounds=64, jobs_per_round=50, total_gpu=70, ragged=True)
    sec = b.submit_sec.copy()
    ns = np.zeros(b.jobs, np.int32) if b.submit_nsec is None else b.submit_nsec.copy()
    rnd = np.repeat(np.arange(b.rounds), np.diff(b.job_off))
    j = np.arange(b.jobs)
    sec[(rnd % 4 == 1) & (j % 6 == 0)] = -1_000_000              # before the epoch
    sec[(rnd % 4 == 2) & (j % 5 == 0)] = (1 << 33) + 5           # beyond the 32-bit field
    sec[(rnd % 4 == 3) & (j % 2 == 0)] = 1_790_000_123           # packable, heavy ties
    ns[(rnd % 4 == 3)] = (j[(rnd % 4 == 3)] % 3) * 7
    b.submit_sec, b.submit_nsec = sec, ns
    _same(ctx, oracle, b)


def test_esrjf_large_rounds(ctx, oracle):
    """The policy of BASELINE config 4 at 1024 jobs per round (bitonic sort + epoch fill)."""
    for algo in ("ElasticSRJF", "ElasticFIFO"):
        b = synth.jobs_batch(algo, rounds=64, jobs_per_round=1024, total_gpu=3000)
        _same(ctx, oracle, b)


def test_ffdl_wide_max_and_table_edge(ctx, oracle):
    """max_proc beyond the dense table (map miss => S = 0) and far beyond 32 (every row of P
    stored, row-by-row backtrack)."""
    b = synth.jobs_batch("FfDLOptimizer", rounds=12, jobs_per_round=24, total_gpu=700)
    b.max_proc = (b.max_proc.astype(np.int64) * np.where(np.arange(b.jobs) % 5 == 0, 40, 2)).astype(np.int32)
    _same(ctx, oracle, b)
    b.total_gpu[:] = 90
    _same(ctx, oracle, b)


@pytest.mark.parametrize("gcap", [31, 32, 33, 34])
@pytest.mark.parametrize("jobs", [1, 2, 3, 4, 5, 7, 8, 9, 63, 64, 65])
def test_ffdl_checkpoint_boundaries(ctx, oracle, gcap, jobs):
    """The backtrack walks a round in segments of 4 rows from every 4th stored row of P when no job
    can take more than 32 GPUs, and row by row from every row otherwise: both sides of that
    switch, every remainder of the row count modulo 4, budgets from 0 to far beyond the demand."""
    b = synth.jobs_batch("FfDLOptimizer", rounds=24, jobs_per_round=jobs, total_gpu=96, ragged=False)
    b.max_proc = np.minimum(b.max_proc * 4, gcap).astype(np.int32)
    b.max_proc[::3] = gcap  # the widest job sits in every round
    b.total_gpu[:] = np.resize(np.array([0, 1, 2, 3, 5, 31, 32, 33, 64, 97, 130, 1000], np.int32), b.rounds)
    _same(ctx, oracle, b)


def test_ffdl_ties_and_flat_tables(ctx, oracle):
    """Flat, zero, negative and NaN-free degenerate speedup tables make most candidates tie: the
    strict '>' keeps the smallest g in the reference, and so must the re-evaluated arg-max."""
    b = synth.jobs_batch("FfDLOptimizer", rounds=40, jobs_per_round=37, total_gpu=150)
    S = b.speedup.reshape(b.jobs, b.speedup_stride).copy()
    kind = np.arange(b.jobs) % 5
    S[kind == 0, 1:] = 1.0                      # flat
    S[kind == 1, :] = 0.0                       # all zero (also +0 vs -0 below)
    S[kind == 2, 1::2] = -0.0
    S[kind == 3, 1:] = -np.abs(S[kind == 3, 1:])  # negative: -10000 init never loses to them
    S[kind == 4, 1:] = np.round(S[kind == 4, 1:])  # coarse values: many equal sums
    b.speedup = S.reshape(-1).astype(np.float32)
    _same(ctx, oracle, b)
    b.total_gpu[:] = 17
    _same(ctx, oracle, b)


@pytest.mark.parametrize("algo", ALGOS)
def test_chunked_pipeline_equals_single(ctx, oracle, algo):
    """A batch large enough to be cut into several pipeline chunks (>= 16 MiB and >= 12 rounds
    per SM each) gives the same answer as the oracle computes in one piece."""
    rounds = 9000 if algo in ("AFS-L", "ElasticTiresias", "FfDLOptimizer") else 30000
    b = synth.jobs_batch(algo, rounds=rounds, jobs_per_round=48, total_gpu=40, ragged=True)
    _same(ctx, oracle, b)


def test_declared_bounds_are_enforced(ctx):
    b = synth.jobs_batch(ALGOS[0], rounds=4, jobs_per_round=16, total_gpu=32)
    b.max_jobs = 8
    g, s = ctx.allocate(b)
    assert (s == 8).all() and (g == 0).all()


def _ffdl_dev(c, b, declared):
    """voda_allocate_dev on a device copy of b with max_max_proc = declared (0: not declared)."""
    import torch
    dev = torch.device("cuda", 0)
    db = b.to_device(dev)
    db.max_max_proc = declared
    g = torch.full((max(1, b.jobs),), -7, dtype=torch.int32, device=dev)
    s = torch.full((b.rounds,), -1, dtype=torch.int32, device=dev)
    c.allocate_dev(db, g, s)
    c.sync()
    return g.cpu().numpy()[:b.jobs], s.cpu().numpy()


@pytest.mark.parametrize("jobs", [1, 4, 5, 8, 9, 61, 256])
def test_ffdl_declared_max_max_proc(ctx, oracle, jobs):
    """The device entry point with the bound on the GPUs one job can take declared (scratch holds
    the checkpoint rows only), not declared, and declared above 32 (every row stored): one answer."""
    b = synth.jobs_batch("FfDLOptimizer", rounds=40, jobs_per_round=jobs, total_gpu=300, ragged=True)
    b.total_gpu[:] = np.resize(np.array([0, 1, 7, 64, 300, 299, 33, 150], np.int32), b.rounds)
    g_ref, s_ref = oracle.allocate(b, dense=True, threads=8)
    top = int(b.max_proc.max())
    assert top <= 32
    for declared in (0, top, 32, 33, 1000):
        g, s = _ffdl_dev(ctx, b, declared)
        assert np.array_equal(s, s_ref), declared
        assert np.array_equal(g, g_ref), declared


def test_ffdl_declared_max_max_proc_is_enforced(ctx, oracle):
    """A round in which a job can take more GPUs than declared reports VODA_ST_BOUND (its jobs get
    0); min(MaxNumProc, total_gpu) is what counts, and the other rounds are unaffected."""
    b = synth.jobs_batch("FfDLOptimizer", rounds=30, jobs_per_round=20, total_gpu=200, ragged=False)
    b.max_proc = np.minimum(b.max_proc, 8).astype(np.int32)
    first = b.job_off[:-1]
    for r, m in ((3, 9), (11, 33), (12, 40), (20, 9), (29, 1 << 20)):
        b.max_proc[first[r] + 2] = m
    b.total_gpu[5] = 6   # no job can take more than 6 here
    b.max_proc[first[5] + 1] = 50  # ... whatever MaxNumProc says
    b.total_gpu[3] = 8   # and round 3 is inside a bound of 8 after all
    g_ref, s_ref = oracle.allocate(b, dense=True, threads=8)
    for declared in (8, 32):
        g, s = _ffdl_dev(ctx, b, declared)
        # which rounds break the bound: some job with min(max_proc, total) > declared
        top = np.array([min(int(b.max_proc[b.job_off[r]:b.job_off[r + 1]].max()), int(b.total_gpu[r]))
                        for r in range(b.rounds)])
        over = top > declared
        assert np.nonzero(over)[0].tolist() == ([11, 12, 20, 29] if declared == 8 else [11, 12, 29])
        assert (s[over] == 8).all()
        assert np.array_equal(s[~over], s_ref[~over])
        per_round = np.repeat(over, np.diff(b.job_off))
        assert (g[per_round] == 0).all()
        assert np.array_equal(g[~per_round], g_ref[~per_round])


def test_ffdl_declared_bound_shrinks_the_scratch(oracle):
    """What the declaration buys: a context that ran a declared batch holds about a quarter of
    the device memory of one that ran the same batch undeclared."""
    import torch
    from vodascheduler_b200 import Context
    b = synth.jobs_batch("FfDLOptimizer", rounds=1500, jobs_per_round=256, total_gpu=1024, ragged=False)
    used = {}
    for declared in (0, 32):  # torch's own (cached) tensors are counted in the first, larger, figure
        torch.cuda.synchronize()
        c = Context(0)
        try:
            free0 = torch.cuda.mem_get_info(0)[0]
            g, s = _ffdl_dev(c, b, declared)
            used[declared] = free0 - torch.cuda.mem_get_info(0)[0]
            if not declared:
                first = (g.copy(), s.copy())
            else:
                assert np.array_equal(g, first[0]) and np.array_equal(s, first[1])
        finally:
            c.close()
    assert used[0] > (1 << 30)
    assert used[32] < 0.35 * used[0], used


def test_ffdl_packed_rows_opt_in(oracle, monkeypatch):
    """VODA_PACK_ROWS=1: the host path ships only the speedup entries FfDL can read (packed
    rows + offsets); results must not change."""
    from vodascheduler_b200 import Context
    monkeypatch.setenv("VODA_PACK_ROWS", "1")
    monkeypatch.setenv("VODA_HOST_THREADS", "3")
    c = Context(0)
    try:
        for b in (synth.jobs_batch("FfDLOptimizer", rounds=4000, jobs_per_round=40, total_gpu=90, ragged=True),
                  synth.quirk_batch("FfDLOptimizer", rounds=500, jobs_per_round=30, total_gpu=50)):
            g_ref, s_ref = oracle.allocate(b, dense=True, threads=8)
            g, s = c.allocate(b)
            assert np.array_equal(s, s_ref) and np.array_equal(g, g_ref)
    finally:
        c.close()


def _tie_heavy(b, seed=7):
    """10 % of the jobs repeat their predecessor's SubmitTime / FirstStartTime / remaining time and a
    fifth of the curves are flat from some count on: the stable-sort fall-backs (input order) and the
    equal-gain / equal-length rules decide most picks."""
    rng = np.random.default_rng(seed)
    n = b.jobs
    dup = np.nonzero(rng.random(n) < 0.10)[0]
    dup = dup[dup > 0]
    for name in ("submit_sec", "first_sec", "first_nsec", "remaining_sec"):
        v = getattr(b, name).copy()
        v[dup] = v[dup - 1]
        setattr(b, name, v)
    S = b.speedup.reshape(n, b.speedup_stride).copy()
    flat = np.nonzero(rng.random(n) < 0.20)[0]
    cap = rng.integers(1, 6, size=flat.size)
    for i, c in zip(flat, cap):
        S[i, c:] = S[i, c]
    b.speedup = np.ascontiguousarray(S.reshape(-1))
    return b


@pytest.mark.parametrize("ties", [False, True], ids=["plain", "tie-heavy"])
def test_ffdl_c2_shape_vs_structure_faithful_oracle(ctx, oracle, ties):
    """BASELINE configs[1] shape (256 jobs x 1024 GPUs) against the STRUCTURE-FAITHFUL restatement
    (string-keyed maps, the reference's own loop nest), not the dense one that shares the kernels'
    derived forms."""
    b = synth.jobs_batch("FfDLOptimizer", rounds=64, jobs_per_round=256, total_gpu=1024)
    if ties:
        b = _tie_heavy(b)
    _same(ctx, oracle, b, dense=False, threads=8)


@pytest.mark.parametrize("ties", [False, True], ids=["plain", "tie-heavy"])
def test_afsl_c3_shape_vs_structure_faithful_oracle(ctx, oracle, ties):
    """BASELINE configs[2] shape (1024 jobs x 4096 GPUs) against the structure-faithful restatement:
    the literal left fold of afsl.go:72-92 with a map lookup per step — the check that the rank
    columns, the first-beat passes and the champion cache of the kernel hold at full size, with heavy
    ties."""
    b = synth.jobs_batch("AFS-L", rounds=64, jobs_per_round=1024, total_gpu=4096)
    if ties:
        b = _tie_heavy(b)
    g, s = _same(ctx, oracle, b, dense=False, threads=8)
    assert (s == 0).all()


@pytest.mark.parametrize("ties", [False, True], ids=["plain", "tie-heavy"])
def test_tiresias_1024_jobs_vs_structure_faithful_oracle(ctx, oracle, ties):
    """The stamp arg-max against the literal double stable sort per pick at 1024 jobs."""
    b = synth.jobs_batch("ElasticTiresias", rounds=32, jobs_per_round=1024, total_gpu=4096)
    if ties:
        b = _tie_heavy(b)
    _same(ctx, oracle, b, dense=False, threads=8)


def test_afsl_more_than_1024_jobs(ctx, oracle):
    """Rounds above 1024 jobs take the four-warp kernel; 1024 and below the one-warp kernel."""
    for jobs, total in ((1500, 3000), (1025, 700)):
        b = synth.jobs_batch("AFS-L", rounds=6, jobs_per_round=jobs, total_gpu=total)
        _same(ctx, oracle, b)


@pytest.mark.parametrize("jobs", [1, 31, 32, 33, 64, 100, 128, 200, 256, 500, 512, 513, 1000, 1024])
def test_afsl_slot_widths(ctx, oracle, jobs):
    """Every register-slot width of the one-warp kernel (1 .. 32 slots per lane), ragged rounds, a
    budget that removes most jobs (compaction of the live list) and one that removes none."""
    for total in (3 * jobs, max(1, jobs // 3)):
        b = synth.jobs_batch("AFS-L", rounds=12, jobs_per_round=jobs, total_gpu=total, ragged=True)
        _same(ctx, oracle, b)
    q = synth.quirk_batch("AFS-L", rounds=40, jobs_per_round=jobs, total_gpu=2 * jobs)
    _same(ctx, oracle, q)


def test_afsl_large_worker_counts(ctx, oracle):
    """Champions holding more workers than the rank table has columns (on-the-fly division), beyond
    the table width (map miss: rem / 0) included."""
    b = synth.jobs_batch("AFS-L", rounds=16, jobs_per_round=96, total_gpu=2500)
    b.max_proc = (b.max_proc.astype(np.int64) * 3).astype(np.int32)
    _same(ctx, oracle, b)
    b = synth.jobs_batch("AFS-L", rounds=8, jobs_per_round=700, total_gpu=9000)
    b.max_proc = np.where(np.arange(b.jobs) % 7 == 0, 60, b.max_proc).astype(np.int32)
    _same(ctx, oracle, b)


@pytest.mark.parametrize("jobs,total", [(40, 90), (300, 700), (1024, 2500)])
def test_afsl_degenerate_tables_and_times(ctx, oracle, jobs, total):
    """Quotients that are +Inf, NaN or negative: all-zero speedup rows (rem / 0 = +Inf, 0 / 0 = NaN),
    zero and negative remaining times, negative and flat table entries, -0.0.  The rank columns must
    order them exactly as the float32 compares of afsl.go:77-90 do (NaN loses every '>=')."""
    b = synth.jobs_batch("AFS-L", rounds=10, jobs_per_round=jobs, total_gpu=total)
    S = b.speedup.reshape(b.jobs, b.speedup_stride).copy()
    rem = b.remaining_sec.copy()
    kind = np.arange(b.jobs) % 11
    S[kind == 0, :] = 0.0                       # every length +Inf (or NaN with rem 0), every ratio NaN
    S[kind == 1, 1::2] = -0.0
    S[kind == 2, 1:] = -np.abs(S[kind == 2, 1:])  # negative lengths
    S[kind == 3, 1:] = 1.0                      # flat: zero gains
    S[kind == 4, 3:] = 0.0                      # the table ends early
    rem[kind == 5] = 0.0
    rem[kind == 6] = -7.5
    rem[(kind == 0) & (np.arange(b.jobs) % 2 == 0)] = 0.0   # 0 / 0
    rem[kind == 7] = np.float32(np.inf)
    b.speedup = S.reshape(-1).astype(np.float32)
    b.remaining_sec = rem.astype(np.float32)
    _same(ctx, oracle, b)
    _same(ctx, oracle, b, dense=False, threads=8)


@pytest.mark.parametrize("algo", ALGOS)
def test_shared_curve_pool_equals_dense(ctx, oracle, algo):
    """voda_alloc_batch.speedup_off: what-if rounds over the same jobs point into ONE pool of curves.
    Host path (the pool crosses PCIe once, chunks stage offsets) and device path give exactly what the
    dense table gives, which is what the oracle gives."""
    import torch
    b = synth.jobs_batch(algo, rounds=700, jobs_per_round=40, total_gpu=64)
    b.total_gpu = (np.arange(700) % 97).astype(np.int32)
    pooled, dense = synth.share_curves(b)
    g_ref, s_ref = oracle.allocate(dense, dense=True, threads=8)
    g, s = ctx.allocate(pooled)
    assert np.array_equal(s, s_ref) and np.array_equal(g, g_ref)
    gp, sp = oracle.allocate(pooled, dense=True, threads=8)     # the oracle reads the pool the same way
    assert np.array_equal(sp, s_ref) and np.array_equal(gp, g_ref)
    dev = torch.device("cuda", 0)
    db = pooled.to_device(dev)
    d_g = torch.zeros(pooled.jobs, dtype=torch.int32, device=dev)
    d_s = torch.full((pooled.rounds,), -1, dtype=torch.int32, device=dev)
    ctx.allocate_dev(db, d_g, d_s)
    ctx.sync()
    assert np.array_equal(d_s.cpu().numpy(), s_ref) and np.array_equal(d_g.cpu().numpy(), g_ref)


def test_shared_curve_pool_large_chunks(ctx, oracle):
    """Enough rounds for several pipeline chunks: every chunk reads the pool copied at the start."""
    for algo, jobs in (("FfDLOptimizer", 48), ("AFS-L", 64), ("ElasticTiresias", 64)):
        b = synth.jobs_batch(algo, rounds=30000, jobs_per_round=jobs, total_gpu=40)
        b.total_gpu = (1 + np.arange(30000) % 80).astype(np.int32)
        pooled, dense = synth.share_curves(b)
        g_ref, s_ref = oracle.allocate(dense, dense=True, threads=8)
        g, s = ctx.allocate(pooled)
        assert np.array_equal(s, s_ref) and np.array_equal(g, g_ref)
        h2d, d2h = ctx.copied_bytes
        assert h2d < dense.nbytes() / 3   # the curves crossed once


@pytest.mark.parametrize("algo", ALGOS)
def test_small_batch_paths_agree(oracle, monkeypatch, algo):
    """A batch of a few KB is read by the kernel straight from the page-locked staging block (no
    copies); a larger small one is gathered into one copy each way; VODA_DIRECT_BYTES=0 turns the
    first into the second.  All three give the oracle's answer, status codes included."""
    from vodascheduler_b200 import Context
    batches = [synth.jobs_batch(algo, rounds=1, jobs_per_round=64, total_gpu=16),
               synth.quirk_batch(algo, rounds=3, jobs_per_round=12, total_gpu=20),
               synth.jobs_batch(algo, rounds=40, jobs_per_round=30, total_gpu=64, ragged=True)]
    for direct in ("16384", "0"):
        monkeypatch.setenv("VODA_DIRECT_BYTES", direct)
        c = Context(0)
        try:
            for b in batches:
                g_ref, s_ref = oracle.allocate(b, dense=True, threads=4)
                for _ in range(2):  # the second call reuses the staging block
                    g, s = c.allocate(b)
                    assert np.array_equal(s, s_ref) and np.array_equal(g, g_ref), (algo, direct)
        finally:
            c.close()
