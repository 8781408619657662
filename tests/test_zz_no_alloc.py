"""voda_ffdl_allocate_static — the entry point for hosts that must not allocate GPU memory
(their HBM is owned by another allocator, or they cannot take the device-wide stall of a
cudaMalloc).  It keeps the FfDL kernel's SOL scratch in a fixed array inside the library
instead of a buffer grown on demand, so it must give exactly the answers of the ordinary
entry point, call after call, while sharing that one array between calls.

Collected last on purpose (file name): a failure here cannot hide other tests in a -x run.
Works alone in a fresh process and after the rest of the suite (it creates its own context
and shares no state with other tests).
"""
import ctypes as C

import numpy as np
import pytest

from vodascheduler_b200 import synth

pytestmark = pytest.mark.gpu


def _run_both(ctx, lib, batch):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)  # the static entry point runs on the current device
    db = batch.to_device(dev)
    ref_g = torch.zeros(batch.jobs, dtype=torch.int32, device=dev)
    ref_s = torch.full((batch.rounds,), -1, dtype=torch.int32, device=dev)
    ctx.allocate_dev(db, ref_g, ref_s)
    ctx.sync()
    got_g = torch.full((batch.jobs,), -7, dtype=torch.int32, device=dev)
    got_s = torch.full((batch.rounds,), -7, dtype=torch.int32, device=dev)
    err = C.create_string_buffer(256)
    s = db.struct()
    rc = lib.voda_ffdl_allocate_static(None, C.byref(s), got_g.data_ptr(), got_s.data_ptr(), err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    return ref_g.cpu().numpy(), ref_s.cpu().numpy(), got_g.cpu().numpy(), got_s.cpu().numpy()


def test_static_scratch_matches_allocating_entry_point_twice(oracle):
    """Calls in a row on different batches (each finds the previous one's bytes in the shared
    array; the third uses a much smaller layout, so stale rows of the larger ones lie under its
    slabs) all equal voda_allocate_dev bit for bit — allocations and statuses — and, so that a
    defect common to both GPU entry points cannot pass, the CPU oracle as well."""
    from vodascheduler_b200 import Context
    ctx = Context(0)
    try:
        lib = ctx._lib
        first = synth.jobs_batch("FfDLOptimizer", rounds=300, jobs_per_round=96, total_gpu=320, ragged=True)
        second = synth.quirk_batch("FfDLOptimizer", rounds=500, jobs_per_round=48, total_gpu=200, seed=99)
        third = synth.jobs_batch("FfDLOptimizer", rounds=40, jobs_per_round=7, total_gpu=12, ragged=True, seed=5)
        for batch in (first, second, third):
            ref_g, ref_s, got_g, got_s = _run_both(ctx, lib, batch)
            assert np.array_equal(got_s, ref_s)
            assert np.array_equal(got_g, ref_g)
            cpu_g, cpu_s = oracle.allocate(batch, dense=True, threads=8)
            assert np.array_equal(got_s, cpu_s)
            assert np.array_equal(got_g, cpu_g)
            if batch is second:
                assert (ref_s != 0).any()  # the quirk batch exercises the panic statuses too
    finally:
        ctx.close()


def test_static_scratch_rejects_oversize_shape():
    """A shape whose single-round scratch exceeds the fixed 32 MiB is refused with a message,
    before anything is launched."""
    import torch
    from vodascheduler_b200 import Context, _abi
    ctx = Context(0)
    try:
        torch.cuda.set_device(0)
        b = synth.jobs_batch("FfDLOptimizer", rounds=2, jobs_per_round=8, total_gpu=16)
        db = b.to_device(torch.device("cuda", 0))
        db.max_jobs, db.max_total_gpu = 2048, 8192  # 2048 rows * 1024 tiles * 32 B = 64 MiB per round
        db.max_max_proc = 0
        out = torch.zeros(b.jobs, dtype=torch.int32, device="cuda")
        st = torch.zeros(b.rounds, dtype=torch.int32, device="cuda")
        err = C.create_string_buffer(256)
        s = db.struct()
        rc = ctx._lib.voda_ffdl_allocate_static(None, C.byref(s), out.data_ptr(), st.data_ptr(), err, len(err))
        assert rc == _abi.E_TOO_LARGE
        assert b"exceeds the static scratch" in err.value
        # with the bound on the GPUs one job can take declared only every 4th row is kept: 16 MiB, it fits
        db.max_max_proc = int(b.max_proc.max())
        assert db.max_max_proc <= 32
        s = db.struct()
        rc = ctx._lib.voda_ffdl_allocate_static(None, C.byref(s), out.data_ptr(), st.data_ptr(), err, len(err))
        assert rc == 0, err.value
        torch.cuda.synchronize()
        g_ref, s_ref = ctx.allocate(b)
        assert np.array_equal(out.cpu().numpy(), g_ref) and np.array_equal(st.cpu().numpy(), s_ref)
    finally:
        ctx.close()
