"""ctypes binding of vodascheduler_b200/libvoda_accel.so (the C ABI of include/voda_accel.h).

Fails loudly: a missing library, a missing symbol or a missing CUDA device raises
EngineError.  There is no fallback path of any kind behind these calls.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _abi
from .batch import AllocBatch, PlaceBatch, PlaceOut

_HERE = os.path.dirname(os.path.abspath(__file__))
# VODA_ACCEL_LIB lets a profiling run point at an experimental build of the same library
_SO = os.environ.get("VODA_ACCEL_LIB") or os.path.join(_HERE, "libvoda_accel.so")


class EngineError(RuntimeError):
    """Infrastructure failure of the engine (not a per-round reference panic)."""


def library_path():
    return _SO


_lib = None


def load_library():
    """dlopen the built CUDA library and bind every symbol the header declares."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise EngineError(
                f"{_SO} is missing: build it with `python -m vodascheduler_b200.build` "
                "(nvcc, sm_100a).  This package has no CPU fallback.")
        try:
            _lib = _abi.bind(C.CDLL(_SO))
        except (OSError, AttributeError) as e:
            raise EngineError(f"cannot load {_SO}: {e}") from e
    return _lib


def status_str(code):
    return load_library().voda_status_str(int(code)).decode()


class Context:
    """voda_ctx: one CUDA device, its streams and scratch.  Not thread safe."""

    def __init__(self, device=0):
        self._lib = load_library()
        h = C.c_void_p()
        rc = self._lib.voda_ctx_create(int(device), C.byref(h))
        if rc != 0:
            raise EngineError(f"voda_ctx_create({device}) failed rc={rc}: "
                              f"{self._lib.voda_last_error(None).decode()}")
        self._h = h
        self.device = int(device)

    # -- plumbing ----------------------------------------------------------------------
    def close(self):
        """Destroys the context and frees the page-locked memory handed out by pinned() /
        pin_batch(): arrays returned by those die with the context."""
        if getattr(self, "_h", None):
            self._lib.voda_ctx_destroy(self._h)
            self._h = None
            for p in getattr(self, "_pins", []):
                self._lib.voda_host_free(p)
            self._pins = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise EngineError(f"{what} failed rc={rc}: {self._lib.voda_last_error(self._h).decode()}")

    @property
    def stream(self):
        """cudaStream_t (int) the *_dev entry points launch on."""
        return self._lib.voda_ctx_stream(self._h)

    def sync(self):
        self._check(self._lib.voda_ctx_sync(self._h), "voda_ctx_sync")

    @property
    def launches(self):
        return int(self._lib.voda_ctx_launch_count(self._h))

    @property
    def copied_bytes(self):
        """(host->device, device->host) bytes of the last host-pointer call."""
        a, b = C.c_int64(0), C.c_int64(0)
        self._lib.voda_ctx_copied_bytes(self._h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    # -- allocation --------------------------------------------------------------------
    def allocate(self, batch: AllocBatch, out_gpus=None, out_status=None):
        """Host buffers in, host buffers out (H2D, kernels, D2H inside the call)."""
        if out_gpus is None:
            out_gpus = np.zeros(max(1, batch.jobs), np.int32)
        if out_status is None:
            out_status = np.full(max(1, batch.rounds), -1, np.int32)
        s = batch.struct()
        rc = self._lib.voda_allocate(self._h, C.byref(s), out_gpus.ctypes.data, out_status.ctypes.data)
        self._check(rc, "voda_allocate")
        return out_gpus[:batch.jobs], out_status[:batch.rounds]

    def allocate_dev(self, dev_batch: AllocBatch, d_out_gpus, d_out_status):
        """Device-resident batch (torch tensors); asynchronous on self.stream."""
        s = dev_batch.struct()
        rc = self._lib.voda_allocate_dev(self._h, C.byref(s), d_out_gpus.data_ptr(), d_out_status.data_ptr())
        self._check(rc, "voda_allocate_dev")

    # -- munkres -----------------------------------------------------------------------
    def munkres_max(self, A):
        A = np.ascontiguousarray(A, dtype=np.int64)
        if A.ndim == 2:
            A = A[None]
        batch, n = A.shape[0], A.shape[1]
        out = np.full(max(1, batch * n), -1, np.int32)
        st = np.full(max(1, batch), -1, np.int32)
        rc = self._lib.voda_munkres_max(self._h, n, batch, A.ctypes.data, out.ctypes.data, st.ctypes.data)
        self._check(rc, "voda_munkres_max")
        return out[:batch * n].reshape(batch, n), st[:batch]

    def munkres_max_dev(self, n, batch, d_A, d_out, d_status):
        rc = self._lib.voda_munkres_max_dev(self._h, n, batch, d_A.data_ptr(), d_out.data_ptr(), d_status.data_ptr())
        self._check(rc, "voda_munkres_max_dev")

    # -- placement ---------------------------------------------------------------------
    def place(self, batch: PlaceBatch, out: PlaceOut | None = None) -> PlaceOut:
        if out is None:
            out = PlaceOut(batch.rounds, batch.nodes, batch.requests)
        s, o = batch.struct(), out.struct()
        self._check(self._lib.voda_place(self._h, C.byref(s), C.byref(o)), "voda_place")
        return out

    def place_dev(self, dev_batch: PlaceBatch, dev_out_struct):
        s = dev_batch.struct()
        self._check(self._lib.voda_place_dev(self._h, C.byref(s), C.byref(dev_out_struct)), "voda_place_dev")

    # -- speedup curves ----------------------------------------------------------------
    def build_curves(self, rec_off, workers, epoch_time_sec, width=34):
        """metrics_collector.py:131-167 for a batch of jobs (host arrays in, host arrays out):
        returns (epoch_time, speedup, efficiency) as [jobs, width] float64 and status [jobs]."""
        rec_off = np.ascontiguousarray(rec_off, np.int32)
        workers = np.ascontiguousarray(workers, np.int32)
        t = np.ascontiguousarray(epoch_time_sec, np.float64)
        jobs = max(0, rec_off.size - 1)
        et, sp, ef = (np.zeros((max(1, jobs), width), np.float64) for _ in range(3))
        st = np.full(max(1, jobs), -1, np.int32)
        s = _abi.CurveBatchStruct(jobs, width, rec_off.ctypes.data, workers.ctypes.data if workers.size else None,
                                  t.ctypes.data if t.size else None)
        self._check(self._lib.voda_build_curves(self._h, C.byref(s), et.ctypes.data, sp.ctypes.data, ef.ctypes.data,
                                                st.ctypes.data), "voda_build_curves")
        return et[:jobs], sp[:jobs], ef[:jobs], st[:jobs]

    # -- scheduler bookkeeping ---------------------------------------------------------
    TICK_FIELDS = ("last_update_ns", "running_ns", "gpu_ns", "total_ns", "waiting_ns", "last_running_ns",
                   "last_gpu_ns", "last_waiting_ns")

    def tick(self, state: dict, now_ns: int, tiresias: bool, thresholds_sec=(3600.0, float("inf")), promote_knob=8):
        """Scheduler.updateTimeMetrics for all jobs of `state` (dict of numpy arrays: status, num_gpu,
        priority int32; TICK_FIELDS int64 nanoseconds), updated in place; returns `changed` [jobs]."""
        n = int(state["status"].size)
        for k in ("status", "num_gpu", "priority"):
            state[k] = np.ascontiguousarray(state[k], np.int32)
        for k in self.TICK_FIELDS:
            state[k] = np.ascontiguousarray(state[k], np.int64)
        changed = np.zeros(max(1, n), np.int32)
        s = _abi.TickBatchStruct()
        s.jobs, s.tiresias, s.queue_num, s.promote_knob, s.now_ns = n, int(bool(tiresias)), len(thresholds_sec), promote_knob, now_ns
        for q, v in enumerate(thresholds_sec):
            s.thresholds_sec[q] = v
        for k in ("status", "num_gpu", "priority") + self.TICK_FIELDS:
            setattr(s, k, state[k].ctypes.data)
        s.changed = changed.ctypes.data
        self._check(self._lib.voda_tick(self._h, C.byref(s)), "voda_tick")
        return changed[:n]

    def compare_results(self, old_gpu, new_gpu, status):
        """Scheduler.compareResults: per job ACT_NONE / HALT / SCALE_IN / SCALE_OUT / START."""
        old_gpu, new_gpu, status = (np.ascontiguousarray(x, np.int32) for x in (old_gpu, new_gpu, status))
        n = int(old_gpu.size)
        act = np.zeros(max(1, n), np.int32)
        self._check(self._lib.voda_compare_results(self._h, n, old_gpu.ctypes.data, new_gpu.ctypes.data,
                                                   status.ctypes.data, act.ctypes.data), "voda_compare_results")
        return act[:n]

    # -- pinned host memory ------------------------------------------------------------
    def pinned(self, shape, dtype):
        """numpy array backed by voda_host_alloc (page-locked) memory."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape))
        p = self._lib.voda_host_alloc(max(1, n * dtype.itemsize))
        if not p:
            raise EngineError("voda_host_alloc failed")
        buf = (C.c_char * max(1, n * dtype.itemsize)).from_address(p)
        arr = np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)
        self._pins = getattr(self, "_pins", [])
        self._pins.append(p)
        return arr

    def free_pinned(self):
        """Give back every buffer handed out by pinned() / pin_batch(); the arrays must not be used
        afterwards."""
        for p in getattr(self, "_pins", []):
            self._lib.voda_host_free(p)
        self._pins = []

    def pin_place_batch(self, batch: PlaceBatch) -> PlaceBatch:
        """Copy of a placement batch whose arrays live in page-locked memory."""
        b = PlaceBatch.__new__(PlaceBatch)
        b.__dict__.update(batch.__dict__)
        for name, v in batch.__dict__.items():
            if isinstance(v, np.ndarray):
                pv = self.pinned(v.shape, v.dtype)
                pv[...] = v
                setattr(b, name, pv)
        return b

    def pinned_place_out(self, rounds, nodes, requests) -> PlaceOut:
        """PlaceOut whose arrays live in page-locked memory."""
        out = PlaceOut(rounds, nodes, requests)
        for name, v in list(out.__dict__.items()):
            if isinstance(v, np.ndarray):
                pv = self.pinned(v.shape, v.dtype)
                pv[...] = v
                setattr(out, name, pv)
        return out

    def pin_batch(self, batch: AllocBatch) -> AllocBatch:
        """Copy of `batch` whose arrays live in page-locked memory."""
        from .batch import _ALLOC_ARRAYS
        b = AllocBatch.__new__(AllocBatch)
        b.__dict__.update(batch.__dict__)
        for name, dt in _ALLOC_ARRAYS:
            v = getattr(batch, name)
            if v is not None:
                pv = self.pinned(v.shape, dt)
                pv[...] = v
                setattr(b, name, pv)
        so = getattr(batch, "speedup_off", None)
        if so is not None:
            pv = self.pinned(so.shape, np.int32)
            pv[...] = so
            b.speedup_off = pv
        return b
