"""Python binding of the CPU oracle (oracle/libvoda_oracle.so).

TEST INFRASTRUCTURE, NOT PRODUCT: imported only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package never imports this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from vodascheduler_b200 import _abi
from vodascheduler_b200.batch import AllocBatch, PlaceBatch, PlaceOut

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libvoda_oracle.so")

_PROTOS = {
    "voda_oracle_allocate": (C.c_int, [C.POINTER(_abi.AllocBatchStruct), C.c_void_p, C.c_void_p, C.c_int32]),
    "voda_oracle_allocate_dense": (C.c_int, [C.POINTER(_abi.AllocBatchStruct), C.c_void_p, C.c_void_p, C.c_int32]),
    "voda_oracle_munkres_max": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]),
    "voda_oracle_munkres_counters": (None, [C.POINTER(C.c_int64)] * 3),
    "voda_oracle_place": (C.c_int, [C.POINTER(_abi.PlaceBatchStruct), C.POINTER(_abi.PlaceOutStruct), C.c_int32]),
    "voda_oracle_afsl_counters": (None, [C.POINTER(C.c_int64)] * 3),
}


def build(force=False):
    """Compile the oracle with the committed Makefile (g++ only; no reference sources)."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    srcs.append(os.path.join(_HERE, "..", "include", "voda_accel.h"))
    stale = force or not os.path.exists(_SO) or any(
        os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-B", "libvoda_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = _abi.bind(C.CDLL(_SO), _PROTOS)
    return _lib


def allocate(batch: AllocBatch, dense=False, threads=1):
    """-> (gpus[jobs] int32, status[rounds] int32)"""
    out = np.zeros(max(1, batch.jobs), np.int32)
    st = np.full(max(1, batch.rounds), -1, np.int32)
    s = batch.struct()
    fn = lib().voda_oracle_allocate_dense if dense else lib().voda_oracle_allocate
    rc = fn(C.byref(s), out.ctypes.data, st.ctypes.data, threads)
    if rc != 0:
        raise RuntimeError(f"oracle allocate failed rc={rc}")
    return out[:batch.jobs], st[:batch.rounds]


def munkres_max(A, threads=1):
    """A: int64 [batch, n, n] -> (col_of_row [batch, n] int32, status [batch])"""
    A = np.ascontiguousarray(A, dtype=np.int64)
    batch, n = (A.shape[0], A.shape[1]) if A.ndim == 3 else (1, A.shape[0])
    out = np.full(max(1, batch * n), -1, np.int32)
    st = np.full(max(1, batch), -1, np.int32)
    rc = lib().voda_oracle_munkres_max(n, batch, A.ctypes.data, out.ctypes.data, st.ctypes.data, threads)
    if rc != 0:
        raise RuntimeError(f"oracle munkres failed rc={rc}")
    return out[:batch * n].reshape(batch, n), st[:batch]


def munkres_counters():
    a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
    lib().voda_oracle_munkres_counters(C.byref(a), C.byref(b), C.byref(c))
    return {"step4": a.value, "step6": b.value, "step5": c.value}


def afsl_counters():
    a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
    lib().voda_oracle_afsl_counters(C.byref(a), C.byref(b), C.byref(c))
    return {"fold_steps": a.value, "champion_changes": b.value, "picks": c.value}


def place(batch: PlaceBatch, threads=1) -> PlaceOut:
    out = PlaceOut(batch.rounds, batch.nodes, batch.requests)
    s, o = batch.struct(), out.struct()
    rc = lib().voda_oracle_place(C.byref(s), C.byref(o), threads)
    if rc != 0:
        raise RuntimeError(f"oracle place failed rc={rc}")
    return out
