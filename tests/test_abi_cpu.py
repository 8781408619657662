"""CPU-only checks of the C ABI library: it loads, exports every symbol the header
declares, answers the pure host helpers, and REFUSES to compute without a GPU."""
import ctypes as C
import os
import re

import pytest

from vodascheduler_b200 import _abi, build as vbuild


@pytest.fixture(scope="module")
def lib():
    so = vbuild.build()
    return _abi.bind(C.CDLL(so))


def test_every_declared_symbol_is_exported(lib):
    hdr = open(os.path.join(os.path.dirname(vbuild.HERE), "include", "voda_accel.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(voda_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/voda_accel.h but not exported"
    assert declared == set(_abi.PROTOTYPES), "ctypes prototypes out of sync with the header"


def test_policy_names_mirror_factory(lib):
    # pkg/algorithm/types.go:26-47
    for i, name in enumerate(_abi.ALGO_NAMES):
        assert lib.voda_algo_from_name(name.encode()) == i
        assert lib.voda_algo_name(i).decode() == name
    assert lib.voda_algo_from_name(b"Optimus") == -1  # errors.New("Not found")
    need = {n: bool(lib.voda_algo_need_job_info(i)) for i, n in enumerate(_abi.ALGO_NAMES)}
    assert need == {"FIFO": False, "ElasticFIFO": False, "SRJF": True, "ElasticSRJF": True,
                    "Tiresias": False, "ElasticTiresias": True, "FfDLOptimizer": True, "AFS-L": True}


def test_status_strings_are_the_reference_panics(lib):
    assert lib.voda_status_str(1).decode() == "Invalid GPU allocations: can't be negative"
    assert lib.voda_status_str(2).decode() == "Invalid GPU allocations: less than job min gpu"
    assert lib.voda_status_str(3).decode() == "Invalid GPU allocations: exceeded job max gpu"
    assert lib.voda_status_str(4).decode() == "Invalid GPU allocations: exceeded total GPUs"
    assert lib.voda_status_str(5).decode() == "infeasible"
    assert lib.voda_status_str(6).decode() == "Not found"


def test_no_cpu_fallback(lib):
    """Without a CUDA device the context cannot be created: the product never computes on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.voda_ctx_create(0, C.byref(h))
    assert rc == _abi.E_CUDA and not h.value
    assert b"no CUDA device" in lib.voda_last_error(None)


def test_product_package_never_uses_oracle():
    """Nothing under vodascheduler_b200/ imports, links or dlopens anything from oracle/."""
    root = os.path.dirname(vbuild.HERE)
    pat = re.compile(r"(^\s*(from|import)\s+oracle\b|libvoda_oracle|voda_oracle_|oracle/|#include\s+\".*oracle)", re.M)
    for dirpath, _, files in os.walk(os.path.join(root, "vodascheduler_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert not pat.search(text), f"{f} reaches into the oracle"


def test_ctypes_mirrors_have_the_compiled_layouts(lib, tmp_path):
    """Every struct of include/voda_accel.h that _abi.py mirrors by hand: same size as the library
    was compiled with, and — through a C program compiled against the header here — the same offset
    for every field."""
    import subprocess
    mirrors = [("voda_alloc_batch", _abi.AllocBatchStruct), ("voda_place_batch", _abi.PlaceBatchStruct),
               ("voda_place_out", _abi.PlaceOutStruct), ("voda_curve_batch", _abi.CurveBatchStruct),
               ("voda_tick_batch", _abi.TickBatchStruct), ("voda_replay_batch", _abi.ReplayBatchStruct),
               ("voda_replay_out", _abi.ReplayOutStruct)]
    for which, (cname, st) in enumerate(mirrors):
        assert lib.voda_struct_size(which) == C.sizeof(st), cname
    assert lib.voda_struct_size(99) == 0
    inc = os.path.join(os.path.dirname(vbuild.HERE), "include")
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "voda_accel.h"', 'int main(void) {']
    for cname, st in mirrors:
        for fname, _ in st._fields_:
            lines.append(f'  printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "offsets.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "offsets"
    subprocess.run(["gcc", "-I", inc, str(src), "-o", str(exe)], check=True)
    got = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, st in mirrors:
        for fname, _ in st._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(st, fname).offset, f"{cname}.{fname}"
