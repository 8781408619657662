"""The speedup-curve builder kernel through the C ABI: bit-exact float64 against the golden
vectors produced by the reference's own code (metrics_collector.py:131-167 run through pandas,
tests/golden/make_curve_kats.py) and against the pinned oracle on random batches."""
import json
import os

import numpy as np
import pytest

from oracle import curves

pytestmark = pytest.mark.gpu
KATS = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "curve_kats.json")))["cases"]


def _dense(d, width):
    out = np.zeros(width)
    for k, v in d.items():
        out[int(k)] = float.fromhex(v)
    return out


def test_golden_vectors_of_the_reference(ctx):
    off = np.cumsum([0] + [len(c["workers"]) for c in KATS]).astype(np.int32)
    w = np.concatenate([np.asarray(c["workers"], np.int32) for c in KATS])
    t = np.concatenate([np.array([float.fromhex(x) for x in c["epoch_time_sec"]]) for c in KATS])
    et, sp, ef, st = ctx.build_curves(off, w, t, width=34)
    assert (st == 0).all()
    for j, c in enumerate(KATS):
        for got, name in ((et[j], "epoch_time"), (sp[j], "speedup"), (ef[j], "efficiency")):
            want = _dense(c[name], 34)
            assert got.tobytes() == want.tobytes(), (j, name, np.nonzero(got != want)[0][:4])


def test_random_batches_against_the_oracle(ctx):
    rng = np.random.default_rng(20261107)
    counts = np.concatenate([[0, 1, 7, 8, 9, 127, 128, 129, 1500], rng.integers(0, 400, size=300)])
    off = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    w = rng.choice([1, 1, 1, 2, 2, 4, 8, 16, 32, 33], size=off[-1]).astype(np.int32)
    t = np.exp(rng.normal(3.0, 1.0, size=off[-1])) / w
    et, sp, ef, st = ctx.build_curves(off, w, t, width=34)
    assert (st == 0).all()
    for j in range(len(counts)):
        a, b = off[j], off[j + 1]
        ret, rsp, ref = curves.build_curves(w[a:b], t[a:b], 34)
        assert et[j].tobytes() == ret.tobytes() and sp[j].tobytes() == rsp.tobytes() and ef[j].tobytes() == ref.tobytes(), j


def test_out_of_range_worker_counts_and_empty_batch(ctx):
    off = np.array([0, 3, 5], np.int32)
    w = np.array([1, 40, 2, 1, -1], np.int32)
    t = np.array([10.0, 1.0, 6.0, 8.0, 3.0])
    et, sp, ef, st = ctx.build_curves(off, w, t, width=34)
    assert st.tolist() == [8, 8]
    assert et[0, 1] == 10.0 and et[0, 2] == 6.0 and sp[0, 2] == 10.0 / 6.0 and ef[0, 2] == (10.0 / 6.0) / 2.0
    et, sp, ef, st = ctx.build_curves(np.zeros(1, np.int32), np.zeros(0, np.int32), np.zeros(0), width=34)
    assert et.shape == (0, 34) and st.size == 0
