"""The curve-builder oracle against golden vectors produced by the reference's own code
(tests/golden/make_curve_kats.py runs metrics_collector.py:131-167): bit-exact float64."""
import json
import os

import numpy as np

from oracle import curves

KATS = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "curve_kats.json")))["cases"]


def _dense(d, width):
    out = np.zeros(width)
    for k, v in d.items():
        out[int(k)] = float.fromhex(v)
    return out


def test_oracle_matches_the_reference_bit_for_bit():
    assert len(KATS) >= 10
    for c in KATS:
        t = [float.fromhex(x) for x in c["epoch_time_sec"]]
        et, sp, ef = curves.build_curves(c["workers"], t)
        w = len(et)
        for got, name in ((et, "epoch_time"), (sp, "speedup"), (ef, "efficiency")):
            want = _dense(c[name], w)
            assert got.tobytes() == want.tobytes(), (name, len(t), np.nonzero(got != want)[0][:4])


def test_pairwise_sum_is_numpys():
    rng = np.random.default_rng(5)
    for n in (0, 1, 7, 8, 9, 15, 16, 127, 128, 129, 255, 256, 1000, 4099):
        a = rng.normal(size=n) * 10.0 ** rng.integers(-3, 6, size=n)
        assert curves.pairwise_sum(a) == float(np.sum(a)), n
