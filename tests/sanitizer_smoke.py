"""Small end-to-end pass over every kernel, meant to be run under
`compute-sanitizer --tool memcheck python tests/sanitizer_smoke.py` on a GPU box
(not collected by pytest).  Sizes are tiny on purpose: the sanitizer is ~100x slower."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vodascheduler_b200 import ALGO_NAMES, Context, synth  # noqa: E402
from oracle import oracle  # noqa: E402

ctx = Context(0)
for algo in ALGO_NAMES:
    for b in (synth.jobs_batch(algo, 6, 70, 150, ragged=True), synth.quirk_batch(algo, 12, 20, 30)):
        g, s = ctx.allocate(b)
        g0, s0 = oracle.allocate(b, dense=True)
        assert np.array_equal(g, g0) and np.array_equal(s, s0), algo
b = synth.jobs_batch("FfDLOptimizer", 3, 40, 300)
b.max_proc = (b.max_proc * 20).astype(np.int32)  # uint16 SOL path + beyond-table g
assert np.array_equal(ctx.allocate(b)[0], oracle.allocate(b, dense=True)[0])
for n in (5, 33, 128):
    A = synth.matrices(3, n, "placement")
    assert np.array_equal(ctx.munkres_max(A)[0], oracle.munkres_max(A)[0])
A = synth.matrices(2, 40, "uniform", hi=200) * 100000
assert np.array_equal(ctx.munkres_max(A)[0], oracle.munkres_max(A)[0])
pb = synth.place_batch(3, 20, 8, 30)
out, ref = ctx.place(pb), oracle.place(pb)
assert np.array_equal(out.phys_of_logical, ref.phys_of_logical)
ctx.close()
print("sanitizer smoke ok")
