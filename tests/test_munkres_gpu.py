"""GPU parity of the batched munkres kernel (voda_munkres_max, through the C ABI) against the
CPU restatement: identical starred cells (tie-breaking included), and the optimum of scipy.
Parity with the real heyfey/munkres module is unpinned (its source is not in the reference)."""
import numpy as np
import pytest
from scipy.optimize import linear_sum_assignment

from helpers import load_kats
from vodascheduler_b200 import synth

pytestmark = pytest.mark.gpu
KATS = load_kats()


@pytest.mark.parametrize("kat", KATS["munkres"], ids=lambda k: f"score{k['score']}-n{len(k['A'])}")
def test_kat(ctx, kat):
    A = np.array(kat["A"], np.int64)
    cols, st = ctx.munkres_max(A)
    assert st[0] == 0 and cols[0].tolist() == kat["cols"]


@pytest.mark.parametrize("n", [1, 2, 3, 7, 31, 32, 33, 64, 100, 128])
@pytest.mark.parametrize("kind,hi", [("uniform", 1), ("uniform", 8), ("uniform", 999), ("placement", 8)])
def test_matches_oracle(ctx, oracle, n, kind, hi):
    A = synth.matrices(24, n, kind=kind, hi=hi, seed=1000 + n)
    ref, _ = oracle.munkres_max(A, threads=8)
    got, st = ctx.munkres_max(A)
    assert (st == 0).all()
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("scale,offset", [(1, 0), (300, 0), (70000, 0), (1 << 33, 0), (1, 1 << 40), (1, -(1 << 40))])
def test_cell_width_ladder(ctx, oracle, scale, offset):
    """uint8 / uint16 / int32 / int64 shared-memory cell types are chosen per matrix by range;
    a batch mixing them must still match the oracle everywhere."""
    A = synth.matrices(16, 48, kind="uniform", hi=200, seed=77) * scale + offset
    A[::2] = synth.matrices(8, 48, kind="uniform", hi=9, seed=78)  # narrow ones in between
    ref, _ = oracle.munkres_max(A, threads=8)
    got, st = ctx.munkres_max(A)
    assert (st == 0).all() and np.array_equal(got, ref)


def test_optimal_and_permutation_at_full_size(ctx):
    """BASELINE config 4 shape (128 x 128): checked through properties that need no oracle."""
    A = synth.matrices(512, 128, kind="placement")
    got, st = ctx.munkres_max(A)
    assert (st == 0).all()
    idx = np.arange(128)
    for k in range(0, 512, 37):
        assert sorted(got[k].tolist()) == list(range(128))
        ri, ci = linear_sum_assignment(A[k], maximize=True)
        assert A[k][idx, got[k]].sum() == A[k][ri, ci].sum()
    # invariance: adding a constant to a whole row does not move the assignment
    B = A + (np.arange(128, dtype=np.int64) * 3)[None, :, None]
    got2, _ = ctx.munkres_max(B)
    assert np.array_equal(got, got2)


def test_large_batch_chunked(ctx, oracle):
    A = synth.matrices(6000, 24, kind="uniform", hi=5)
    ref, _ = oracle.munkres_max(A, threads=8)
    got, st = ctx.munkres_max(A)
    assert np.array_equal(got, ref) and (st == 0).all()


def test_rejects_oversize(ctx):
    from vodascheduler_b200 import EngineError
    with pytest.raises(EngineError):
        ctx.munkres_max(np.zeros((1, 129, 129), np.int64))


@pytest.mark.parametrize("n", [1, 2, 3, 5, 8, 16, 31, 32, 63, 64, 100, 128])
def test_mixed_sign_scores_wrap_like_the_reference(ctx, oracle, n):
    """MaxInt64 - score wraps in the reference when a row holds negative and non-negative scores
    (tests/test_munkres_wrap_cpu.py pins what then comes out); such matrices go to the widest kernel,
    which computes in the same wrapping arithmetic.  Mixed batches: plain matrices in between."""
    rng = np.random.default_rng(300 + n)
    for hi in (1, 8, 255, 999, 70000, 1 << 40):
        A = synth.matrices(20, n, kind="uniform", hi=hi, seed=500 + n)
        A[::3] -= int(rng.integers(1, hi + 1))   # every third matrix gets negative entries
        A[1::6] = -A[1::6] - 1                    # and some are negative throughout
        got, st = ctx.munkres_max(A)
        ref, _ = oracle.munkres_max(A, threads=8)
        assert (st == 0).all()
        assert np.array_equal(got, ref), (n, hi)
