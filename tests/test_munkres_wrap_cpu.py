"""The reference's ComputeMunkresMax forms MaxInt64 - score and subtracts row minima in Go's wrapping
int64 arithmetic.  For a matrix whose rows each hold scores of one sign that is the ordinary
max-weight assignment; a row mixing negative and non-negative scores wraps and the six steps then run
on whatever comes out.  This test pins the C++ oracle (compiled with -fwrapv) to a second, pure-Python
restatement of the same steps that wraps explicitly, on mixed-sign matrices."""
import numpy as np

from oracle import oracle

M64 = (1 << 63) - 1


def wrap(x):
    return ((x + (1 << 63)) % (1 << 64)) - (1 << 63)


def munkres_max_py(A):
    n = len(A)
    a = [[wrap(M64 - int(A[i][j])) for j in range(n)] for i in range(n)]
    for i in range(n):  # step 1
        mn = min(a[i])
        a[i] = [wrap(v - mn) for v in a[i]]
    mark = [[0] * n for _ in range(n)]  # 1 star, 2 prime
    rc, cc = [False] * n, [False] * n
    for i in range(n):  # step 2
        for j in range(n):
            if a[i][j] == 0 and not rc[i] and not cc[j]:
                mark[i][j] = 1
                rc[i] = cc[j] = True
    rc, cc = [False] * n, [False] * n
    while True:
        for i in range(n):  # step 3
            for j in range(n):
                if mark[i][j] == 1:
                    cc[j] = True
        if sum(cc) >= n:
            break
        while True:  # step 4, with step 6 whenever no uncovered zero is left
            z = next(((i, j) for i in range(n) for j in range(n) if a[i][j] == 0 and not rc[i] and not cc[j]), None)
            if z is None:
                mn = min(a[i][j] for i in range(n) for j in range(n) if not rc[i] and not cc[j])
                for i in range(n):
                    for j in range(n):
                        if rc[i]:
                            a[i][j] = wrap(a[i][j] + mn)
                        if not cc[j]:
                            a[i][j] = wrap(a[i][j] - mn)
                continue
            i, j = z
            mark[i][j] = 2
            sc = next((c for c in range(n) if mark[i][c] == 1), -1)
            if sc >= 0:
                rc[i] = True
                cc[sc] = False
            else:
                break
        path = [z]  # step 5
        while True:
            r = next((r for r in range(n) if mark[r][path[-1][1]] == 1), -1)
            if r < 0:
                break
            path.append((r, path[-1][1]))
            c = next(c for c in range(n) if mark[r][c] == 2)
            path.append((r, c))
        for (r, c) in path:
            mark[r][c] = 0 if mark[r][c] == 1 else 1
        rc, cc = [False] * n, [False] * n
        for r in range(n):
            for c in range(n):
                if mark[r][c] == 2:
                    mark[r][c] = 0
    return [next(j for j in range(n) if mark[i][j] == 1) for i in range(n)]


def test_oracle_wraps_like_the_reference_on_mixed_sign_rows():
    oracle.build()
    rng = np.random.default_rng(11)
    mixed = 0
    for it in range(120):
        n = int(rng.integers(1, 9))
        hi = int(rng.choice([3, 9, 300, 70000, 1 << 40]))
        A = rng.integers(-hi, hi + 1, size=(n, n)).astype(np.int64)
        if it % 3 == 0:
            A = np.abs(A)  # one sign: the plain case
        mixed += int(((A < 0).any(axis=1) & (A >= 0).any(axis=1)).any())
        got, st = oracle.munkres_max(A[None], threads=1)
        assert got[0].tolist() == munkres_max_py(A.tolist()), A
    assert mixed >= 50
