"""Writes tests/golden/kats.json — the known-answer tests of SURVEY.md Appendix B and C.

Provenance: the expected values are HAND-TRACED from the reference's Go source text
(pkg/algorithm/*.go, pkg/placement/placement_manager.go:415-487) and, for munkres, from the
six-step order restated in oracle/oracle_munkres.cpp.  They were not produced by running the
reference (no Go toolchain on this image) and not by running this repo's code: this script
only serialises the literals below.  If a KAT ever disagrees with the Go text, the Go text wins.
"""
import json
import os

C = [0, 1, 1.9, 2.7, 3.4, 4.0, 4.5, 4.9, 5.2, 5.4]
D = [0, 1, 1.5, 1.8, 2.0, 2.1, 2.15, 2.17, 2.18, 2.18]
LIN = list(range(34))
OK, LESS_MIN, EXC_MAX, NONTERM = 0, 2, 3, 7


def job(name, submit=0, first=0, prio=0, np_=1, mn=1, mx=1, rem=0.0, curve=LIN):
    return dict(name=name, submit=submit, first=first, priority=prio, num_proc=np_, min=mn,
                max=mx, remaining=rem, speedup=curve)


fifo3 = [job("A", submit=1, mn=1, mx=4, rem=30), job("B", submit=2, mn=2, mx=2, rem=10),
         job("C", submit=3, mn=2, mx=8, rem=20)]
srjf3 = [dict(j) for j in fifo3]
srjf3b = [dict(j) for j in fifo3]
srjf3b[1]["max"] = 3
fifo4 = [job("A", submit=1, mn=3, mx=4), job("B", submit=2, mn=4, mx=4), job("C", submit=3, mn=1, mx=8)]
tir = [job("A", first=5, prio=0, np_=2, mn=1, mx=4, curve=C), job("B", first=3, prio=1, np_=2, mn=1, mx=4, curve=D),
       job("C", first=4, prio=0, np_=4, mn=2, mx=6, curve=C), job("D", first=1, prio=1, np_=1, mn=1, mx=2, curve=LIN)]
ffdl3 = [job("A", submit=1, mn=1, mx=4, curve=C), job("B", submit=2, mn=1, mx=4, curve=D),
         job("C", submit=3, mn=1, mx=6, curve=C)]
ffdl2 = [job("A", submit=1, mn=1, mx=4), job("B", submit=2, mn=1, mx=4)]
# MinNumProc == 0 (S[0] = 0.5 so that the first gain S[0]/0 is +Inf, not NaN).  Hand trace, total 3:
# basic portion A(np 5) does not fit, B gets 1 (free 2, gain S[2]-S[1] = 1); greedy: A (+Inf) takes its
# minimum 0 -> r 0, free 2, gain S[1]-S[0] = 0.5; B (1) -> 2 = max, removed, free 1; A (0.5) takes 0
# again -> r, free, gain all unchanged: the Go loop repeats this pick forever -> guard status 7.
E = [0.5, 1, 1.9, 2.7, 3.4]
tir_min0 = [job("A", first=1, prio=0, np_=5, mn=0, mx=4, curve=E), job("B", first=2, prio=0, np_=1, mn=1, mx=2)]
afsl3 = [job("A", submit=1, mn=1, mx=4, rem=100, curve=C), job("B", submit=2, mn=1, mx=4, rem=50, curve=D),
         job("C", submit=3, mn=1, mx=6, rem=200, curve=C)]

alloc = [
    dict(kat="1", algo="ElasticFIFO", jobs=fifo3, total=8, status=OK, expect=dict(A=3, B=2, C=3)),
    dict(kat="2", algo="ElasticSRJF", jobs=srjf3, total=8, status=EXC_MAX, expect=None),
    dict(kat="3", algo="ElasticSRJF", jobs=srjf3b, total=8, status=OK, expect=dict(A=2, B=3, C=3)),
    dict(kat="4a", algo="ElasticFIFO", jobs=fifo4, total=5, status=OK, expect=dict(A=4, B=0, C=1)),
    dict(kat="4b", algo="ElasticFIFO", jobs=fifo4, total=6, status=LESS_MIN, expect=None),
    dict(kat="5", algo="ElasticTiresias", jobs=tir, total=8, status=OK, expect=dict(A=2, B=1, C=4, D=1)),
    dict(kat="6", algo="ElasticTiresias", jobs=tir, total=12, status=OK, expect=dict(A=4, B=2, C=4, D=2)),
    dict(kat="7", algo="ElasticTiresias", jobs=tir, total=5, status=OK, expect=dict(A=0, B=0, C=4, D=1)),
    dict(kat="15", algo="ElasticTiresias", jobs=tir_min0, total=3, status=NONTERM, expect=None),
    dict(kat="8", algo="FfDLOptimizer", jobs=ffdl3, total=8, status=OK, expect=dict(A=4, B=1, C=3)),
    dict(kat="9", algo="FfDLOptimizer", jobs=ffdl3, total=2, status=OK, expect=dict(A=1, B=1, C=0)),
    dict(kat="10", algo="FfDLOptimizer", jobs=ffdl2, total=6, status=OK, expect=dict(A=4, B=2)),
]
for total, exp in [(1, dict(A=0, B=1, C=0)), (2, dict(A=1, B=1, C=0)), (3, dict(A=1, B=1, C=1)),
                   (5, dict(A=2, B=2, C=1)), (8, dict(A=3, B=2, C=3)), (14, dict(A=4, B=4, C=6)),
                   (20, dict(A=4, B=4, C=6))]:
    alloc.append(dict(kat=f"11/{total}", algo="AFS-L", jobs=afsl3, total=total, status=OK, expect=exp))

bestfit = [
    dict(kat="12", nodes=[4, 4], requests=[4, 3, 1], status=OK,
         layout=[[0, 0, 4], [1, 1, 3], [1, 2, 1]], free=[0, 0]),
    dict(kat="13", nodes=[8, 8], requests=[6, 6, 4], status=OK,
         layout=[[0, 0, 6], [1, 1, 6], [0, 2, 2], [1, 2, 4]], free=[0, -2]),
    dict(kat="14", nodes=[8, 8], requests=[10, 3], status=NONTERM, layout=None, free=None),
]

munkres = [
    dict(A=[[0, 0], [0, 0]], cols=[0, 1], score=0),
    dict(A=[[4, 0, 0], [0, 4, 0], [4, 4, 0]], cols=[2, 1, 0], score=8),
    dict(A=[[2, 2, 0], [2, 2, 0], [0, 0, 3]], cols=[0, 1, 2], score=7),
    dict(A=[[1, 2, 3], [2, 4, 6], [3, 6, 9]], cols=[0, 1, 2], score=14),
    dict(A=[[8, 0, 0, 0], [0, 0, 8, 0], [0, 4, 0, 4], [0, 4, 0, 4]], cols=[0, 2, 1, 3], score=24),
    dict(A=[[5, 5, 5], [5, 5, 5], [5, 5, 5]], cols=[0, 1, 2], score=15),
]

out = dict(provenance=__doc__.strip(), alloc=alloc, bestfit=bestfit, munkres=munkres)
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "kats.json"), "w") as f:
    json.dump(out, f, indent=1)
