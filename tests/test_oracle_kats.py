"""The CPU oracle against the known-answer tests of SURVEY.md Appendix B / C
(tests/golden/kats.json, hand-traced from the reference's Go text).  CPU only."""
import numpy as np
import pytest

from helpers import batch_from_jobs, load_kats, place_batch_from

KATS = load_kats()


@pytest.mark.parametrize("kat", KATS["alloc"], ids=lambda k: f"KAT-{k['kat']}-{k['algo']}")
@pytest.mark.parametrize("dense", [False, True], ids=["faithful", "dense"])
def test_alloc_kat(oracle, kat, dense):
    b = batch_from_jobs(kat["algo"], [(kat["jobs"], kat["total"])])
    gpus, st = oracle.allocate(b, dense=dense)
    assert st[0] == kat["status"]
    if kat["expect"] is not None:
        got = {j["name"]: int(g) for j, g in zip(kat["jobs"], gpus)}
        assert got == kat["expect"]


@pytest.mark.parametrize("kat", KATS["bestfit"], ids=lambda k: f"KAT-{k['kat']}")
def test_bestfit_kat(oracle, kat):
    out = oracle.place(place_batch_from([dict(nodes=kat["nodes"], requests=kat["requests"])]))
    assert out.status[0] == kat["status"]
    if kat["layout"] is not None:
        cnt = int(out.lay_cnt[0])
        got = [[int(out.lay_node[t]), int(out.lay_job[t]), int(out.lay_workers[t])] for t in range(cnt)]
        assert got == kat["layout"]
        assert out.free_slots[:len(kat["nodes"])].tolist() == kat["free"]
        # no previous layout -> all-zero score matrix -> identity binding (first zero per row)
        assert out.phys_of_logical[:len(kat["nodes"])].tolist() == list(range(len(kat["nodes"])))


@pytest.mark.parametrize("kat", KATS["munkres"], ids=lambda k: f"score{k['score']}-n{len(k['A'])}")
def test_munkres_kat(oracle, kat):
    A = np.array(kat["A"], np.int64)
    cols, st = oracle.munkres_max(A[None])
    assert st[0] == 0
    assert cols[0].tolist() == kat["cols"]
    assert int(A[np.arange(len(A)), cols[0]].sum()) == kat["score"]
