"""GPU: the same hand-derived semantic cases through the C ABI -- against the expected answers AND the oracle."""
import numpy as np
import pytest

import semantic_cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", semantic_cases.cases(), ids=lambda c: c[0])
def test_gpu_hand_derived(engine, oracle, gen, case):
    name, pm, X, P, rep, expected = case
    m = engine.compile(pm)
    pb = gen.pack_present(P)
    r = engine.score(m, X, pb, rep)
    o = oracle.OracleModel(pm).predict(X, pb, rep)
    assert np.array_equal(r["status"], o["status"]), (name, r["status"], o["status"])
    ok = o["status"] == 0
    if name in semantic_cases.TOLERANT:
        np.testing.assert_allclose(r["score"][ok], o["score"][ok], rtol=1e-9)
    else:
        assert np.array_equal(r["score"][ok], o["score"][ok]), (name, r["score"], o["score"])
    for i, (st, score, aux) in enumerate(expected):
        assert r["status"][i] == st, (name, i, r["status"])
        if st == 0 and name in semantic_cases.TOLERANT:
            assert abs(r["score"][i] - score) <= 1e-9 * abs(score)
        elif st == 0:
            assert r["score"][i] == score
        if aux is not None:
            assert r["aux"][i] == aux, (name, i, r["aux"], o["aux"])
