"""CPU: the oracle against hand-derived answers for tree / mining / preparation / clustering semantics
(families the reference's own tests do not pin)."""
import numpy as np
import pytest

import semantic_cases


@pytest.mark.parametrize("case", semantic_cases.cases(), ids=lambda c: c[0])
def test_oracle_hand_derived(oracle, gen, case):
    name, pm, X, P, rep, expected = case
    m = oracle.OracleModel(pm)
    r = m.predict(X, gen.pack_present(P), rep)
    for i, (st, score, aux) in enumerate(expected):
        assert r["status"][i] == st, (name, i, r["status"], r["score"])
        if st == 0 and name in semantic_cases.TOLERANT:
            assert abs(r["score"][i] - score) <= semantic_cases.TOLERANT[name] * abs(score), (name, i, r["score"])
        elif st == 0:
            assert r["score"][i] == score, (name, i, r["score"])
        else:
            assert np.isnan(r["score"][i])
        if aux is not None:
            assert r["aux"][i] == aux, (name, i, r["aux"])
