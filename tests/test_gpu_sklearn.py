"""GPU: cross-checks against arithmetic this repository did not write.  scikit-learn fits a random forest (predict_proba =
average of the leaves' class distributions), a multinomial logistic regression and three MLPs (relu / tanh / logistic);
test-side emitters (tests/hand_pmml.py) write them as PMML; the CUDA path scores them through the C ABI and must give
scikit-learn's own labels (except where scikit-learn's top-2 probabilities differ by less than 1e-12) and, for the families
whose kernels report probabilities, scikit-learn's probabilities within 1e-6 relative.  The reference holds no fixture for
these families; this is the independent check for configs 3 and both halves of config 5."""
import numpy as np
import pytest

import hand_pmml

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def suite():
    pytest.importorskip("sklearn")
    return hand_pmml.sklearn_suite()


@pytest.mark.parametrize("k", range(5))
def test_gpu_matches_scikit_learn(engine, oracle, suite, k):
    name, pm, X, proba = suite[k]
    m = engine.compile(pm)
    has_probs = not name.startswith("rf")         # forest kernels report the label only
    r = engine.score(m, X, n_probs=proba.shape[1] if has_probs else 0)
    hand_pmml.check_against_sklearn(name, r, proba, has_probs)
    # and the same answers as the oracle, bit for bit on labels
    o = oracle.OracleModel(pm).predict(X, n_probs=proba.shape[1], nthreads=4)
    assert np.array_equal(r["status"], o["status"]) and np.array_equal(r["score"], o["score"]), name
    # a batch large enough for the throughput kernels (tiles over all SMs) must agree with the small one
    reps = 40
    Xb = np.tile(X, (reps, 1))
    rb = engine.score(m, Xb)
    assert np.array_equal(rb["score"], np.tile(r["score"], reps)) and (rb["status"] == 0).all(), name
