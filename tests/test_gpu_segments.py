"""GPU: Segment predicates other than True/False and the selectFirst / median / max segmentation methods
(JPMML MiningModelEvaluator behind S/api/PmmlModel.scala:159-160; SURVEY.md App. A "MiningModel / Segmentation").
Seeded ensembles from the C4 / C5 generators get predicates spliced onto a share of their segments; scores, statuses
and per-segment node ids must equal the oracle's bit for bit, through every forest path the model qualifies for."""
import re

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def splice_predicates(pm: bytes, F: int, seed: int, share=0.4, keep_first=True, keep_last=False) -> bytes:
    """Replace the <True/> of a random share of the segments with a simple / compound / set predicate on the input fields."""
    rng = np.random.default_rng(seed)
    segs = [mm.start() for mm in re.finditer(rb'<Segment id="\d+"[^>]*><True/>', pm)]
    out, prev = [], 0
    for k, _ in enumerate(segs):
        mm = re.compile(rb'<Segment id="\d+"[^>]*><True/>').match(pm, segs[k])
        out.append(pm[prev:mm.end() - len(b"<True/>")]); prev = mm.end()
        if (keep_first and k == 0) or (keep_last and k == len(segs) - 1) or rng.random() >= share:
            out.append(b"<True/>"); continue
        f, g = (int(v) for v in rng.integers(0, F, size=2)); v = float(np.float32(rng.normal()))
        kind = int(rng.integers(0, 5))
        if kind == 0: p = f'<SimplePredicate field="x{f}" operator="greaterThan" value="{v!r}"/>'
        elif kind == 1: p = f'<SimplePredicate field="x{f}" operator="lessOrEqual" value="{v!r}"/>'
        elif kind == 2: p = f'<SimplePredicate field="x{f}" operator="isNotMissing"/>'
        elif kind == 3: p = (f'<CompoundPredicate booleanOperator="or"><SimplePredicate field="x{f}" operator="lessThan" value="{v!r}"/>'
                             f'<SimplePredicate field="x{g}" operator="isMissing"/></CompoundPredicate>')
        else: p = (f'<CompoundPredicate booleanOperator="and"><SimplePredicate field="x{f}" operator="greaterOrEqual" value="{v!r}"/>'
                   f'<SimplePredicate field="x{g}" operator="lessThan" value="0.5"/></CompoundPredicate>')
        out.append(p.encode())
    out.append(pm[prev:])
    return b"".join(out)


def same(r, o, what, aux=True):
    assert np.array_equal(r["status"], o["status"]), what
    assert np.array_equal(r["score"].view(np.int64), o["score"].view(np.int64)), what
    if aux:
        assert np.array_equal(r["aux"], o["aux"]), what


@pytest.mark.parametrize("method", ["sum", "weightedSum", "average", "weightedAverage", "median", "max", "selectFirst"])
def test_regression_segments_with_predicates(engine, oracle, gen, monkeypatch, method):
    T, depth, F = 41, 5, 12
    pm, _ = gen.gbt_pmml(T=T, depth=depth, F=F, seed=311, method=method, segment_weights=method.startswith("weighted"))
    pm = splice_predicates(pm, F, seed=5, share=0.9 if method == "selectFirst" else 0.4, keep_first=method != "selectFirst", keep_last=True)
    assert len(re.findall(rb"<True/>\s*<TreeModel", pm)) < T
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    # sum / weightedSum with an always-active segment: a skipped segment adds +0.0, so the ensemble keeps the on-chip forms
    assert ("format=complete" in m.describe) == (method in ("sum", "weightedSum")), m.describe
    X, pres, pb = gen.events(20000, F, 312, 0.15)
    o = om.predict(X, pb, node_ids=True, nthreads=8)
    assert len(np.unique(o["score"][o["status"] == 0])) > 100
    monkeypatch.setenv("FJP_FOREST_COMPACT", "0")
    r = engine.score(m, X, pb, node_ids=True)
    same(r, o, method)
    assert np.array_equal(r["node_ids"], o["node_ids"]), method
    assert (o["node_ids"] == -1).any()                       # some segments did sit out
    r2 = engine.score(m, X, pb)                              # the path without node ids (one-kernel walk where it applies)
    same(r2, o, method)
    if method in ("sum", "weightedSum"):
        monkeypatch.setenv("FJP_FOREST_COMPACT", "2")        # rank-key pre-pass + compact walk
        before = engine.launch_count
        c = engine.score(m, X, pb)
        assert engine.launch_count - before >= 2
        same(c, o, method + "/compact")


def test_no_always_active_segment_stays_general_and_can_come_up_empty(engine, oracle, gen):
    T, F = 3, 6
    pm, _ = gen.gbt_pmml(T=T, depth=3, F=F, seed=313, method="sum")
    pm = splice_predicates(pm, F, seed=6, share=1.0, keep_first=False)
    assert not re.search(rb"<True/>\s*<TreeModel", pm)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "format=general" in m.describe
    X, pres, pb = gen.events(20000, F, 314, 0.5)
    o = om.predict(X, pb, node_ids=True, nthreads=8); r = engine.score(m, X, pb, node_ids=True)
    same(r, o, "sum/no-always-active")
    assert np.array_equal(r["node_ids"], o["node_ids"])
    empty = (o["node_ids"] == -1).all(axis=1)
    assert empty.any() and (o["status"][empty] == 3).all()   # no active segment: no result


@pytest.mark.parametrize("method", ["majorityVote", "weightedMajorityVote", "average", "selectFirst"])
def test_classification_segments_with_predicates(engine, oracle, gen, method):
    T, F = 33, 10
    pm, _ = gen.rf_pmml(T=T, depth=4, F=F, n_labels=4, seed=315, method=method, segment_weights=method.startswith("weighted"))
    pm = splice_predicates(pm, F, seed=7, share=0.8 if method == "selectFirst" else 0.4, keep_first=method != "selectFirst", keep_last=True)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(20000, F, 316, 0.15)
    o = om.predict(X, pb, node_ids=True, nthreads=8); r = engine.score(m, X, pb, node_ids=True)
    same(r, o, method, aux=method == "selectFirst")          # aggregated votes: aux is the winner's index, not compared here
    assert np.array_equal(r["node_ids"], o["node_ids"]), method
    assert len(np.unique(o["score"][o["status"] == 0])) >= 3


def test_radial_basis_layer(engine, oracle, gen):
    """NeuralNetwork with a radialBasis hidden layer (width / altitude at network level) against the oracle: labels
    bit-exact, probabilities within 1e-6 relative (the tolerance north_star states for floating point)."""
    pm, _ = gen.nn_pmml(n_in=16, hidden=(32,), n_out=4, seed=911, activation="radialBasis")
    pm = pm.replace(b'activationFunction="radialBasis">', b'activationFunction="radialBasis" width="3.0" altitude="1.01">', 1)
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    X, pres, pb = gen.events(20000, 16, 912, 0.01)
    o = om.predict(X, pb, n_probs=4, nthreads=8); r = engine.score(m, X, pb, n_probs=4)
    assert np.array_equal(r["status"], o["status"]) and (o["status"] == 0).sum() > 15000 and (o["status"] == 3).sum() > 100
    ok = o["status"] == 0
    assert np.array_equal(r["score"][ok], o["score"][ok])
    np.testing.assert_allclose(r["probs"][ok], o["probs"][ok], rtol=1e-6)
    assert len(np.unique(o["score"][ok])) >= 2


def test_local_transformations_of_nested_trees(engine, oracle, gen):
    """Every tree of the ensemble defines its own DerivedField -- all of them named "d" -- in its LocalTransformations; the names are
    scoped to their trees, the values are computed once per event by the derived-field pre-pass.  Against the oracle and numpy."""
    from hand_pmml import doc, schema, tree_model
    T, F, n = 24, 6, 30000
    rng = np.random.default_rng(77)
    names = [f"x{i}" for i in range(F)]
    thr = np.round(rng.normal(0, 1.2, T), 3); va = np.round(rng.normal(0, 1, T), 3); vb = np.round(rng.normal(0, 1, T), 3); dflt = rng.integers(0, 2, T)
    segs = []
    for k in range(T):
        i, j = k % F, (k + 1) % F
        lt = (f'<LocalTransformations><DerivedField name="d" optype="continuous" dataType="double"><Apply function="+">'
              f'<FieldRef field="x{i}"/><FieldRef field="x{j}"/></Apply></DerivedField></LocalTransformations>')
        nodes = (f'<Node id="1" defaultChild="{2 + dflt[k]}"><True/><Node id="2" score="{float(va[k])!r}"><SimplePredicate field="d" operator="lessOrEqual" value="{float(thr[k])!r}"/></Node>'
                 f'<Node id="3" score="{float(vb[k])!r}"><SimplePredicate field="d" operator="greaterThan" value="{float(thr[k])!r}"/></Node></Node>')
        segs.append(f'<Segment id="{k + 1}"><True/>{tree_model(names, nodes, strategy="defaultChild", target=None, extra=lt)}</Segment>')
    pm = doc([(nm, "continuous", "double") for nm in names],
             f'<MiningModel functionName="regression">{schema(names)}<Segmentation multipleModelMethod="sum">{"".join(segs)}</Segmentation></MiningModel>')
    m, om = engine.compile(pm), oracle.OracleModel(pm)
    assert "derived=24" in m.describe, m.describe
    X, pres, pb = gen.events(n, F, 78, 0.1)
    o = om.predict(X, pb, nthreads=8); r = engine.score(m, X, pb)
    assert np.array_equal(r["status"], o["status"]) and (o["status"] == 0).all()
    assert np.array_equal(r["score"].view(np.int64), o["score"].view(np.int64))
    want = np.zeros(n)
    for k in range(T):
        i, j = k % F, (k + 1) % F
        d = X[:, i] + X[:, j]; known = pres[:, i] & pres[:, j]
        left = np.where(known, d <= thr[k], dflt[k] == 0)
        want = want + np.where(left, va[k], vb[k])
    assert np.array_equal(r["score"], want)
