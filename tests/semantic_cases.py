"""Adversarial small documents with hand-derived answers (PMML 4.3 spec read the way JPMML 1.3.x implements it,
SURVEY.md Appendix A).  Each case: (name, pmml bytes, X rows, present mask or None, replace or None,
expected list of (status, score-or-None, aux-or-None))."""
import math

import numpy as np

from hand_pmml import AB, SMALL_TREE, doc, schema, tree_model

NAN = float("nan")


def _rows(*rows):
    X = np.array([[0.0 if v is None else v for v in r] for r in rows], dtype=np.float64)
    P = np.array([[v is not None for v in r] for r in rows], dtype=bool)
    return X, P


def cases():
    out = []
    X, P = _rows((0, 5), (2, -1), (2, 3), (None, 1), (2, None), (1, 0), (NAN, -1))
    t = lambda **kw: doc(AB, tree_model(["a", "b"], SMALL_TREE, **kw))
    #                         a<=1     R,b<0    R,b>=0   a missing          b missing        a==1     NaN is greatest: a>1
    out.append(("tree/none/null", t(strategy="none"), X, P, None,
                [(0, 1.5, 2), (0, -2.0, 4), (0, 7.0, 5), (3, None, None), (3, None, None), (0, 1.5, 2), (0, -2.0, 4)]))
    out.append(("tree/none/last", t(strategy="none", no_true="returnLastPrediction"), X, P, None,
                [(0, 1.5, 2), (0, -2.0, 4), (0, 7.0, 5), (0, 10.0, 1), (0, 20.0, 3), (0, 1.5, 2), (0, -2.0, 4)]))
    out.append(("tree/defaultChild", t(strategy="defaultChild"), X, P, None,
                [(0, 1.5, 2), (0, -2.0, 4), (0, 7.0, 5), (0, 7.0, 5), (0, -2.0, 4), (0, 1.5, 2), (0, -2.0, 4)]))
    out.append(("tree/lastPrediction", t(strategy="lastPrediction"), X, P, None,
                [(0, 1.5, 2), (0, -2.0, 4), (0, 7.0, 5), (0, 10.0, 1), (0, 20.0, 3), (0, 1.5, 2), (0, -2.0, 4)]))
    out.append(("tree/nullPrediction", t(strategy="nullPrediction"), X, P, None,
                [(0, 1.5, 2), (0, -2.0, 4), (0, 7.0, 5), (3, None, None), (3, None, None), (0, 1.5, 2), (0, -2.0, 4)]))
    # replaceNan fills the absent slot before preparation (api/PmmlModel.scala:148)
    out.append(("tree/replaceNan", t(strategy="nullPrediction"), X, P, 0.5,
                [(0, 1.5, 2), (0, -2.0, 4), (0, 7.0, 5), (0, 1.5, 2), (0, 7.0, 5), (0, 1.5, 2), (0, -2.0, 4)]))
    # Targets: clamp to [min,max] first, then rescale, then cast
    tg = '<Targets><Target field="y" min="0" max="5" rescaleFactor="2" rescaleConstant="0.25" castInteger="floor"/></Targets>'
    out.append(("tree/targets", t(strategy="defaultChild", extra=tg), X, P, None,
                [(0, 3.0, 2), (0, 0.0, 4), (0, 10.0, 5), (0, 10.0, 5), (0, 0.0, 4), (0, 3.0, 2), (0, 0.0, 4)]))
    # equality / set membership of missing, isMissing
    eq = ('<Node id="1"><True/>'
          '<Node id="2" score="1"><SimplePredicate field="a" operator="isMissing"/></Node>'
          '<Node id="3" score="2"><SimplePredicate field="a" operator="equal" value="3"/></Node>'
          '<Node id="4" score="3"><SimplePredicate field="b" operator="notEqual" value="0"/></Node>'
          '<Node id="5" score="4"><True/></Node></Node>')
    X2, P2 = _rows((None, 1), (3, 1), (2, 1), (2, 0), (2, None), (-0.0, 0.0), (2, -0.0))
    #   b == 0 tests use Double.equals: -0.0 != 0.0
    out.append(("tree/equal-missing", doc(AB, tree_model(["a", "b"], eq)), X2, P2, None,
                [(0, 1.0, 2), (0, 2.0, 3), (0, 3.0, 4), (0, 4.0, 5), (0, 4.0, 5), (0, 4.0, 5), (0, 3.0, 4)]))
    # compound / set predicates, three-valued logic (strategy none: UNKNOWN counts as "not true")
    sp = lambda f, op, v=None: f'<SimplePredicate field="{f}" operator="{op}"' + (f' value="{v}"/>' if v is not None else '/>')
    comp = ('<Node id="1"><True/>'
            f'<Node id="2" score="1"><CompoundPredicate booleanOperator="and">{sp("a", "lessThan", 1)}{sp("b", "lessThan", 1)}</CompoundPredicate></Node>'
            f'<Node id="3" score="2"><CompoundPredicate booleanOperator="or">{sp("a", "greaterThan", 5)}{sp("b", "isMissing")}</CompoundPredicate></Node>'
            '<Node id="4" score="3"><SimpleSetPredicate field="a" booleanOperator="isIn"><Array n="2" type="real">2 3</Array></SimpleSetPredicate></Node>'
            f'<Node id="5" score="4"><CompoundPredicate booleanOperator="surrogate">{sp("a", "greaterOrEqual", 4)}{sp("b", "greaterOrEqual", 0)}</CompoundPredicate></Node>'
            f'<Node id="6" score="5"><CompoundPredicate booleanOperator="xor">{sp("a", "greaterThan", 0)}{sp("b", "greaterThan", 0)}</CompoundPredicate></Node>'
            '<Node id="7" score="6"><True/></Node></Node>')
    Xc, Pc = _rows((0, 0), (0, 2), (0, None), (None, 0), (3, 5), (6, 2), (4.5, -1), (1, -1), (-1, 2), (1, 1), (None, None), (None, 2))
    out.append(("tree/compound-and-set", doc(AB, tree_model(["a", "b"], comp)), Xc, Pc, None,
                [(0, 1.0, 2), (0, 5.0, 6), (0, 2.0, 3), (0, 4.0, 5), (0, 3.0, 4), (0, 2.0, 3), (0, 4.0, 5), (0, 5.0, 6), (0, 5.0, 6), (0, 6.0, 7),
                 (0, 2.0, 3), (0, 4.0, 5)]))
    notin = ('<Node id="1"><True/><Node id="2" score="1"><SimpleSetPredicate field="a" booleanOperator="isNotIn"><Array n="3" type="real">1 2 3</Array></SimpleSetPredicate></Node>'
             '<Node id="3" score="2"><True/></Node></Node>')
    Xn, Pn = _rows((2, 0), (2.5, 0), (None, 0))
    out.append(("tree/isNotIn-nullPrediction", doc(AB, tree_model(["a", "b"], notin, strategy="nullPrediction")), Xn, Pn, None,
                [(0, 2.0, 3), (0, 1.0, 2), (3, None, None)]))
    # classification tree: labels become doubles only when they parse; integer-typed target never extracts
    cls = ('<Node id="1"><True/><Node id="2" score="7"><SimplePredicate field="a" operator="lessThan" value="0"/></Node>'
           '<Node id="3" score="setosa"><True/></Node></Node>')
    X3, P3 = _rows((-1, 0), (1, 0))
    tgt = ("y", "categorical", "string", ("7", "setosa"))
    out.append(("tree/classification-labels", doc(AB, tree_model(["a", "b"], cls, function="classification"), target=tgt), X3, P3, None,
                [(0, 7.0, 2), (4, None, 3)]))
    tgt_i = ("y", "categorical", "integer", ("7", "8"))
    out.append(("tree/integer-target", doc(AB, tree_model(["a", "b"], cls.replace("setosa", "8"), function="classification"), target=tgt_i), X3, P3, None,
                [(4, None, 2), (4, None, 3)]))
    out.append(("tree/float-regression-target", doc(AB, tree_model(["a", "b"], SMALL_TREE, strategy="defaultChild"), target=("y", "continuous", "float", ())),
                X[:3], P[:3], None, [(4, None, 2), (4, None, 4), (4, None, 5)]))
    # Math.round: half goes up also below zero (-2 * 1.25 = -2.5 -> -2; C's round() would give -3)
    tr = '<Targets><Target field="y" rescaleFactor="1.25" castInteger="round"/></Targets>'
    out.append(("tree/targets-round-negative-half", t(strategy="defaultChild", extra=tr), X[:3], P[:3], None, [(0, 2.0, 2), (0, -2.0, 4), (0, 9.0, 5)]))
    # a tree over a string-typed field loads; a Double offered for that field fails preparation (status 2, as for the
    # reference's k-means string fixture), an absent slot makes the test UNKNOWN
    sfields = [("a", "categorical", "string"), ("b", "continuous", "double")]
    stree = ('<Node id="1" score="0" defaultChild="3"><True/><Node id="2" score="1"><SimplePredicate field="a" operator="equal" value="x"/></Node>'
             '<Node id="3" score="2"><SimpleSetPredicate field="a" booleanOperator="isIn"><Array n="2" type="string">x y</Array></SimpleSetPredicate></Node>'
             '<Node id="4" score="3"><SimplePredicate field="b" operator="lessThan" value="0"/></Node></Node>')
    Xs, Ps = _rows((1, 0), (None, -1), (None, 1), (None, None))
    out.append(("tree/string-field-none", doc(sfields, tree_model(["a", "b"], stree, strategy="none")), Xs, Ps, None,
                [(2, None, None), (0, 3.0, 4), (3, None, None), (3, None, None)]))
    out.append(("tree/string-field-defaultChild", doc(sfields, tree_model(["a", "b"], stree, strategy="defaultChild")), Xs, Ps, None,
                [(2, None, None), (0, 2.0, 3), (0, 2.0, 3), (0, 2.0, 3)]))
    out.append(("tree/string-field-last", doc(sfields, tree_model(["a", "b"], stree, strategy="none", no_true="returnLastPrediction")), Xs, Ps, None,
                [(2, None, None), (0, 3.0, 4), (0, 0.0, 1), (0, 0.0, 1)]))
    # every ACTIVE field is prepared before the model runs (S/api/PmmlModel.scala:143-152): an invalid value fails the event
    # even when the model never reads that field
    ivu = '<Interval closure="closedOpen" leftMargin="0" rightMargin="10"/>'
    ufields = [("a", "continuous", "double", ivu), ("b", "continuous", "double")]
    only_b = ('<Node id="1"><True/><Node id="2" score="1"><SimplePredicate field="b" operator="lessThan" value="0"/></Node><Node id="3" score="2"><True/></Node></Node>')
    Xu, Pu = _rows((5, -1), (20, -1), (None, 1), (-3, 1))
    out.append(("prep/unused-field-tree", doc(ufields, tree_model(["a", "b"], only_b)), Xu, Pu, None, [(0, 1.0, 2), (2, None, None), (0, 2.0, 3), (2, None, None)]))
    ureg = (f'<RegressionModel modelName="r" functionName="regression">{schema(["a", "b"])}'
            '<RegressionTable intercept="1"><NumericPredictor name="b" coefficient="2"/></RegressionTable></RegressionModel>')
    out.append(("prep/unused-field-regression", doc(ufields, ureg), Xu, Pu, None, [(0, -1.0, None), (2, None, None), (0, 3.0, None), (2, None, None)]))
    # ---- input preparation (JPMML FieldValueUtil.prepareInputValue)
    iv = '<Interval closure="closedOpen" leftMargin="0" rightMargin="10"/>'
    fields = [("a", "continuous", "double", iv), ("b", "continuous", "double")]
    X4, P4 = _rows((5, 1), (10, 1), (-1, 1), (0, 1), (None, 1))
    mk = lambda attr: doc(fields, tree_model(["a", "b"], SMALL_TREE, strategy="nullPrediction", attrs={"a": attr}))
    out.append(("prep/returnInvalid", mk(""), X4, P4, None, [(0, 7.0, 5), (2, None, None), (2, None, None), (0, 1.5, 2), (3, None, None)]))
    out.append(("prep/asIs", mk('invalidValueTreatment="asIs"'), X4, P4, None, [(0, 7.0, 5), (0, 7.0, 5), (0, 1.5, 2), (0, 1.5, 2), (3, None, None)]))
    out.append(("prep/asMissing+replacement", mk('invalidValueTreatment="asMissing" missingValueReplacement="0.5"'), X4, P4, None,
                [(0, 7.0, 5), (0, 1.5, 2), (0, 1.5, 2), (0, 1.5, 2), (0, 1.5, 2)]))
    out.append(("prep/outliers-extreme", doc(AB, tree_model(["a", "b"], SMALL_TREE, strategy="nullPrediction",
                                                             attrs={"a": 'outliers="asExtremeValues" lowValue="1.5" highValue="3"'})),
                X4, P4, None, [(0, 7.0, 5), (0, 7.0, 5), (0, 7.0, 5), (0, 7.0, 5), (3, None, None)]))
    out.append(("prep/outliers-missing", doc(AB, tree_model(["a", "b"], SMALL_TREE, strategy="nullPrediction",
                                                             attrs={"a": 'outliers="asMissingValues" lowValue="-5" highValue="8"'})),
                X4, P4, None, [(0, 7.0, 5), (3, None, None), (0, 1.5, 2), (0, 1.5, 2), (3, None, None)]))
    vals = '<Value value="1"/><Value value="2"/><Value value="9" property="missing"/><Value value="8" property="invalid"/>'
    fields_i = [("a", "categorical", "integer", vals), ("b", "continuous", "double")]
    X5, P5 = _rows((1, 1), (2, 1), (3, 1), (9, 1), (8, 1), (1.5, 1))
    out.append(("prep/integer-values", doc(fields_i, tree_model(["a", "b"], SMALL_TREE, strategy="nullPrediction")), X5, P5, None,
                [(0, 1.5, 2), (0, 7.0, 5), (2, None, None), (3, None, None), (2, None, None), (2, None, None)]))
    # ---- ensembles
    seg = lambda i, w, nodes: f'<Segment id="{i}" weight="{w}"><True/>{tree_model(["a", "b"], nodes, strategy="defaultChild", target=None)}</Segment>'
    t2 = ('<Node id="1" defaultChild="2"><True/><Node id="2" score="100"><SimplePredicate field="b" operator="lessOrEqual" value="0"/></Node>'
          '<Node id="3" score="200"><SimplePredicate field="b" operator="greaterThan" value="0"/></Node></Node>')
    mm = lambda method: doc(AB, f'<MiningModel functionName="regression">{schema(["a", "b"])}<Segmentation multipleModelMethod="{method}">'
                            f'{seg(1, 1, SMALL_TREE)}{seg(2, 3, t2)}</Segmentation></MiningModel>')
    X6, P6 = _rows((0, 5), (2, -1), (None, None))
    out.append(("mining/sum", mm("sum"), X6, P6, None, [(0, 201.5, None), (0, 98.0, None), (0, 98.0, None)]))
    out.append(("mining/average", mm("average"), X6, P6, None, [(0, 100.75, None), (0, 49.0, None), (0, 49.0, None)]))
    out.append(("mining/weightedSum", mm("weightedSum"), X6, P6, None, [(0, 601.5, None), (0, 298.0, None), (0, 298.0, None)]))
    out.append(("mining/weightedAverage", mm("weightedAverage"), X6, P6, None, [(0, 150.375, None), (0, 74.5, None), (0, 74.5, None)]))
    # majority vote: tie between "1" and "0" -> the label seen first in segment order wins
    cseg = lambda i, w, lab_l, lab_r: (f'<Segment id="{i}" weight="{w}"><True/>' + tree_model(["a", "b"],
                                       f'<Node id="1" defaultChild="3"><True/><Node id="2" score="{lab_l}"><SimplePredicate field="a" operator="lessOrEqual" value="0"/></Node>'
                                       f'<Node id="3" score="{lab_r}"><True/></Node></Node>', function="classification", strategy="defaultChild", target=None) + '</Segment>')
    vote = lambda method: doc(AB, f'<MiningModel functionName="classification">{schema(["a", "b"])}<Segmentation multipleModelMethod="{method}">'
                              f'{cseg(1, 1, "1", "0")}{cseg(2, 1, "0", "1")}{cseg(3, 5, "2", "2")}{cseg(4, 1, "2", "0")}</Segmentation></MiningModel>',
                              target=("y", "categorical", "string", ("0", "1", "2")))
    X7, P7 = _rows((-1, 0), (1, 0))
    #  a<=0: votes 1,0,2,2 -> "2"(2 votes).   a>0: votes 0,1,2,0 -> "0"(2 votes)
    out.append(("mining/majorityVote", vote("majorityVote"), X7, P7, None, [(0, 2.0, None), (0, 0.0, None)]))
    #  weighted: a<=0: 1:1, 0:1, 2:6 -> "2";  a>0: 0:2, 1:1, 2:5 -> "2"
    out.append(("mining/weightedMajorityVote", vote("weightedMajorityVote"), X7, P7, None, [(0, 2.0, None), (0, 2.0, None)]))
    tie = lambda: doc(AB, f'<MiningModel functionName="classification">{schema(["a", "b"])}<Segmentation multipleModelMethod="majorityVote">'
                      f'{cseg(1, 1, "1", "0")}{cseg(2, 1, "0", "1")}</Segmentation></MiningModel>', target=("y", "categorical", "string", ("0", "1")))
    out.append(("mining/vote-tie-first-seen", tie(), X7, P7, None, [(0, 1.0, None), (0, 0.0, None)]))
    # ---- Segment predicates: only segments whose predicate is TRUE take part (UNKNOWN counts as not TRUE)
    gt = lambda f, v: f'<SimplePredicate field="{f}" operator="greaterThan" value="{v}"/>'
    miss_b = '<SimplePredicate field="b" operator="isMissing"/>'
    const = lambda v: f'<Node id="1" score="{v}"><True/></Node>'
    pseg = lambda i, w, pred, nodes: f'<Segment id="{i}" weight="{w}">{pred}{tree_model(["a", "b"], nodes, strategy="defaultChild", target=None)}</Segment>'
    pm_ = lambda method, *segs: doc(AB, f'<MiningModel functionName="regression">{schema(["a", "b"])}<Segmentation multipleModelMethod="{method}">'
                                        + "".join(segs) + '</Segmentation></MiningModel>')
    s_all, s_agt1, s_bmiss = pseg(1, 1, "<True/>", SMALL_TREE), pseg(2, 3, gt("a", 1), t2), pseg(3, 1, miss_b, const(1000))
    Xp, Pp = _rows((0, 5), (2, -1), (None, None), (2, None))
    #   S1 (always): 1.5, -2, -2, -2.   S2 (a > 1): -, 100, - (UNKNOWN), 100.   S3 (b missing): -, -, 1000, 1000
    out.append(("mining/segment-predicates-sum", pm_("sum", s_all, s_agt1, s_bmiss), Xp, Pp, None,
                [(0, 1.5, None), (0, 98.0, None), (0, 998.0, None), (0, 1098.0, None)]))
    out.append(("mining/segment-predicates-average", pm_("average", s_all, s_agt1, s_bmiss), Xp, Pp, None,
                [(0, 1.5, None), (0, 49.0, None), (0, 499.0, None), (0, 366.0, None)]))
    out.append(("mining/segment-predicates-weightedAverage", pm_("weightedAverage", s_all, s_agt1, s_bmiss), Xp, Pp, None,
                [(0, 1.5, None), (0, 74.5, None), (0, 499.0, None), (0, (-2.0 + 300.0 + 1000.0) / 5.0, None)]))
    # no always-active segment: an event none of the predicates accepts has no result at all
    out.append(("mining/no-active-segment", pm_("sum", s_agt1, s_bmiss), Xp, Pp, None,
                [(3, None, None), (0, 100.0, None), (0, 1000.0, None), (0, 1100.0, None)]))
    # selectFirst: the first segment whose predicate holds decides alone (aux = that tree's node id)
    out.append(("mining/selectFirst", pm_("selectFirst", s_agt1, s_bmiss, s_all), Xp, Pp, None,
                [(0, 1.5, 2), (0, 100.0, 2), (0, 1000.0, 1), (0, 100.0, 2)]))
    out.append(("mining/selectFirst-none", pm_("selectFirst", s_agt1, s_bmiss), Xp, Pp, None,
                [(3, None, None), (0, 100.0, 2), (0, 1000.0, 1), (0, 100.0, 2)]))
    pcseg = lambda i, pred, lab_l, lab_r: cseg(i, 1, lab_l, lab_r).replace("<True/>", pred, 1)
    sfc = doc(AB, f'<MiningModel functionName="classification">{schema(["a", "b"])}<Segmentation multipleModelMethod="selectFirst">'
                  f'{pcseg(1, gt("b", 0), "1", "0")}{pcseg(2, "<True/>", "2", "0")}</Segmentation></MiningModel>', target=("y", "categorical", "string", ("0", "1", "2")))
    Xf, Pf = _rows((-1, 1), (-1, 0), (1, None), (1, 3))
    out.append(("mining/selectFirst-classification", sfc, Xf, Pf, None, [(0, 1.0, 2), (0, 2.0, 2), (0, 0.0, 3), (0, 0.0, 3)]))
    # median (odd count: the middle value; even count: lower + (upper - lower) / 2) and max over the active segments
    s_c3 = pseg(4, 1, "<True/>", const(3))
    s_c1000_agt1 = pseg(3, 1, gt("a", 1), const(1000))
    plain_t2 = pseg(2, 1, "<True/>", t2)
    #   (0,5): 1.5, 200, -, 3 -> sorted 1.5 3 200;   (2,-1): -2, 100, 1000, 3 -> sorted -2 3 100 1000
    out.append(("mining/median", pm_("median", s_all, plain_t2, s_c1000_agt1, s_c3), Xp[:2], Pp[:2], None, [(0, 3.0, None), (0, 51.5, None)]))
    out.append(("mining/max", pm_("max", s_all, plain_t2, s_c1000_agt1, s_c3), Xp[:2], Pp[:2], None, [(0, 200.0, None), (0, 1000.0, None)]))
    # a root Node whose own predicate is not TRUE: null prediction, whatever the strategies say
    rooted = (f'<Node id="1" score="9">{gt("a", 0)}<Node id="2" score="1"><SimplePredicate field="b" operator="lessThan" value="0"/></Node></Node>')
    Xr, Pr = _rows((1, -1), (-1, -1), (None, -1), (1, 1))
    out.append(("tree/root-predicate", doc(AB, tree_model(["a", "b"], rooted, strategy="lastPrediction", no_true="returnLastPrediction")), Xr, Pr, None,
                [(0, 1.0, 2), (3, None, None), (3, None, None), (0, 9.0, 1)]))
    # ---- LocalTransformations inside nested models: a DerivedField is visible in its own model only -- both trees call theirs "d"
    apply2 = lambda fn, x, y: f'<Apply function="{fn}">{x}{y}</Apply>'
    fref = lambda f: f'<FieldRef field="{f}"/>'
    kdbl = lambda v: f'<Constant dataType="double">{v}</Constant>'
    dfield = lambda n, ex: f'<DerivedField name="{n}" optype="continuous" dataType="double">{ex}</DerivedField>'
    lt1 = f'<LocalTransformations>{dfield("d", apply2("*", fref("a"), kdbl(2)))}</LocalTransformations>'
    lt2 = f'<LocalTransformations>{dfield("d", apply2("+", fref("b"), kdbl(1)))}{dfield("e", apply2("*", fref("d"), fref("d")))}</LocalTransformations>'
    nt1 = ('<Node id="1" defaultChild="3"><True/><Node id="2" score="10"><SimplePredicate field="d" operator="lessOrEqual" value="3"/></Node>'
           '<Node id="3" score="20"><SimplePredicate field="d" operator="greaterThan" value="3"/></Node></Node>')
    nt2 = ('<Node id="1" defaultChild="3"><True/><Node id="2" score="1"><SimplePredicate field="e" operator="greaterThan" value="4"/></Node>'
           '<Node id="3" score="2"><SimplePredicate field="e" operator="lessOrEqual" value="4"/></Node></Node>')
    nseg = lambda i, pred, lt, nodes: f'<Segment id="{i}">{pred}{tree_model(["a", "b"], nodes, strategy="defaultChild", target=None, extra=lt)}</Segment>'
    nested = pm_("sum", nseg(1, "<True/>", lt1, nt1), nseg(2, "<True/>", lt2, nt2), pseg(3, 1, gt("a", 1), const(100)))
    #   (1,5): d1 = 2 -> 10; d2 = 6, e = 36 -> 1; a > 1 false.   (2,-1): d1 = 4 -> 20; d2 = 0, e = 0 -> 2; +100.
    #   (None,None): both derived values missing -> defaultChild: 20 + 2; predicate UNKNOWN.   (2,5): 20 + 1 + 100
    Xl, Pl = _rows((1, 5), (2, -1), (None, None), (2, 5))
    out.append(("mining/nested-local-transformations", nested, Xl, Pl, None, [(0, 11.0, None), (0, 122.0, None), (0, 22.0, None), (0, 121.0, None)]))
    # ---- clustering variants: 1-D centres 0 and 10 plus a second field; cluster ids "5" and "x"
    def clu(measure, compare="absDiff", extra="", ids=("5", "x"), weights=""):
        body = (f'<ClusteringModel modelName="c" functionName="clustering" modelClass="centerBased" numberOfClusters="2">'
                f'<MiningSchema><MiningField name="a"/><MiningField name="b"/><MiningField name="y" usageType="predicted"/></MiningSchema>'
                f'<ComparisonMeasure kind="distance" compareFunction="{compare}"><{measure}/></ComparisonMeasure>'
                f'<ClusteringField field="a"{weights}/><ClusteringField field="b"/>{extra}'
                f'<Cluster id="{ids[0]}"><Array n="2" type="real">0 0</Array></Cluster><Cluster id="{ids[1]}"><Array n="2" type="real">10 1</Array></Cluster></ClusteringModel>')
        return doc(AB, body, target=("y", "categorical", "string", ()))
    X8, P8 = _rows((1, 0), (9, 0), (5, 0.5), (None, 0.9), (None, None))
    # ids: first cluster "5" -> 5.0; second "x" does not parse -> status 4; exact tie (5,0.5) -> first cluster
    out.append(("cluster/squaredEuclidean-ids", clu("squaredEuclidean"), X8, P8, None, [(0, 5.0, 1), (4, None, 2), (0, 5.0, 1), (4, None, 2), (0, 5.0, 1)]))
    out.append(("cluster/chebychev", clu("chebychev", ids=("1", "2")), X8, P8, None, [(0, 1.0, 1), (0, 2.0, 2), (0, 1.0, 1), (0, 2.0, 2), (0, 1.0, 1)]))
    out.append(("cluster/cityBlock-weighted", clu("cityBlock", ids=("1", "2"), weights=' fieldWeight="0.01"'), X8, P8, None,
                [(0, 1.0, 1), (0, 1.0, 1), (0, 1.0, 1), (0, 2.0, 2), (0, 1.0, 1)]))
    # ---- GeneralRegressionModel: p0 = intercept, p1 = a, p2 = a^2 * [b == 1]; beta = 0.5, 2, -1
    def greg(model_type, function="regression", link="", cats=(), pcells=None, target_dtype="double"):
        pc = pcells or '<PCell parameterName="p0" beta="0.5"/><PCell parameterName="p1" beta="2"/><PCell parameterName="p2" beta="-1"/>'
        body = (f'<GeneralRegressionModel modelName="g" functionName="{function}" modelType="{model_type}"{link}>{schema(["a", "b"])}'
                '<ParameterList><Parameter name="p0"/><Parameter name="p1"/><Parameter name="p2"/></ParameterList>'
                '<FactorList><Predictor name="b"/></FactorList><CovariateList><Predictor name="a"/></CovariateList>'
                '<PPMatrix><PPCell value="1" predictorName="a" parameterName="p1"/><PPCell value="2" predictorName="a" parameterName="p2"/>'
                '<PPCell value="1" predictorName="b" parameterName="p2"/></PPMatrix>'
                f'<ParamMatrix>{pc}</ParamMatrix></GeneralRegressionModel>')
        tgt = ("y", "categorical", target_dtype, cats) if cats else ("y", "continuous", "double", ())
        return doc(AB, body, target=tgt)
    X9, P9 = _rows((2, 1), (2, 0), (-1, 1), (None, 1), (3, None))
    out.append(("greg/regression-identity", greg("regression"), X9, P9, None, [(0, 0.5, None), (0, 4.5, None), (0, -2.5, None), (3, None, None), (3, None, None)]))
    out.append(("greg/link-log", greg("generalLinear", link=' linkFunction="log"'), X9, P9, None,
                [(0, math.exp(0.5), None), (0, math.exp(4.5), None), (0, math.exp(-2.5), None), (3, None, None), (3, None, None)]))
    out.append(("greg/link-logit-offset", greg("generalLinear", link=' linkFunction="logit" offsetValue="1"'), X9, P9, None,
                [(0, 1 / (1 + math.exp(-1.5)), None), (0, 1 / (1 + math.exp(-5.5)), None), (0, 1 / (1 + math.exp(1.5)), None), (3, None, None), (3, None, None)]))
    # multinomial: eta("0") = 0.5 + 2a - a^2[b==1], eta("1") = 1 (p0 only, beta 1), "2" is the reference (eta 0)
    pc = ('<PCell targetCategory="0" parameterName="p0" beta="0.5"/><PCell targetCategory="0" parameterName="p1" beta="2"/>'
          '<PCell targetCategory="0" parameterName="p2" beta="-1"/><PCell targetCategory="1" parameterName="p0" beta="1"/>')
    #  (2,1): eta0 = 0.5 -> "1" wins (eta 1);  (2,0): eta0 = 4.5 -> "0";  (-1,1): eta0 = -2.5 -> "1";  a = 0.25,b=0: eta0 = 1.0 == eta1 -> tie, first ("0")
    X10, P10 = _rows((2, 1), (2, 0), (-1, 1), (0.25, 0), (None, 0))
    out.append(("greg/multinomial", greg("multinomialLogistic", function="classification", cats=("0", "1", "2"), pcells=pc, target_dtype="string"), X10, P10, None,
                [(0, 1.0, 1), (0, 0.0, 0), (0, 1.0, 1), (0, 0.0, 0), (3, None, None)]))
    # ---- derived fields (TransformationDictionary + LocalTransformations) feeding a one-predictor regression y = 1.0 * d
    fr = lambda f: f'<FieldRef field="{f}"/>'
    cst = lambda v: f'<Constant dataType="double">{v}</Constant>'
    # DefineFunction: scale(x, k) = x * (k + 1) with k's own mapMissingTo inside the body; f32(v) = scale(v, 0.1) as a float
    fns = ('<DefineFunction name="scale" optype="continuous" dataType="double"><ParameterField name="x"/><ParameterField name="k" dataType="double"/>'
           f'<Apply function="*">{fr("x")}<Apply function="+"><FieldRef field="k" mapMissingTo="10"/>{cst(1)}</Apply></Apply></DefineFunction>'
           '<DefineFunction name="f32" optype="continuous" dataType="float"><ParameterField name="v"/>'
           f'<Apply function="scale">{fr("v")}{cst(0.1)}</Apply></DefineFunction>')
    td = ('<TransformationDictionary>' + fns +
          '<DerivedField name="d1" optype="continuous" dataType="double"><NormContinuous field="a"><LinearNorm orig="0" norm="0"/><LinearNorm orig="10" norm="1"/></NormContinuous></DerivedField>'
          f'<DerivedField name="d2" optype="continuous" dataType="double"><Apply function="+">{fr("a")}{cst(2)}</Apply></DerivedField>'
          f'<DerivedField name="d3" optype="continuous" dataType="double"><Apply function="if"><Apply function="greaterThan">{fr("a")}{cst(1)}</Apply>'
          f'<Apply function="*">{fr("b")}{cst(3)}</Apply>{cst(-1)}</Apply></DerivedField>'
          f'<DerivedField name="d7" optype="continuous" dataType="double"><Apply function="scale">{fr("a")}{fr("b")}</Apply></DerivedField>'
          f'<DerivedField name="d8" optype="continuous" dataType="double"><Apply function="scale" mapMissingTo="-7">{fr("a")}{fr("b")}</Apply></DerivedField>'
          f'<DerivedField name="d9" optype="continuous" dataType="double"><Apply function="f32" defaultValue="5">{fr("a")}</Apply></DerivedField>'
          '</TransformationDictionary>')
    lt = ('<LocalTransformations>'
          '<DerivedField name="d4" optype="continuous" dataType="double"><Discretize field="b" mapMissingTo="7" defaultValue="99">'
          '<DiscretizeBin binValue="-5"><Interval closure="openOpen" rightMargin="0"/></DiscretizeBin>'
          '<DiscretizeBin binValue="5"><Interval closure="closedClosed" leftMargin="0" rightMargin="5"/></DiscretizeBin></Discretize></DerivedField>'
          '<DerivedField name="d5" optype="continuous" dataType="double"><MapValues outputColumn="o" defaultValue="0"><FieldColumnPair field="a" column="i"/>'
          '<InlineTable><row><i>1</i><o>100</o></row><row><i>2</i><o>200</o></row></InlineTable></MapValues></DerivedField>'
          f'<DerivedField name="d6" optype="continuous" dataType="double"><Apply function="max" mapMissingTo="-100">{fr("d2")}{fr("d3")}</Apply></DerivedField>'
          '</LocalTransformations>')
    f32 = lambda v: float(np.float32(v))
    def derived_doc(name):
        body = (f'<RegressionModel modelName="r" functionName="regression">{schema(["a", "b"])}{lt}'
                f'<RegressionTable intercept="0"><NumericPredictor name="{name}" coefficient="1"/></RegressionTable></RegressionModel>')
        d = doc(AB, body).decode()
        return d.replace("</DataDictionary>", "</DataDictionary>" + td).encode()
    Xd, Pd = _rows((2, 1), (0.5, 4), (None, 1), (2, None), (20, -3))
    M = (3, None, None)
    expect = {"d1": [(0, 0.2, None), (0, 0.05, None), M, (0, 0.2, None), (0, 2.0, None)],
              "d2": [(0, 4.0, None), (0, 2.5, None), M, (0, 4.0, None), (0, 22.0, None)],
              "d3": [(0, 3.0, None), (0, -1.0, None), M, M, (0, -9.0, None)],
              "d4": [(0, 5.0, None), (0, 5.0, None), (0, 5.0, None), (0, 7.0, None), (0, -5.0, None)],
              "d5": [(0, 200.0, None), (0, 0.0, None), M, (0, 200.0, None), (0, 0.0, None)],
              "d6": [(0, 4.0, None), (0, 2.5, None), (0, -100.0, None), (0, -100.0, None), (0, 22.0, None)],
              "d7": [(0, 4.0, None), (0, 2.5, None), M, (0, 22.0, None), (0, -40.0, None)],
              "d8": [(0, 4.0, None), (0, 2.5, None), (0, -7.0, None), (0, -7.0, None), (0, -40.0, None)],
              "d9": [(0, f32(2 * (0.1 + 1)), None), (0, f32(0.5 * (0.1 + 1)), None), (0, 5.0, None), (0, f32(2 * (0.1 + 1)), None), (0, f32(20 * (0.1 + 1)), None)]}
    for name, exp in expect.items():
        out.append((f"derived/{name}", derived_doc(name), Xd, Pd, None, exp))
    # ---- NeuralNetwork with a radial basis layer: h = exp(fanIn ln(altitude) - sum (x - w)^2 / (2 width^2)), no bias;
    #      h1: centre (0,0), width 1 (network), altitude 1;  h2: centre (1,1), width 2 (neuron), altitude 2 (layer);  y = 0.5 + h1 + 2 h2
    inp = lambda i, f: f'<NeuralInput id="i{i}"><DerivedField optype="continuous" dataType="double"><FieldRef field="{f}"/></DerivedField></NeuralInput>'
    rbf = (f'<NeuralNetwork modelName="rbf" functionName="regression" activationFunction="radialBasis" width="1">{schema(["a", "b"])}'
           f'<NeuralInputs numberOfInputs="2">{inp(0, "a")}{inp(1, "b")}</NeuralInputs>'
           '<NeuralLayer numberOfNeurons="2" altitude="2">'
           '<Neuron id="h1" bias="99" altitude="1"><Con from="i0" weight="0"/><Con from="i1" weight="0"/></Neuron>'
           '<Neuron id="h2" width="2"><Con from="i0" weight="1"/><Con from="i1" weight="1"/></Neuron></NeuralLayer>'
           '<NeuralLayer numberOfNeurons="1" activationFunction="identity"><Neuron id="o" bias="0.5"><Con from="h1" weight="1"/><Con from="h2" weight="2"/></Neuron></NeuralLayer>'
           '<NeuralOutputs numberOfOutputs="1"><NeuralOutput outputNeuron="o"><DerivedField optype="continuous" dataType="double"><FieldRef field="y"/></DerivedField></NeuralOutput></NeuralOutputs>'
           '</NeuralNetwork>')
    Xn_, Pn_ = _rows((0, 0), (1, 1), (2, 0), (None, 1))
    out.append(("neural/radialBasis", doc(AB, rbf), Xn_, Pn_, None,
                [(0, 0.5 + 1.0 + 8.0 * math.exp(-0.25), None), (0, 0.5 + math.exp(-1.0) + 8.0, None),
                 (0, 0.5 + math.exp(-2.0) + 8.0 * math.exp(-0.25), None), (3, None, None)]))
    return out


# cases whose expected scores go through exp(): compared with a relative tolerance instead of bit-exactly
TOLERANT = {"greg/link-log": 1e-12, "greg/link-logit-offset": 1e-12, "neural/radialBasis": 1e-12}
