"""Small hand-written PMML documents with answers derivable by hand, plus a scikit-learn -> PMML exporter.

The reference holds no fixture for trees / mining / regression / neural networks, so these are the
known-answer tests for those families."""
import numpy as np


def doc(fields, body, target=("y", "continuous", "double", ())):
    out = ['<PMML version="4.3" xmlns="http://www.dmg.org/PMML-4_3"><Header/>', f'<DataDictionary numberOfFields="{len(fields) + 1}">']
    for f in fields:
        name, optype, dtype = f[:3]
        inner = f[3] if len(f) > 3 else ""
        out.append(f'<DataField name="{name}" optype="{optype}" dataType="{dtype}">{inner}</DataField>')
    if target:
        vals = "".join(f'<Value value="{v}"/>' for v in target[3])
        out.append(f'<DataField name="{target[0]}" optype="{target[1]}" dataType="{target[2]}">{vals}</DataField>')
    out.append("</DataDictionary>")
    out.append(body)
    out.append("</PMML>")
    return "\n".join(out).encode()


def schema(names, target="y", attrs=None):
    attrs = attrs or {}
    s = "<MiningSchema>" + (f'<MiningField name="{target}" usageType="target"/>' if target else "")
    for n in names:
        s += f'<MiningField name="{n}" {attrs.get(n, "")}/>'
    return s + "</MiningSchema>"


def tree_model(names, nodes_xml, function="regression", strategy="none", no_true="returnNullPrediction", target="y", attrs=None, extra=""):
    return (f'<TreeModel modelName="t" functionName="{function}" missingValueStrategy="{strategy}" noTrueChildStrategy="{no_true}">'
            f'{schema(names, target, attrs)}{extra}{nodes_xml}</TreeModel>')


# A 3-leaf regression tree used by several tests:
#   root(score 10) -> [a <= 1 -> leaf L(1.5)] [a > 1 -> node R(score 20) -> [b < 0 -> leaf RL(-2)] [b >= 0 -> leaf RR(7)]]
SMALL_TREE = ('<Node id="1" score="10" defaultChild="3"><True/>'
              '<Node id="2" score="1.5"><SimplePredicate field="a" operator="lessOrEqual" value="1"/></Node>'
              '<Node id="3" score="20" defaultChild="4"><SimplePredicate field="a" operator="greaterThan" value="1"/>'
              '<Node id="4" score="-2"><SimplePredicate field="b" operator="lessThan" value="0"/></Node>'
              '<Node id="5" score="7"><SimplePredicate field="b" operator="greaterOrEqual" value="0"/></Node>'
              '</Node></Node>')
AB = [("a", "continuous", "double"), ("b", "continuous", "double")]


def sklearn_gbr_to_pmml(gb, F):
    """GradientBoostingRegressor -> MiningModel(sum) of TreeModels (float32 thresholds as sklearn uses them), with the
    init estimator's constant as rescaleConstant and the learning rate folded into the leaf scores."""
    names = [f"x{i}" for i in range(F)]
    out = ['<PMML version="4.3" xmlns="http://www.dmg.org/PMML-4_3"><Header/>', f'<DataDictionary numberOfFields="{F + 1}">']
    out += [f'<DataField name="{n}" optype="continuous" dataType="double"/>' for n in names]
    out.append('<DataField name="y" optype="continuous" dataType="double"/></DataDictionary>')
    out.append('<MiningModel functionName="regression"><MiningSchema><MiningField name="y" usageType="target"/>')
    out += [f'<MiningField name="{n}"/>' for n in names]
    out.append("</MiningSchema>")
    init = float(gb.init_.constant_.ravel()[0])
    out.append(f'<Targets><Target field="y" rescaleConstant="{init!r}"/></Targets><Segmentation multipleModelMethod="sum">')
    lr = gb.learning_rate
    for k, est in enumerate(gb.estimators_.ravel()):
        t = est.tree_

        def emit(i, pred):
            if t.children_left[i] < 0:
                return f'<Node score="{float(lr * t.value[i].ravel()[0])!r}">{pred}</Node>'
            f, thr = names[t.feature[i]], float(t.threshold[i])
            left = emit(t.children_left[i], f'<SimplePredicate field="{f}" operator="lessOrEqual" value="{thr!r}"/>')
            right = emit(t.children_right[i], f'<SimplePredicate field="{f}" operator="greaterThan" value="{thr!r}"/>')
            return f"<Node>{pred}{left}{right}</Node>"

        out.append(f'<Segment id="{k}"><True/><TreeModel functionName="regression"><MiningSchema>' + "".join(f'<MiningField name="{n}"/>' for n in names)
                   + f'</MiningSchema>{emit(0, "<True/>")}</TreeModel></Segment>')
    out.append("</Segmentation></MiningModel></PMML>")
    return "\n".join(out).encode()


# ---- scikit-learn exporters for the independent-arithmetic cross-checks (tests/test_gpu_sklearn.py) --------------------
def _dd(names, target, labels=None):
    out = ['<PMML version="4.3" xmlns="http://www.dmg.org/PMML-4_3"><Header/>', f'<DataDictionary numberOfFields="{len(names) + 1}">']
    out += [f'<DataField name="{n}" optype="continuous" dataType="double"/>' for n in names]
    if labels is None:
        out.append(f'<DataField name="{target}" optype="continuous" dataType="double"/></DataDictionary>')
    else:
        out.append(f'<DataField name="{target}" optype="categorical" dataType="string">' + "".join(f'<Value value="{v}"/>' for v in labels) + '</DataField></DataDictionary>')
    return out


def _ms(names, target="y"):
    return "<MiningSchema>" + (f'<MiningField name="{target}" usageType="target"/>' if target else "") + "".join(f'<MiningField name="{n}"/>' for n in names) + "</MiningSchema>"


def sklearn_rf_to_pmml(rf, F):
    """RandomForestClassifier -> MiningModel(average) of classification TreeModels whose leaves carry the class
    distribution of the training records (what predict_proba averages)."""
    names = [f"x{i}" for i in range(F)]; labels = [str(int(c)) for c in rf.classes_]
    out = _dd(names, "y", labels)
    out.append(f'<MiningModel functionName="classification">{_ms(names)}<Segmentation multipleModelMethod="average">')
    for k, est in enumerate(rf.estimators_):
        t = est.tree_

        def emit(i, pred):
            if t.children_left[i] < 0:
                v = t.value[i].ravel(); best = int(np.argmax(v))
                sd = "".join(f'<ScoreDistribution value="{labels[c]}" recordCount="{float(v[c])!r}"/>' for c in range(len(labels)))
                return f'<Node score="{labels[best]}">{pred}{sd}</Node>'
            f, thr = names[t.feature[i]], float(t.threshold[i])
            left = emit(t.children_left[i], f'<SimplePredicate field="{f}" operator="lessOrEqual" value="{thr!r}"/>')
            right = emit(t.children_right[i], f'<SimplePredicate field="{f}" operator="greaterThan" value="{thr!r}"/>')
            return f"<Node>{pred}{left}{right}</Node>"

        out.append(f'<Segment id="{k}"><True/><TreeModel functionName="classification">{_ms(names, None)}{emit(0, "<True/>")}</TreeModel></Segment>')
    out.append("</Segmentation></MiningModel></PMML>")
    return "\n".join(out).encode()


def sklearn_logreg_to_pmml(lr, F):
    """Multinomial LogisticRegression -> RegressionModel with one RegressionTable per class and the softmax link."""
    names = [f"x{i}" for i in range(F)]; labels = [str(int(c)) for c in lr.classes_]
    out = _dd(names, "y", labels)
    out.append(f'<RegressionModel functionName="classification" normalizationMethod="softmax">{_ms(names)}')
    for c, lab in enumerate(labels):
        out.append(f'<RegressionTable intercept="{float(lr.intercept_[c])!r}" targetCategory="{lab}">'
                   + "".join(f'<NumericPredictor name="{n}" coefficient="{float(lr.coef_[c, i])!r}"/>' for i, n in enumerate(names)) + '</RegressionTable>')
    out.append("</RegressionModel></PMML>")
    return "\n".join(out).encode()


def sklearn_mlp_to_pmml(mlp, F):
    """MLPClassifier (3+ classes: softmax output) -> NeuralNetwork; hidden activation relu / tanh / logistic."""
    act = {"relu": "rectifier", "tanh": "tanh", "logistic": "logistic", "identity": "identity"}[mlp.activation]
    assert mlp.out_activation_ == "softmax"
    names = [f"x{i}" for i in range(F)]; labels = [str(int(c)) for c in mlp.classes_]
    out = _dd(names, "y", labels)
    out.append(f'<NeuralNetwork functionName="classification" activationFunction="{act}">{_ms(names)}<NeuralInputs numberOfInputs="{F}">')
    out += [f'<NeuralInput id="0,{i}"><DerivedField optype="continuous" dataType="double"><FieldRef field="{n}"/></DerivedField></NeuralInput>' for i, n in enumerate(names)]
    out.append("</NeuralInputs>")
    L = len(mlp.coefs_)
    for l, (W, b) in enumerate(zip(mlp.coefs_, mlp.intercepts_), 1):
        attr = ' activationFunction="identity" normalizationMethod="softmax"' if l == L else ""
        out.append(f'<NeuralLayer numberOfNeurons="{W.shape[1]}"{attr}>')
        for j in range(W.shape[1]):
            out.append(f'<Neuron id="{l},{j}" bias="{float(b[j])!r}">' + "".join(f'<Con from="{l - 1},{i}" weight="{float(W[i, j])!r}"/>' for i in range(W.shape[0])) + "</Neuron>")
        out.append("</NeuralLayer>")
    out.append(f'<NeuralOutputs numberOfOutputs="{len(labels)}">')
    out += [f'<NeuralOutput outputNeuron="{L},{j}"><DerivedField optype="categorical" dataType="string"><NormDiscrete field="y" value="{lab}"/></DerivedField></NeuralOutput>'
            for j, lab in enumerate(labels)]
    out.append("</NeuralOutputs></NeuralNetwork></PMML>")
    return "\n".join(out).encode()


def sklearn_suite():
    """(name, pmml, X, labels as doubles, probabilities or None) for models FITTED BY scikit-learn: arithmetic the builders
    of this repository did not write.  Inputs are float32-representable so that scikit-learn's float32 tree inputs and the
    PMML's double fields see the same numbers."""
    from sklearn.ensemble import RandomForestClassifier
    from sklearn.linear_model import LogisticRegression
    from sklearn.neural_network import MLPClassifier
    import warnings
    rng = np.random.default_rng(11)
    F, C = 12, 4
    X = rng.standard_normal((4000, F)).astype(np.float32).astype(np.float64)
    Wt = rng.standard_normal((F, C)); y = np.argmax(X @ Wt + rng.normal(0, 1.0, (4000, C)), axis=1)
    Xq = rng.standard_normal((6000, F)).astype(np.float32).astype(np.float64)
    out = []
    rf = RandomForestClassifier(n_estimators=25, max_depth=7, random_state=0).fit(X, y)
    out.append(("rf-predict_proba", sklearn_rf_to_pmml(rf, F), Xq, rf.predict_proba(Xq)))
    lr = LogisticRegression(max_iter=300).fit(X, y)
    out.append(("multinomial-logreg", sklearn_logreg_to_pmml(lr, F), Xq, lr.predict_proba(Xq)))
    for act in ("relu", "tanh", "logistic"):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            mlp = MLPClassifier(hidden_layer_sizes=(24, 16), activation=act, max_iter=60, random_state=1).fit(X, y)
        out.append((f"mlp-{act}", sklearn_mlp_to_pmml(mlp, F), Xq, mlp.predict_proba(Xq)))
    return out


def check_against_sklearn(name, result, proba, has_probs):
    """labels equal sklearn's argmax wherever sklearn's own top-2 gap is not a rounding tie; probabilities within 1e-6."""
    assert (result["status"] == 0).all(), name
    srt = np.sort(proba, axis=1); clear = srt[:, -1] - srt[:, -2] > 1e-12
    assert clear.mean() > 0.99
    assert np.array_equal(result["score"][clear], np.argmax(proba, axis=1)[clear].astype(np.float64)), name
    if has_probs:
        np.testing.assert_allclose(result["probs"], proba, rtol=1e-6, atol=1e-12, err_msg=name)
