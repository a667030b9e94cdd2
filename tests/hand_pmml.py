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
