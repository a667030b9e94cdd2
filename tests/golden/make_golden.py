#!/usr/bin/env python3
"""Regenerates tests/golden/ from the read-only reference checkout (run in the build container only).

Inputs  : /root/reference/flink-jpmml-assets/src/main/resources/kmeans*.xml  (the reference's only fixtures)
Outputs : tests/golden/iris_*.pmml   - the same documents re-serialised (comments dropped, one line, attribute
                                        and element order kept) so that the parity tests can run on the GPU box,
                                        where /root/reference does not exist;
          tests/golden/golden_vectors.json - the known-answer vectors the reference's own tests assert, each with
                                        the spec file:line it was read from.

Nothing here is executed by the test-suite; the committed outputs are.
"""
import json
import os
import re
import xml.etree.ElementTree as ET

SRC = "/root/reference/flink-jpmml-assets/src/main/resources"
HERE = os.path.dirname(os.path.abspath(__file__))
T = "flink-jpmml-scala/src/test/scala/io/radicalbit/flink/pmml/scala"

NAMES = {
    "kmeans.xml": "iris_kmeans43.pmml",
    "kmeans41.xml": "iris_kmeans41.pmml",
    "kmeans40.xml": "iris_kmeans40.pmml",
    "kmeans42.xml": "iris_kmeans42.pmml",
    "kmeans32.xml": "iris_kmeans32.pmml",
    "kmeans_nooutput.xml": "iris_kmeans_nooutput.pmml",
    "kmeans_nooutput_notarget.xml": "iris_kmeans_nooutput_notarget.pmml",
    "kmeans_stringfields.xml": "iris_kmeans_stringfields.pmml",
    "kmeans_empty.xml": "iris_kmeans_empty.pmml",
}


def compact(text: str) -> str:
    text = re.sub(r"<!--.*?-->", "", text, flags=re.S).strip()
    if not text:
        return "<!-- empty document: the reference's kmeans_empty.xml holds no root element -->\n"
    for prefix, uri in re.findall(r'xmlns:?(\w*)="([^"]+)"', text):
        ET.register_namespace(prefix, uri)
    root = ET.fromstring(text)
    for el in root.iter():
        if el.text is not None and not el.text.strip():
            el.text = None
        if el.tail is not None and not el.tail.strip():
            el.tail = None
    return ET.tostring(root, encoding="unicode") + "\n"


def main():
    for src, dst in NAMES.items():
        with open(os.path.join(SRC, src)) as f:
            out = compact(f.read())
        with open(os.path.join(HERE, dst), "w") as f:
            f.write(out)
    fields = ["sepal_length", "sepal_width", "petal_length", "petal_width"]
    vectors = {
        "active_fields": {"value": fields, "cite": f"{T}/api/converter/VectorConverterSpec.scala:76-78"},
        "predict": [
            {"model": "iris_kmeans43.pmml", "kind": "dense", "values": [1.0, 1.0, 1.0, 1.0], "replace": None,
             "score": 3.0, "cite": f"{T}/api/PmmlModelSpec.scala:50-55"},
            {"model": "iris_kmeans43.pmml", "kind": "sparse", "size": 4, "indices": [0, 1, 2, 3],
             "values": [1.0, 2.0, 3.0, 4.0], "replace": None, "score": 4.0,
             "cite": f"{T}/api/PmmlModelSpec.scala:57-62"},
            {"model": "iris_kmeans43.pmml", "kind": "sparse", "size": 4, "indices": [0, 2], "values": [1.0, 2.0],
             "replace": 0.0, "score": 3.0, "cite": f"{T}/api/PmmlModelSpec.scala:64-69"},
            {"model": "iris_kmeans43.pmml", "kind": "sparse", "size": 4, "indices": [0, 2], "values": [1.0, 2.0],
             "replace": None, "score": 3.0, "cite": f"{T}/api/PmmlModelSpec.scala:71-76"},
            {"model": "iris_kmeans43.pmml", "kind": "dense", "values": [1.0, 2.0, 3.0], "replace": None,
             "score": None, "cite": f"{T}/api/PmmlModelSpec.scala:78-83"},
            {"model": "iris_kmeans_nooutput.pmml", "kind": "dense", "values": [1.0, 1.0, 1.0, 1.0], "replace": None,
             "score": 3.0, "cite": f"{T}/RichConnectedStreamSpec.scala (default model) / PmmlModelSpec.scala:40-46"},
            {"model": "iris_kmeans_nooutput_notarget.pmml", "kind": "dense", "values": [1.0, 1.0, 1.0, 1.0],
             "replace": None, "score": None, "cite": f"{T}/RichConnectedStreamSpec.scala:97-116"},
            {"model": "iris_kmeans_stringfields.pmml", "kind": "dense", "values": [1.0, 4.0, -1.0, 3.0],
             "replace": None, "score": None, "cite": f"{T}/api/PmmlModelSpec.scala:131-136"},
        ],
        "loadable": {"value": ["iris_kmeans43.pmml", "iris_kmeans42.pmml", "iris_kmeans41.pmml", "iris_kmeans40.pmml"],
                     "cite": f"{T}/api/PmmlModelSpec.scala:329-347"},
        "not_loadable": {"value": ["iris_kmeans_empty.pmml"], "cite": f"{T}/RichDataStreamSpec.scala:98-105"},
        "model_id": {"uuid": "2bc91c20-bf64-4860-ba11-de97a95d05cf", "version": 23, "bad": "23441-_23",
                     "cite": f"{T}/models/core/ModelIdSpec.scala:26-52"},
    }
    with open(os.path.join(HERE, "golden_vectors.json"), "w") as f:
        json.dump(vectors, f, indent=1)
        f.write("\n")


if __name__ == "__main__":
    main()
