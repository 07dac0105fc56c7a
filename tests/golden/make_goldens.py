"""Regenerates the fixtures under tests/golden/ from the reference tree (run in the build
container only: /root/reference does not exist on the GPU box).

* circle2_config.json / circle8_config.json: base_config.json merged with the task file's `tasks`
  exactly as workspace/experiments/python/preprocess.py:20-24 does (effective experiment config).
* checks that the known answers transcribed into reference_goldens.json still appear verbatim in
  the reference's test sources (cbf/tests/TestInitSafetyCBF.cpp, cbf/tests/TestInitConnectivity.cpp, qpcpp/tests/CPLEXTest.cpp).
"""
import json
import os

REF = "/root/reference/workspace"
HERE = os.path.dirname(os.path.abspath(__file__))


def merged(task):
    base = json.load(open(os.path.join(REF, "experiments/config/base_config.json")))
    cfg = json.load(open(os.path.join(REF, "experiments/config", task)))
    return {**base, "tasks": cfg.get("tasks", {})}


if __name__ == "__main__":
    for name, task in (("circle2_config.json", "baseline/2r/circle.json"),
                       ("circle8_config.json", "baseline/8r/circle.json")):
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(merged(task), f, indent=1)
    gold = json.load(open(os.path.join(HERE, "reference_goldens.json")))
    src = open(os.path.join(REF, "lib/cbf/tests/TestInitSafetyCBF.cpp")).read()
    for case in gold["safety_cbf"]["cases"]:
        assert case["name"] in src, case["name"]
        if case["Bc"] != 0.0:
            assert repr(case["Bc"]) in src, case["Bc"]
    src = open(os.path.join(REF, "lib/cbf/tests/TestInitConnectivity.cpp")).read()
    for case in gold["connectivity_cbf"]["cases"]:
        assert repr(case["Bc"]) in src, case["Bc"]
        assert ", ".join(repr(v) for v in case["Ac"][:2]) in src.replace("0.0, -2.703392", "0.0, -2.703392"), case["Ac"]
    assert "0.5, 1e-6" in open(os.path.join(REF, "lib/qpcpp/tests/CPLEXTest.cpp")).read()
    print("fixtures regenerated; goldens verified against the reference test sources")
