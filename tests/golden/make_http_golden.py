"""Record what the UNMODIFIED reference app answers (REF/main.py under fastapi's TestClient, with
the notebook's pickle regenerated offline) for a fixed set of /predict requests.  Run in this
container (needs /root/reference); writes tests/golden/http_golden.json.

    python tests/golden/make_http_golden.py
"""
import importlib.util
import json
import os
import pickle
import shutil
import sys
import tempfile
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import synth  # noqa: E402

REQUESTS = [
    {"sepal_length": 5.1, "sepal_width": 3.5, "petal_length": 1.4, "petal_width": 0.2},
    {"sepal_length": 6.0, "sepal_width": 2.9, "petal_length": 4.5, "petal_width": 1.5},
    {"sepal_length": 6.3, "sepal_width": 3.3, "petal_length": 6.0, "petal_width": 2.5},
    {"sepal_length": 5.9, "sepal_width": 3.2, "petal_length": 4.8, "petal_width": 1.8},
    {"sepal_length": "6.0", "sepal_width": 2.2, "petal_length": 5.0, "petal_width": 1.5},   # str coerced
    {"sepal_length": 7, "sepal_width": 3, "petal_length": 5, "petal_width": 2, "extra": "ignored"},
    {"sepal_length": 0.0, "sepal_width": 0.0, "petal_length": 0.0, "petal_width": 0.0},
    {"sepal_length": 100.0, "sepal_width": -50.0, "petal_length": 30.0, "petal_width": 9.0},
    {"sepal_length": 5.1, "sepal_width": 3.5, "petal_length": 1.4},                          # missing -> 422
    {"sepal_length": "abc", "sepal_width": 3.5, "petal_length": 1.4, "petal_width": 0.2},    # 422
    {},                                                                                        # 422
]


def main():
    warnings.simplefilter("ignore")
    from fastapi.testclient import TestClient

    clf, _ = synth.fit_notebook_model()
    tmp = tempfile.mkdtemp()
    shutil.copy("/root/reference/main.py", os.path.join(tmp, "main.py"))
    with open(os.path.join(tmp, "LRClassifier.pkl"), "wb") as fh:
        pickle.dump(clf, fh)
    os.chdir(tmp)
    spec = importlib.util.spec_from_file_location("ref_main", os.path.join(tmp, "main.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    client = TestClient(mod.app)
    out = []
    for body in REQUESTS:
        r = client.post("/predict", json=body)
        out.append({"request": body, "status": r.status_code, "json": r.json()})
    routes = sorted((r.path, sorted(r.methods)) for r in mod.app.routes if hasattr(r, "methods"))
    with open(os.path.join(HERE, "http_golden.json"), "w") as fh:
        json.dump({"responses": out, "routes": routes}, fh, indent=1)
    print(json.dumps(out[:2], indent=1))
    print(routes)


if __name__ == "__main__":
    main()
