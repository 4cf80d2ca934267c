"""What the reference's /predict handler does per request, restated for timing on a host where
/root/reference does not exist (the GPU box): unpickle the model file, score the one row twice through
scikit-learn (`predict`, then `predict_proba(...).max()`), on the event loop, one uvicorn worker
(REF/main.py:16-27, REF/README.md:16).  Not the reference's file: tools/loadtest.py times the unmodified
REF/main.py wherever it is present (this container) and this restatement beside it, so the two can be compared.

    uvicorn ref_handler_app:app        (cwd must hold LRClassifier.pkl)
"""
import pickle

from fastapi import FastAPI
from pydantic import BaseModel

app = FastAPI()


class Iris(BaseModel):
    sepal_length: float
    sepal_width: float
    petal_length: float
    petal_width: float


@app.post("/predict")
async def predict(iris: Iris):
    fields = iris.dict()                                 # the reference goes through the (deprecated, warning-raising) dict view
    with open("LRClassifier.pkl", "rb") as fh:           # every request, as the reference does
        model = pickle.load(fh)
    row = [[fields["sepal_length"], fields["sepal_width"], fields["petal_length"], fields["petal_width"]]]
    return {"prediction": model.predict(row)[0], "probability": model.predict_proba(row).max()}
