"""Test-only stand-in for xgboost==0.7.post4 gblinear (requirements.txt:38; predict.py:150-166):
Booster.predict = base_score + bias + w.x (dump layout: predict_by_cluster.py:73-75).  ``load_model`` reads an
``.npz`` with ``weights`` [n_features], ``bias`` and ``base_score``."""
import numpy as np


class DMatrix:
    def __init__(self, data, **kw):
        self.data = np.asarray(data)


class Booster:
    def __init__(self, params=None, **kw):
        self.w = None

    def load_model(self, path):
        z = np.load(path)
        self.w, self.b, self.base = z["weights"], float(z["bias"]), float(z["base_score"])

    def predict(self, d):
        return (self.base + self.b + d.data @ self.w).astype(np.float32)
