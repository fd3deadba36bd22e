"""Loader for xgboost ``gblinear`` expression models (predict.py:165-166 ``xgb.Booster().load_model``).

xgboost==0.7.post4 is a pinned third-party dependency of the reference (requirements.txt:38) that is not in
this image, and the reference's ``models/*.save`` files are not in its repository, so nothing here can be
checked against a real model file: the binary layout below restates xgboost 0.7's published serialisation and is
pinned only as far as that goes offline — tests/test_predict_host_cpu.py builds a file byte by byte from the
struct definitions, written down there with their source locations:

    src/learner.cc   LearnerImpl::Save   = LearnerModelParam (136 bytes: float base_score, unsigned num_feature,
                                           int num_class, int contain_extra_attrs, int contain_eval_metrics,
                                           int reserved[29]) + dmlc string name_obj_ + dmlc string name_gbm_ +
                                           gbm_->Save + optional attribute / metric vectors
                     LearnerImpl::Load     accepts an optional 4-byte "binf" prefix
    src/gbm/gblinear.cc  GBLinearModel::Save = GBLinearModelParam (136 bytes: unsigned num_feature,
                                           int num_output_group, int reserved[32]) + dmlc vector<float> weight,
                                           weight[i * num_output_group + gid], biases after the num_feature rows
    dmlc-core serializer: string / vector = uint64 length + raw little-endian elements

The text-dump form (``bst.dump_model``, train.py:158) is pinned against the reference's own parser of it
(predict_by_cluster.py:73-75); the ``.npz`` form is what the oracle's xgboost shim reads (oracle/shims/xgboost.py).

A gblinear model scores ``base_score + bias + w . x`` (dump layout: predict_by_cluster.py:73-75).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np


@dataclass
class GBLinearModel:
    weights: np.ndarray  # float32 [n_features]
    bias: float
    base_score: float = 0.5

    @property
    def offset(self) -> float:
        return float(self.base_score) + float(self.bias)


def load_npz(path: str) -> GBLinearModel:
    z = np.load(path)
    return GBLinearModel(np.ascontiguousarray(z["weights"], dtype=np.float32), float(z["bias"]),
                         float(z["base_score"]) if "base_score" in z.files else 0.5)


def load_text_dump(path: str, base_score: float = 0.5) -> GBLinearModel:
    """``bst.dump_model(path)`` of a gblinear booster: ``booster[0]:`` / ``bias:`` / value / ``weight:`` / values."""
    bias, weights, mode = [], [], None
    for line in open(path):
        t = line.strip()
        if not t or t.startswith("booster"):
            continue
        if t.startswith("bias"):
            mode = bias
        elif t.startswith("weight"):
            mode = weights
        elif mode is not None:
            mode.append(float(t))
    if len(bias) != 1 or not weights:
        raise ValueError(f"{path}: not a single-output gblinear dump")
    return GBLinearModel(np.asarray(weights, dtype=np.float32), bias[0], base_score)


def save_text_dump(path: str, model: GBLinearModel) -> None:
    """``bst.dump_model(path)`` of a gblinear booster (train.py:158): ``booster[0]:`` then what ``get_dump()[0]``
    returns — ``bias:`` / value / ``weight:`` / one value per line (the layout predict_by_cluster.py:73-75 parses:
    ``dump[1]`` is the bias, ``dump[3:]`` the weights).  Values are written with 9 significant digits, enough to
    round-trip float32."""
    with open(path, "w") as f:
        f.write("booster[0]:\nbias:\n%.9g\nweight:\n" % float(np.float32(model.bias)))
        f.write("\n".join("%.9g" % float(w) for w in np.asarray(model.weights, dtype=np.float32)))
        f.write("\n")


def load_binary(path: str) -> GBLinearModel:
    """xgboost 0.7 binary model (layout: module docstring; no real model file exists offline to pin it against)."""
    b = open(path, "rb").read()
    off = 4 if b[:4] == b"binf" else 0
    base_score, num_feature, num_class, extra_attrs, eval_metrics = struct.unpack_from("<fIiii", b, off)
    off += 136  # LearnerModelParam: 5 fields + int reserved[29]
    names = []
    for _ in range(2):  # objective name, booster name
        (ln,) = struct.unpack_from("<Q", b, off)
        names.append(b[off + 8:off + 8 + ln].decode("ascii", "replace"))
        off += 8 + ln
    if names[1] != "gblinear":
        raise ValueError(f"{path}: booster is {names[1]!r}, not gblinear")
    gb_features, groups = struct.unpack_from("<Ii", b, off)
    off += 136  # GBLinearModelParam: 2 fields + int reserved[32]
    (n,) = struct.unpack_from("<Q", b, off)
    off += 8
    if groups != 1 or n != (gb_features + 1) * groups:
        raise ValueError(f"{path}: unexpected gblinear shape (features {gb_features}, groups {groups}, weights {n})")
    w = np.frombuffer(b, dtype="<f4", count=n, offset=off)
    return GBLinearModel(np.ascontiguousarray(w[:gb_features]), float(w[gb_features]), float(base_score))


def save_binary(path: str, model: GBLinearModel, objective: str = "reg:linear") -> None:
    """Writer of the same layout (round-trip tests, synthetic ``modellist`` files)."""
    nf = int(model.weights.size)
    out = struct.pack("<fIiii", model.base_score, nf, 0, 0, 0) + b"\x00" * (136 - 20)
    for name in (objective, "gblinear"):
        out += struct.pack("<Q", len(name)) + name.encode("ascii")
    out += struct.pack("<Ii", nf, 1) + b"\x00" * (136 - 8)
    out += struct.pack("<Q", nf + 1) + np.asarray(model.weights, dtype="<f4").tobytes() + struct.pack("<f", model.bias)
    open(path, "wb").write(out)


def load_model(path: str) -> GBLinearModel:
    path = path.strip()
    if path.endswith(".npz"):
        return load_npz(path)
    head = open(path, "rb").read(16)
    if head.startswith(b"booster") or head.startswith(b"bias"):
        return load_text_dump(path)
    return load_binary(path)
