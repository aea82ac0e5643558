"""Model checkpoints without pickle: hyper-parameters + flat tree + bin edges in one ``.npz``.

The reference has no explicit checkpointing; fitted estimators are plain-picklable
 and that still works here.  This module adds a portable format that can be
loaded without importing arbitrary classes and that round-trips the exact tree
(``export_text`` / ``predict`` identical).
"""

from __future__ import annotations

import io
import json
from typing import Union

import numpy as np

from ..models.flat_tree import FlatTree

__all__ = ["save_model", "load_model"]


def save_model(clf, path_or_file: Union[str, io.IOBase]) -> None:
    """Write a fitted estimator to ``path_or_file``."""
    flat = clf._flat()
    params = {k: v for k, v in clf.get_params().items() if k != "comm"}
    meta = {
        "class": type(clf).__name__,
        "params": params,
        "n_features": int(clf.n_features_),
        "classes": np.asarray(clf.classes_).tolist(),
        "engine": getattr(clf, "engine_", "cpu"),
    }
    np.savez_compressed(
        path_or_file,
        meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8),
        tree=np.frombuffer(flat.to_bytes(), dtype=np.uint8),
    )


def load_model(path_or_file):
    """Load an estimator written by :func:`save_model`."""
    from ..tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

    z = np.load(path_or_file, allow_pickle=False)
    meta = json.loads(bytes(z["meta"]).decode())
    cls = {"DecisionTreeClassifier": DecisionTreeClassifier,
           "ParallelDecisionTreeClassifier": ParallelDecisionTreeClassifier}[meta["class"]]
    clf = cls(**meta["params"])
    clf.n_features_ = clf.n_features_in_ = meta["n_features"]
    clf.classes_ = np.asarray(meta["classes"])
    clf.engine_ = meta["engine"]
    clf.flat_tree_ = FlatTree.from_bytes(bytes(z["tree"]))
    return clf
