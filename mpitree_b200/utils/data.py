"""Seeded synthetic data sets of the BASELINE.json shapes (no files, no network).

``synthetic_codes`` yields pre-quantised ``uint8`` features (256 bins) with labels that depend
on a handful of informative features plus noise, so trees have real structure instead of
fitting pure noise.  Works with numpy (host) or torch (device) generators.
"""

from __future__ import annotations

from typing import Tuple

import numpy as np

__all__ = ["synthetic_codes", "synthetic_continuous", "synthetic_mixed", "notebook_workload"]

_W = np.array([1.0, -0.8, 0.6, 0.5, -0.4, 0.3, 0.25, -0.2])


def synthetic_codes(n: int, n_features: int = 64, n_classes: int = 10, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """``(X uint8 (n, F) in [0, 256), y int64 (n,))`` — the 100M x 64 / 256 bins / 10 classes family."""
    rng = np.random.default_rng(seed)
    X = rng.integers(0, 256, size=(n, n_features), dtype=np.uint8)
    k = min(len(_W), n_features)
    score = (X[:, :k].astype(np.float64) - 127.5) @ _W[:k] + 40.0 * rng.normal(size=n)
    qs = np.quantile(score, np.linspace(0, 1, n_classes + 1)[1:-1])
    return X, np.digitize(score, qs).astype(np.int64)


def synthetic_continuous(n: int, n_features: int = 32, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Two-class continuous problem (the 10M x 32 family): float32 normal features."""
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, n_features)).astype(np.float32)
    s = X[:, 0] + 0.5 * X[:, 5 % n_features] * X[:, 9 % n_features] - 0.3 * X[:, 17 % n_features]
    y = (s + 0.8 * rng.normal(size=n) > 0).astype(np.int64)
    return X, y


def synthetic_mixed(n: int, n_continuous: int = 64, n_categorical: int = 64, seed: int = 0):
    """Mixed categorical + continuous problem (the 50M x 128 family) as an object-free float matrix:
    continuous columns are floats, categorical columns are small integer level codes."""
    rng = np.random.default_rng(seed)
    Xc = rng.normal(size=(n, n_continuous)).astype(np.float32)
    levels = [4 + (j % 16) * 4 for j in range(n_categorical)]
    Xk = np.stack([rng.integers(0, L, size=n) for L in levels], axis=1).astype(np.float32)
    X = np.concatenate([Xc, Xk], axis=1)
    c = lambda j: Xc[:, j % n_continuous]  # noqa: E731
    k = lambda j: Xk[:, j % n_categorical]  # noqa: E731
    s = c(3) + 1.5 * (k(6) % 3 == 0) - (k(26) > 10) + 2.0 * ((c(10) > 0.5) & (k(36) < 5))
    y = np.digitize(s + 1.2 * rng.normal(size=n), [-1.0, 0.0, 1.0, 2.0]).astype(np.int64)
    return X, y


def notebook_workload(n: int):
    """The reference notebook's timing workload: ``X = arange(n)``, ``y = arange(n)`` (``experiments.ipynb:198-209``)."""
    return np.arange(n).reshape(-1, 1), np.arange(n)
