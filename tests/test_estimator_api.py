"""Estimator surface, sklearn plumbing and the documented quirk policy."""
import pickle

import numpy as np
import pytest
from sklearn.base import clone, is_classifier

from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier


def test_ctor_keyword_only_and_params():
    with pytest.raises(TypeError):
        DecisionTreeClassifier(3)
    clf = DecisionTreeClassifier(max_depth=3, min_samples_split=4)
    p = clf.get_params()
    assert p["max_depth"] == 3 and p["min_samples_split"] == 4
    assert repr(DecisionTreeClassifier()) == "DecisionTreeClassifier()"
    clf.set_params(max_depth=7)
    assert clone(clf).max_depth == 7
    assert is_classifier(clf)


def test_fitted_attributes(iris):
    clf = DecisionTreeClassifier(max_depth=2, device="cpu").fit(iris.data, iris.target)
    assert clf.n_features_ == 4 and list(clf.classes_) == [0, 1, 2]
    assert clf.tree_.count.dtype == np.int64 and isinstance(clf.tree_.threshold, float)
    assert isinstance(clf.tree_.value, np.int64)
    proba = clf.predict_proba(iris.data)
    assert proba.shape == (150, 3) and proba.dtype == np.int64 and proba.sum(1).min() > 1
    assert clf.predict(iris.data).dtype == np.int64
    assert clf.get_depth() == 2 and clf.get_n_leaves() == 3
    clone_ = pickle.loads(pickle.dumps(clf))
    assert np.array_equal(clone_.predict(iris.data), clf.predict(iris.data))


def test_unfitted_raises():
    from sklearn.exceptions import NotFittedError

    with pytest.raises(NotFittedError):
        DecisionTreeClassifier().predict([[0.0]])
    with pytest.raises(NotFittedError):
        DecisionTreeClassifier().export_text()


def test_label_validation():
    X = np.arange(6, dtype=float).reshape(-1, 1)
    with pytest.raises(TypeError):
        DecisionTreeClassifier(device="cpu").fit(X, np.array([0.5, 1, 0, 1, 0, 1]))
    with pytest.raises(ValueError):
        DecisionTreeClassifier(device="cpu").fit(X, np.array([-1, 1, 0, 1, 0, 1]))
    with pytest.raises((TypeError, ValueError)):
        DecisionTreeClassifier(device="cpu").fit(X, np.array(list("ababab")))


def test_nan_rejected():
    X = np.array([[0.0], [np.nan], [1.0]])
    with pytest.raises(ValueError, match="NaN"):
        DecisionTreeClassifier(device="cpu").fit(X, [0, 1, 0])


def test_constant_column_defect_is_fixed():
    # reference: RecursionError / IndexError (constant feature 0 chosen at zero gain)
    X = np.array([[1.0, 0.0], [1.0, 1.0], [1.0, 0.0], [1.0, 1.0]])
    y = np.array([0, 1, 1, 0])
    clf = DecisionTreeClassifier(device="cpu").fit(X, y)
    assert int(clf.tree_.value) == 1  # only feature 1 has two distinct values
    assert all(n.count.sum() > 0 for n in clf.tree_.iter_preorder())


def test_all_rows_identical_is_leaf():
    X = np.ones((5, 3))
    y = np.array([0, 1, 0, 1, 1])
    clf = DecisionTreeClassifier(device="cpu").fit(X, y)
    assert clf.tree_.is_leaf and int(clf.tree_.value) == 1
    assert clf.predict(X).tolist() == [1] * 5


def test_gappy_labels_predict():
    X = np.arange(8, dtype=float).reshape(-1, 1)
    y = np.array([0, 0, 5, 5, 0, 0, 5, 5])
    clf = DecisionTreeClassifier(device="cpu").fit(X, y)
    assert clf.predict(X).tolist() == y.tolist()
    assert clf.predict_proba(X).shape == (8, 6)


def test_leaf_tie_lowest_label():
    X = np.zeros((4, 1))
    clf = DecisionTreeClassifier(device="cpu").fit(X, [2, 1, 1, 2])
    assert int(clf.tree_.value) == 1


def test_gini_extension(iris):
    clf = DecisionTreeClassifier(criterion="gini", max_depth=3, device="cpu").fit(iris.data, iris.target)
    assert clf.score(iris.data, iris.target) > 0.95
    with pytest.raises(ValueError):
        DecisionTreeClassifier(criterion="mse", device="cpu").fit(iris.data, iris.target)


def test_quantile_binning_many_values():
    rng = np.random.default_rng(0)
    X = rng.normal(size=(5000, 3))
    y = (X[:, 1] > 0.25).astype(int)
    clf = DecisionTreeClassifier(max_depth=3, max_bins=64, device="cpu").fit(X, y)
    assert int(clf.tree_.value) == 1 and abs(clf.tree_.threshold - 0.25) < 0.1
    assert clf.tree_.threshold in X[:, 1]  # a data value, never a midpoint
    assert clf.score(X, y) > 0.97


def test_world_attrs_serial():
    clf = ParallelDecisionTreeClassifier()
    assert clf.WORLD_RANK == 0 and clf.WORLD_SIZE == 1
    assert ParallelDecisionTreeClassifier.WORLD_COMM.Get_size() == 1
    sub = clf._get_cyclic_dist(clf.WORLD_COMM, n_blocks=2)
    assert sub.size == 1


def test_parallel_serial_world_equals_serial(iris):
    a = ParallelDecisionTreeClassifier(max_depth=3, device="cpu").fit(iris.data, iris.target)
    b = DecisionTreeClassifier(max_depth=3, device="cpu").fit(iris.data, iris.target)
    assert a.export_text() == b.export_text()
    assert "comm" in a.get_params() and pickle.loads(pickle.dumps(a)).export_text() == b.export_text()


def test_predict_rejects_wrong_feature_count(iris):
    clf = DecisionTreeClassifier(max_depth=2, device="cpu").fit(iris.data, iris.target)
    with pytest.raises(ValueError, match="expecting 4 features"):
        clf.predict(iris.data[:, :3])
    with pytest.raises(ValueError, match="expecting 4 features"):
        clf.predict_proba(np.hstack([iris.data, iris.data[:, :1]]))


def test_zero_edge_of_a_quantile_binned_column_without_zero_rows():
    """More distinct values than bins, both signed zeros present: a node can split at the edge 0 without
    holding a zero itself (this used to raise StopIteration while restoring the printed sign)."""
    rng = np.random.default_rng(0)
    X = rng.normal(size=(3000, 8)).round(2)
    y = (X[:, 0] + X[:, 1] * X[:, 2] + rng.normal(scale=0.7, size=3000) > 0).astype(int) + (X[:, 3] > 1)
    assert np.signbit(X[X == 0]).any() and not np.signbit(X[X == 0]).all()
    clf = DecisionTreeClassifier(device="cpu").fit(X, y)
    assert clf.score(X, y) > 0.95
