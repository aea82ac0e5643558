"""Exact-tree parity against the unmodified reference (the oracle), CPU engine."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from mpitree_b200.tree import DecisionTreeClassifier


def _trees_equal(a, b):
    """Compare our Node graph with the reference's Node graph field by field."""
    stack = [(a, b)]
    while stack:
        x, y = stack.pop()
        assert x.is_leaf == y.is_leaf
        assert int(x.value) == int(y.value)
        assert x.depth == y.depth
        mine, theirs = np.asarray(x.count), np.asarray(y.count)
        # gappy labels: the reference's count has a node-dependent length (defect);
        # ours is zero-padded to max(label) + 1
        assert np.array_equal(mine[: len(theirs)], theirs) and not mine[len(theirs):].any()
        if not x.is_leaf:
            assert x.threshold == y.threshold
            assert type(x.threshold) is type(y.threshold)
            stack.append((x.left, y.left))
            stack.append((x.right, y.right))


@pytest.mark.parametrize("cols", [slice(0, 2), slice(0, 4), slice(2, 4)])
@pytest.mark.parametrize("depth", [0, 1, 3, 5, None])
def test_iris_matches_reference(ref, iris, cols, depth):
    X, y = iris.data[:, cols], iris.target
    ours = DecisionTreeClassifier(max_depth=depth, device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier(max_depth=depth).fit(X, y)
    assert ours.export_text() == theirs.export_text()
    _trees_equal(ours.tree_, theirs.tree_)
    assert np.array_equal(ours.predict(X), theirs.predict(X).astype(np.int64))
    assert np.array_equal(ours.predict_proba(X), theirs.predict_proba(X).astype(np.int64))
    assert ours.score(X, y) == theirs.score(X, y)
    assert ours.n_features_ == theirs.n_features_
    assert np.array_equal(ours.classes_, theirs.classes_)


def test_iris_scores(iris):
    y = iris.target
    assert round(DecisionTreeClassifier(max_depth=3, device="cpu").fit(iris.data[:, :2], y).score(iris.data[:, :2], y), 4) == 0.8067
    assert round(DecisionTreeClassifier(max_depth=3, device="cpu").fit(iris.data, y).score(iris.data, y), 4) == 0.9733
    root = DecisionTreeClassifier(max_depth=3, device="cpu").fit(iris.data, y).tree_
    assert (int(root.value), root.threshold) == (2, 1.9)


@pytest.mark.parametrize("mss", [2, 5, 20, 60])
def test_min_samples_split_matches_reference(ref, iris, mss):
    X, y = iris.data, iris.target
    ours = DecisionTreeClassifier(min_samples_split=mss, device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier(min_samples_split=mss).fit(X, y)
    assert ours.export_text() == theirs.export_text()


def test_xor_zero_gain_split(ref):
    X = np.array([[0, 0], [0, 1], [1, 0], [1, 1]] * 3, dtype=float)
    y = np.array([0, 1, 1, 0] * 3)
    ours = DecisionTreeClassifier(device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier().fit(X, y)
    assert ours.export_text() == theirs.export_text()
    assert int(ours.tree_.value) == 0 and ours.tree_.threshold == 0.0


def test_notebook_timing_workload(ref):
    # experiments.ipynb:198-209: n distinct classes, one feature
    for n in (1, 11, 41):
        X = np.arange(n).reshape(-1, 1)
        y = np.arange(n)
        ours = DecisionTreeClassifier(device="cpu").fit(X, y)
        theirs = ref.DecisionTreeClassifier().fit(X, y)
        assert ours.export_text() == theirs.export_text()


@st.composite
def small_dataset(draw):
    n = draw(st.integers(2, 40))
    f = draw(st.integers(1, 4))
    k = draw(st.integers(2, 4))
    vals = draw(st.integers(2, 6))
    X = draw(st.lists(st.lists(st.integers(0, vals - 1), min_size=f, max_size=f), min_size=n, max_size=n))
    y = draw(st.lists(st.integers(0, k - 1), min_size=n, max_size=n))
    return np.array(X, dtype=float), np.array(y)


@settings(max_examples=120, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(data=small_dataset(), depth=st.sampled_from([1, 2, 3, 6]), mss=st.sampled_from([2, 2, 4]))
def test_random_integer_datasets_match_reference(ref, data, depth, mss):
    """Tie-heavy tiny data sets: every successful reference fit must be reproduced exactly."""
    X, y = data
    try:
        theirs = ref.DecisionTreeClassifier(max_depth=depth, min_samples_split=mss).fit(X, y)
    except (IndexError, RecursionError):
        # reference defect (constant column chosen at zero gain): ours must still fit
        DecisionTreeClassifier(max_depth=depth, min_samples_split=mss, device="cpu").fit(X, y)
        return
    ours = DecisionTreeClassifier(max_depth=depth, min_samples_split=mss, device="cpu").fit(X, y)
    assert ours.export_text() == theirs.export_text()
    _trees_equal(ours.tree_, theirs.tree_)
    try:
        expected = theirs.predict(X)
    except ValueError:  # reference defect with gappy labels
        return
    assert np.array_equal(ours.predict(X), expected)


@settings(max_examples=25, deadline=None, derandomize=True, suppress_health_check=list(HealthCheck))
@given(seed=st.integers(0, 10_000), n=st.integers(20, 120), k=st.integers(2, 9))
def test_random_continuous_match_reference(ref, seed, n, k):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, 3)).round(2)  # rounding leaves negative zeros in: their printed sign must match too
    y = rng.integers(0, k, size=n)
    y[:k] = np.arange(k)  # contiguous labels so the reference's predict works
    ours = DecisionTreeClassifier(max_depth=4, device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier(max_depth=4).fit(X, y)
    assert ours.export_text() == theirs.export_text()
    assert np.array_equal(ours.predict_proba(X), theirs.predict_proba(X))


@pytest.mark.parametrize("seed", [8, 9, 24, 1284, 100023])
def test_signed_zero_thresholds_print_like_the_reference(ref, seed):
    """A column with both -0.0 and 0.0: the reference prints whichever its per-node np.unique keeps."""
    rng = np.random.default_rng(seed)
    n = int(rng.integers(40, 120))
    X = rng.normal(size=(n, 3)).round(1)
    X[rng.random((n, 3)) < 0.15] = 0.0
    X[rng.random((n, 3)) < 0.15] = -0.0
    y = rng.integers(0, 3, size=n)
    ours = DecisionTreeClassifier(device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier().fit(X, y)
    assert "0.00]" in theirs.export_text()
    assert ours.export_text() == theirs.export_text()
    assert ours.export_text(precision=1) == theirs.export_text(precision=1)


def test_object_arrays_keep_the_reference_semantics(ref):
    """Numeric-looking strings are ordered as text, object ints keep int thresholds (found by a random sweep)."""
    X = np.array([["10"], ["9"], ["100"], ["9"], ["10"], ["8"]], dtype=object)
    y = np.array([0, 1, 0, 1, 0, 1])
    ours = DecisionTreeClassifier(device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier().fit(X, y)
    assert ours.tree_.threshold == theirs.tree_.threshold == "100"
    assert np.array_equal(ours.predict(X), theirs.predict(X))
    rng = np.random.default_rng(3)
    X = rng.integers(0, 4, size=(40, 2)).astype(object)
    y = rng.integers(0, 3, size=40)
    ours = DecisionTreeClassifier(device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier().fit(X, y)
    assert ours.export_text() == theirs.export_text()
    _trees_equal(ours.tree_, theirs.tree_)  # includes type(threshold) is int
    mixed = np.array([[1, 2.5], [2, 0.5], [3, 1.5], [1, 2.0]], dtype=object)  # int column next to a float column
    ym = np.array([0, 1, 1, 0])
    _trees_equal(DecisionTreeClassifier(device="cpu").fit(mixed, ym).tree_, ref.DecisionTreeClassifier().fit(mixed, ym).tree_)


@pytest.mark.parametrize("depth,max_depth", [(0, 2), (1, 2), (2, 2), (3, 2), (1, None)])
def test_make_tree_called_directly_matches_reference(ref, depth, max_depth):
    """``_make_tree(X, y, parent=, depth=)`` as a public-ish entry point (reference ``:93-166``).

    ``depth`` only feeds the ``max_depth == depth`` stop test (a start below max_depth never stops on
    depth); ``Node.depth`` follows the parent chain; thresholds of a numeric array are numpy scalars.
    """
    rng = np.random.default_rng(11)
    X = rng.normal(size=(60, 3)).round(1)
    y = rng.integers(0, 3, size=60)
    ours, theirs = DecisionTreeClassifier(max_depth=max_depth, device="cpu"), ref.DecisionTreeClassifier(max_depth=max_depth)
    for clf in (ours, theirs):
        clf.classes_, clf.n_features_ = np.unique(y), 3
    a, b = ours._make_tree(X, y, depth=depth), theirs._make_tree(X, y, depth=depth)
    _trees_equal(a, b)
    assert a.depth == b.depth == 0 and a.parent is None
    if not b.is_leaf:
        assert type(a.threshold) is type(b.threshold) is np.float64
    from mpitree_b200.tree._base import Node

    RefNode = type(theirs._make_tree(X, y))  # the reference's own Node class
    pa, pb = Node(value=0, threshold=0.0, depth=4), RefNode(value=0, threshold=0.0, depth=4)
    a, b = ours._make_tree(X, y, parent=pa, depth=depth), theirs._make_tree(X, y, parent=pb, depth=depth)
    _trees_equal(a, b)
    assert a.depth == b.depth == 5 and a.parent is pa
    for f in range(3):
        ga, gb = ours._compute_information_gain(X, y, f), theirs._compute_information_gain(X, y, f)
        assert ga == gb and type(ga[1]) is type(gb[1])


def test_string_features_ordinal(ref):
    X = np.array([["apple", "x"], ["banana", "y"], ["cherry", "x"], ["date", "y"], ["elder", "x"], ["fig", "y"]], dtype=object)
    y = np.array([0, 0, 0, 1, 1, 1])
    ours = DecisionTreeClassifier(device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier().fit(X, y)
    assert int(ours.tree_.value) == int(theirs.tree_.value)
    assert ours.tree_.threshold == theirs.tree_.threshold == "cherry"
    assert np.array_equal(ours.predict(X), theirs.predict(X))


def test_internal_helpers_match_reference(ref, iris):
    X, y = iris.data, iris.target
    ours, theirs = DecisionTreeClassifier(device="cpu"), ref.DecisionTreeClassifier()
    assert ours._compute_entropy(y) == theirs._compute_entropy(y)
    Xo = X.astype(object)
    for f in range(4):
        g1, t1 = ours._compute_information_gain(X, y, f)
        g2, t2 = theirs._compute_information_gain(Xo, y, f)
        assert g1 == g2 and t1 == t2
    ours.classes_ = theirs.classes_ = np.unique(y)
    ours.n_features_ = theirs.n_features_ = 4
    sub = ours._make_tree(X[:100], y[:100], depth=1)
    sub_ref = theirs._make_tree(Xo[:100], y[:100], depth=1)
    # Node.depth follows the parent chain (none here); the depth argument only feeds the stop rule
    assert sub.depth == sub_ref.depth == 0 and int(sub.value) == int(sub_ref.value) and sub.threshold == sub_ref.threshold


@pytest.mark.parametrize("kw", [dict(max_depth=-1), dict(max_depth=2.0), dict(max_depth=2.5), dict(min_samples_split=2.5),
                                dict(min_samples_split=7.0), dict(min_samples_split=0), dict(max_depth=0)])
def test_unusual_parameter_values_act_like_the_reference(ref, kw):
    """max_depth is only ever compared with ``==`` and min_samples_split with ``<`` (reference ``:121-122``)."""
    rng = np.random.default_rng(0)
    X = rng.integers(0, 5, size=(40, 3)).astype(float)
    y = rng.integers(0, 3, size=40)
    assert DecisionTreeClassifier(device="cpu", **kw).fit(X, y).export_text() == ref.DecisionTreeClassifier(**kw).fit(X, y).export_text()


def test_wide_bins_keep_the_search_exact_beyond_256_distinct_values(ref):
    """``max_bins > 256``: every distinct value stays a candidate (reference ``:73``) through virtual one-byte
    columns (``BinMapper.expand_wide``); the default 256 quantile bins cannot reproduce the reference here."""
    rng = np.random.default_rng(0)
    n = 700
    X = np.round(rng.normal(size=(n, 3)), 3)  # ~700 distinct values in two columns
    X[:, 2] = rng.integers(0, 5, size=n)
    y = (X[:, 0] + 0.5 * X[:, 1] > 0.1).astype(np.int64) + (X[:, 2] > 2)
    theirs = ref.DecisionTreeClassifier(max_depth=4).fit(X, y)
    ours = DecisionTreeClassifier(max_depth=4, device="cpu", max_bins=1024).fit(X, y)
    assert ours.export_text() == theirs.export_text()
    assert np.array_equal(ours.predict(X), theirs.predict(X))
    coarse = DecisionTreeClassifier(max_depth=4, device="cpu").fit(X, y)
    assert coarse.export_text() != theirs.export_text()  # (what the wide mode is for)
    # thread ranks with row shards agree with the serial fit
    from mpitree_b200.parallel.comm import run_threaded
    from mpitree_b200.tree import ParallelDecisionTreeClassifier

    def work(comm):
        lo, hi = n * comm.rank // comm.size, n * (comm.rank + 1) // comm.size
        clf = ParallelDecisionTreeClassifier(max_depth=4, device="cpu", max_bins=1024, data_layout="sharded", comm=comm)
        return clf.fit(X[lo:hi], y[lo:hi]).export_text()

    assert set(run_threaded(3, work)) == {theirs.export_text()}
