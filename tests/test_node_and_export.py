"""Object format + text format (§2.4)."""
import pickle

import numpy as np
import pytest

from mpitree_b200.models.export import render_text
from mpitree_b200.models.flat_tree import FlatTree
from mpitree_b200.tree import BranchType, DecisionTreeClassifier, Node

from .goldens import IRIS2_DEPTH3, IRIS2_DEPTH5_P1


def test_branch_glyphs():
    assert [b.value for b in BranchType] == ["┌──", "├──", "└──"]
    assert BranchType.ROOT.name == "ROOT"


def test_node_defaults_and_repr():
    n = Node(value=np.int64(0), threshold=5.5, count=np.array([50, 50, 50]))
    assert repr(n) == "Node(value=np.int64(0), threshold=5.5, depth=0, count=array([50, 50, 50]))"
    assert n.is_leaf and n.children == [] and n.depth == 0 and n._btype is BranchType.ROOT
    with pytest.raises(TypeError):
        Node(0)  # keyword-only


def test_depth_from_parent_and_children():
    root = Node(value=1, threshold=2.0)
    a = Node(value=0, parent=root)
    b = Node(value=1, parent=root)
    root.left, root.right = a, b
    assert a.depth == b.depth == 1
    assert root.children == [a, b] and not root.is_leaf


def test_lt_tags_siblings():
    root = Node(value=0, threshold=1.0)
    l, r = Node(value=0, parent=root), Node(value=1, parent=root)
    root.left, root.right = l, r
    assert sorted(root.children) == [l, r]
    assert (l._btype, r._btype) == (BranchType.INTERIOR_LIKE, BranchType.LEAF_LIKE)
    # interior right child is listed first
    r.left, r.right = Node(value=0, parent=r), Node(value=1, parent=r)
    assert sorted(root.children) == [r, l]
    assert (l._btype, r._btype) == (BranchType.LEAF_LIKE, BranchType.INTERIOR_LIKE)


def test_goldens(iris):
    X, y = iris.data[:, :2], iris.target
    clf = DecisionTreeClassifier(max_depth=3, device="cpu").fit(X, y)
    assert clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names) == IRIS2_DEPTH3
    clf = DecisionTreeClassifier(max_depth=5, device="cpu").fit(X, y)
    got = clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names, precision=1)
    assert got == IRIS2_DEPTH5_P1


def test_default_labels_and_single_leaf():
    X = np.array([[0.0], [1.0], [2.0], [3.0]])
    y = np.array([0, 0, 1, 1])
    clf = DecisionTreeClassifier(device="cpu").fit(X, y)
    assert clf.export_text() == "┌── feature_0\n│  ├── class: 0 [<= 1.00]\n│  └── class: 1 [> 1.00]"
    leaf = DecisionTreeClassifier(max_depth=0, device="cpu").fit(X, y)
    assert leaf.export_text() == "┌── class: 0"


def test_node_pickle_roundtrip(iris):
    clf = DecisionTreeClassifier(max_depth=2, device="cpu").fit(iris.data, iris.target)
    clone = pickle.loads(pickle.dumps(clf))
    assert clone.export_text() == clf.export_text()
    root = pickle.loads(pickle.dumps(clf.tree_))
    assert render_text(root) == clf.export_text()


def test_flat_roundtrip(iris):
    clf = DecisionTreeClassifier(max_depth=4, device="cpu").fit(iris.data, iris.target)
    flat = clf.flat_tree_
    again = FlatTree.from_nodes(flat.to_nodes())
    assert again.structurally_equal(flat)
    blob = FlatTree.from_bytes(flat.to_bytes())
    assert blob.structurally_equal(flat)
    assert render_text(blob.to_nodes()) == clf.export_text()
    assert np.array_equal(flat.apply_raw(iris.data), clf.apply(iris.data))


def test_string_threshold_prints():
    X = np.array([["a"], ["b"], ["c"], ["d"]], dtype=object)
    y = np.array([0, 0, 1, 1])
    clf = DecisionTreeClassifier(device="cpu").fit(X, y)
    assert "[<= b]" in clf.export_text()
    assert clf.predict(X).tolist() == [0, 0, 1, 1]
