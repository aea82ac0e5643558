"""Structure-of-arrays tree (``FlatTree``): the device/wire format.

The training engines and the batched predict kernel work on flat arrays indexed
by node id (breadth-first numbering, root = 0).  ``FlatTree`` converts to and
from the ``Node`` object graph that the reference exposes as ``clf.tree_``
(``mpitree/tree/_base.py:22-57``); the conversion preserves the reference's
field types: ``value`` is ``np.int64``, ``threshold`` is a plain Python object
taken from the training data, ``count`` an ``int64`` ndarray
(``mpitree/tree/decision_tree.py:124-128`` leaf, ``:143-148`` interior).

It is also what replaces the reference's pickled-subtree ``allgather``
(``decision_tree.py:456-466``): ranks exchange rows of these arrays, a few dozen
bytes per node, instead of pickled object graphs.
"""

from __future__ import annotations

import io
import json
from dataclasses import dataclass, field
from typing import Any, List, Optional

import numpy as np

from ..tree._base import Node

__all__ = ["FlatTree"]


@dataclass
class FlatTree:
    """Flat node arrays, all of length ``n_nodes``.

    Attributes
    ----------
    feature : int32, split feature or -1 for a leaf
    bin : int32, split bin code (rows with ``code <= bin`` go left) or -1
    left, right : int32, child node ids or -1
    parent : int32, parent node id, -1 for the root
    depth : int32
    counts : int64 ``(n_nodes, n_class_codes)``, per-class training rows
    gain : float64, information gain of the split (nan for leaves)
    thresholds : list, raw-value threshold per node (``None`` for leaves)
    class_labels : int64 ``(n_class_codes,)``, class code -> original label
    """

    feature: np.ndarray
    bin: np.ndarray
    left: np.ndarray
    right: np.ndarray
    parent: np.ndarray
    depth: np.ndarray
    counts: np.ndarray
    gain: np.ndarray
    thresholds: List[Any] = field(default_factory=list)
    class_labels: Optional[np.ndarray] = None

    def __getstate__(self):
        state = dict(self.__dict__)
        state.pop("_packed", None)  # device-side cache of the predict kernel, rebuilt on demand
        return state

    # ------------------------------------------------------------------ basic
    @property
    def n_nodes(self) -> int:
        return int(self.feature.shape[0])

    @property
    def n_leaves(self) -> int:
        return int((self.feature < 0).sum())

    @property
    def max_depth(self) -> int:
        return int(self.depth.max()) if self.n_nodes else 0

    def is_leaf(self, i: int) -> bool:
        return self.feature[i] < 0

    def leaf_label_codes(self) -> np.ndarray:
        """argmax class *code* per node (ties -> lowest code, like ``bincount().argmax()``)."""
        return self.counts.argmax(axis=1)

    # ------------------------------------------------------- label expansion
    def expanded_counts(self) -> np.ndarray:
        """Counts indexed by original label value (length ``max_label + 1``).

        The reference's ``count`` is ``np.bincount(y, minlength=len(classes_))``
        (``decision_tree.py:126``), i.e. indexed by label value.  For labels
        ``0..C-1`` this is the identity; for gappy labels it keeps a constant
        width so ``predict`` works (documented fix of a reference defect).
        """
        labels = self.class_labels
        if labels is None:
            return self.counts
        width = int(labels.max()) + 1 if len(labels) else 0
        if width == self.counts.shape[1] and np.array_equal(labels, np.arange(width)):
            return self.counts
        out = np.zeros((self.n_nodes, width), dtype=np.int64)
        out[:, labels] = self.counts
        return out

    # ------------------------------------------------------------ Node graph
    def to_nodes(self) -> Node:
        """Materialise the ``Node`` object graph (iterative, BFS order)."""
        counts = self.expanded_counts()
        nodes: List[Optional[Node]] = [None] * self.n_nodes
        for i in range(self.n_nodes):
            p = int(self.parent[i])
            parent = nodes[p] if p >= 0 else None
            if self.feature[i] < 0:
                value = np.int64(counts[i].argmax())
                node = Node(value=value, count=counts[i].copy(), parent=parent)
            else:
                node = Node(
                    value=np.int64(self.feature[i]),
                    threshold=self.thresholds[i],
                    count=counts[i].copy(),
                    parent=parent,
                )
            nodes[i] = node
            if parent is not None:
                if int(self.left[p]) == i:
                    parent.left = node
                else:
                    parent.right = node
        return nodes[0]

    @classmethod
    def from_nodes(cls, root: Node, n_classes: Optional[int] = None) -> "FlatTree":
        """Flatten a ``Node`` graph (BFS numbering).  ``bin`` is set to -1."""
        order: List[Node] = []
        index = {}
        queue = [root]
        while queue:
            nxt = []
            for n in queue:
                index[id(n)] = len(order)
                order.append(n)
                if not n.is_leaf:
                    nxt.extend((n.left, n.right))
            queue = nxt
        m = len(order)
        width = n_classes or max(len(np.atleast_1d(n.count)) for n in order)
        ft = cls(
            feature=np.full(m, -1, np.int32),
            bin=np.full(m, -1, np.int32),
            left=np.full(m, -1, np.int32),
            right=np.full(m, -1, np.int32),
            parent=np.full(m, -1, np.int32),
            depth=np.zeros(m, np.int32),
            counts=np.zeros((m, width), np.int64),
            gain=np.full(m, np.nan),
            thresholds=[None] * m,
            class_labels=np.arange(width, dtype=np.int64),
        )
        root_depth = root.depth
        for i, n in enumerate(order):
            c = np.atleast_1d(np.asarray(n.count, dtype=np.int64))
            ft.counts[i, : len(c)] = c
            ft.depth[i] = n.depth - root_depth
            if n is not root and n.parent is not None:
                ft.parent[i] = index[id(n.parent)]
            if not n.is_leaf:
                ft.feature[i] = int(n.value)
                ft.thresholds[i] = n.threshold
                ft.left[i] = index[id(n.left)]
                ft.right[i] = index[id(n.right)]
        return ft

    # ------------------------------------------------------------- numerics
    def numeric_thresholds(self, dtype=np.float64) -> np.ndarray:
        """Thresholds as a float array (nan for leaves); raises for string thresholds."""
        out = np.full(self.n_nodes, np.nan, dtype=dtype)
        for i, t in enumerate(self.thresholds):
            if t is not None:
                out[i] = t
        return out

    def apply_codes(self, codes: np.ndarray) -> np.ndarray:
        """Leaf id per row for a matrix of *bin codes* (vectorised host walk)."""
        return self._walk(codes, self.bin, lambda col, thr: col <= thr)

    def apply_raw(self, X: np.ndarray) -> np.ndarray:
        """Leaf id per row for raw feature values (``x[f] <= threshold`` goes left)."""
        if X.dtype == object:
            thr = np.empty(self.n_nodes, dtype=object)
            thr[:] = self.thresholds
        else:
            thr = self.numeric_thresholds()
        return self._walk(X, thr, lambda col, t: col <= t)

    def _walk(self, X, thr, le) -> np.ndarray:
        n = X.shape[0]
        node = np.zeros(n, dtype=np.int64)
        active = np.flatnonzero(self.feature[node] >= 0)
        while active.size:
            cur = node[active]
            f = self.feature[cur]
            go_left = np.asarray(le(X[active, f], thr[cur]), dtype=bool)
            node[active] = np.where(go_left, self.left[cur], self.right[cur])
            active = active[self.feature[node[active]] >= 0]
        return node

    # --------------------------------------------------------- (de)serialise
    def to_bytes(self) -> bytes:
        """Portable serialisation (npz + json thresholds); no pickle involved."""
        buf = io.BytesIO()
        thr = json.dumps([None if t is None else _jsonable(t) for t in self.thresholds])
        np.savez_compressed(
            buf,
            feature=self.feature,
            bin=self.bin,
            left=self.left,
            right=self.right,
            parent=self.parent,
            depth=self.depth,
            counts=self.counts,
            gain=self.gain,
            class_labels=self.class_labels if self.class_labels is not None else np.zeros(0, np.int64),
            thresholds=np.frombuffer(thr.encode(), dtype=np.uint8),
        )
        return buf.getvalue()

    @classmethod
    def from_bytes(cls, data: bytes) -> "FlatTree":
        z = np.load(io.BytesIO(data), allow_pickle=False)
        thr = json.loads(bytes(z["thresholds"]).decode())
        labels = z["class_labels"]
        return cls(
            feature=z["feature"],
            bin=z["bin"],
            left=z["left"],
            right=z["right"],
            parent=z["parent"],
            depth=z["depth"],
            counts=z["counts"],
            gain=z["gain"],
            thresholds=thr,
            class_labels=labels if labels.size else None,
        )

    # wire form between ranks of one fit: the arrays as they are (the communicator pickles them);
    # a subtree exchanged mid-fit carries bin codes, never thresholds, and no compression pass
    _WIRE = ("feature", "bin", "left", "right", "parent", "depth", "counts", "gain")

    def to_wire(self) -> dict:
        return {k: getattr(self, k) for k in self._WIRE}

    @classmethod
    def from_wire(cls, d: dict) -> "FlatTree":
        return cls(thresholds=[None] * len(d["feature"]), **{k: d[k] for k in cls._WIRE})

    def structurally_equal(self, other: "FlatTree") -> bool:
        return (
            self.n_nodes == other.n_nodes
            and np.array_equal(self.feature, other.feature)
            and np.array_equal(self.left, other.left)
            and np.array_equal(self.right, other.right)
            and np.array_equal(self.counts, other.counts)
            and all(_same(a, b) for a, b in zip(self.thresholds, other.thresholds))
        )


def _jsonable(t):
    if isinstance(t, (np.generic,)):
        return t.item()
    return t


def _same(a, b) -> bool:
    if a is None or b is None:
        return a is None and b is None
    return a == b
