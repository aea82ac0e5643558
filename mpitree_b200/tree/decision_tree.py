"""Estimators: ``DecisionTreeClassifier`` and ``ParallelDecisionTreeClassifier``.

Public surface and behaviour follow the reference
(``mpitree/tree/decision_tree.py``): keyword-only ``max_depth`` /
``min_samples_split`` (``:33-35``), ``fit`` (``:168-190`` serial, ``:340-362``
parallel), ``predict_proba`` returning the reached leaf's raw class counts
(``:192-227``), ``predict`` = arg-max of those counts (``:229-248``),
``export_text`` (``:250-307``), fitted attributes ``n_features_``, ``classes_``,
``tree_`` (``:186-189``), plus what ``BaseEstimator``/``ClassifierMixin`` give
(``get_params``, ``set_params``, ``clone``, ``score``).

What is different is everything underneath: features are binned once, trees are
grown level by level from count histograms, on B200s by hand-written sm_100a
kernels (:mod:`mpitree_b200.ops.cuda_engine`), on hosts by the numpy engine
(:mod:`mpitree_b200.ops.cpu_engine`).  The parallel class shards *rows* across
ranks (one rank per GPU) instead of replicating the data set and splitting the
communicator by subtree.

Deliberate, documented deviations from reference *defects* (outputs are
unchanged wherever the reference succeeds; see DESIGN.md "Quirk policy"):

* a split whose right side would be empty is never chosen (the reference then
  recurses forever or raises ``IndexError``, ``:150-164``);
* ``count`` always has ``max(label) + 1`` entries, so gappy labels predict
  (the reference's ``predict`` raises, ``:126`` / ``:227``);
* ``export_text`` prints string thresholds instead of raising (``:276``);
* base-class order makes ``sklearn.base.is_classifier`` true (``:17``);
* importing the package needs neither MPI nor a GPU (``:313`` imports mpi4py in
  the class body);
* ``predict`` / ``predict_proba`` raise ``ValueError`` when the input has a different
  number of features than ``fit`` saw (the reference has no check, ``:206``).
"""

from __future__ import annotations

import dataclasses

import numbers
from typing import Optional

import numpy as np
from sklearn.base import BaseEstimator, ClassifierMixin
from sklearn.utils.validation import check_array, check_is_fitted, check_X_y

from ..models.export import render_text
from ..models.flat_tree import FlatTree
from ..ops.binning import BinMapper
from ..ops.cpu_engine import SplitParams, best_split, exact_entropy
from ..parallel import comm as _comm
from ..parallel.scheduler import fit_flat_tree_host, fit_flat_tree_replicated, resolve_engine
from ._base import Node

__all__ = ["DecisionTreeClassifier", "ParallelDecisionTreeClassifier"]

#: replicated inputs with at most this many rows are fitted redundantly on every rank (no messages)
REPLICATED_SERIAL_ROWS = 4096
#: ... and from this many rows on the subtrees below the redundantly built top are dealt out to the ranks
REPLICATED_SPLIT_MIN_ROWS = 64
#: host torch inputs up to this many rows get the reference's printed sign for zero thresholds
ZERO_SIGN_ROWS_MAX = 1 << 20


def _is_torch_tensor(x) -> bool:
    return type(x).__module__.startswith("torch") and hasattr(x, "is_cuda")


def _validate_labels(y: np.ndarray) -> np.ndarray:
    """Labels must be non-negative integers, as ``np.bincount`` demands (``:125``)."""
    y = np.asarray(y)
    if y.dtype == object:
        try:
            as_int = y.astype(np.int64)
        except (TypeError, ValueError):
            raise TypeError("class labels must be non-negative integers") from None
        if not np.array_equal(as_int, y):
            raise TypeError("class labels must be non-negative integers")
        y = as_int
    if y.dtype == bool:
        y = y.astype(np.int64)
    if not np.issubdtype(y.dtype, np.integer):
        raise TypeError(
            f"Cannot cast array data from dtype('{y.dtype}') to dtype('int64') according to the rule 'safe'"
        )
    if y.size and y.min() < 0:
        raise ValueError("'list' argument must have no negative elements")
    return y.astype(np.int64, copy=False)


def _restore_zero_signs(flat: FlatTree, X: np.ndarray, comm=None) -> None:
    """Give a zero threshold the sign the reference would print (``-0.00`` vs ``0.00``).

    ``-0.0 == 0.0``, so both land in one bin and every decision is unaffected; only the printed
    threshold differs.  The reference takes its thresholds from ``np.unique`` of the *node's* object
    column (``:73``), which keeps whichever zero its sort puts first; the same call on the same rows
    in the same order reproduces that choice.  Only does anything when a split column really holds
    both signs (rare: rounded negative values).  With several ranks ``X`` is this rank's shard and
    the node's column is assembled in rank order (one small all-gather to find out whether any
    column is affected, then one per tree level that has an affected node); every rank takes the
    same decisions, so the collectives match up.
    """
    multi = comm is not None and comm.size > 1
    if X.dtype.kind != "f":
        return
    nodes = [i for i in range(flat.n_nodes) if flat.feature[i] >= 0 and flat.thresholds[i] == 0]
    if not nodes:
        return
    feats = sorted({int(flat.feature[i]) for i in nodes})
    signs = {}
    for f in feats:
        neg = np.signbit(X[X[:, f] == 0, f])
        signs[f] = (bool(neg.any()), bool((~neg).any()))
    if multi:
        everyone = comm.allgather_object(signs)
        signs = {f: (any(e[f][0] for e in everyone), any(e[f][1] for e in everyone)) for f in feats}
    nodes = [i for i in nodes if all(signs[int(flat.feature[i])])]
    if not nodes:
        return
    thr = np.array([np.nan if t is None else float(t) for t in flat.thresholds])
    at = np.zeros(X.shape[0], dtype=np.int64)  # node every row currently sits in, advanced level by level
    todo = {int(flat.depth[i]): [] for i in nodes}
    for i in nodes:
        todo[int(flat.depth[i])].append(i)
    for depth in range(max(todo) + 1):
        if depth in todo:
            cols = {i: X[at == i, int(flat.feature[i])] for i in todo[depth]}
            if multi:
                everyone = comm.allgather_object(cols)
                cols = {i: np.concatenate([e[i] for e in everyone]) for i in todo[depth]}
            for i, col in cols.items():
                # (a quantile-binned column can have an edge of 0 in a node that holds no zero itself)
                flat.thresholds[i] = next((v for v in np.unique(col.astype(object)) if v == 0), flat.thresholds[i])
        inner = flat.feature[at] >= 0
        rows = np.flatnonzero(inner)
        go_left = X[rows, flat.feature[at[rows]]] <= thr[at[rows]]
        at[rows] = np.where(go_left, flat.left[at[rows]], flat.right[at[rows]])


def _as_feature_matrix(X: np.ndarray) -> np.ndarray:
    """Keep numeric dtypes numeric; leave genuine object/string columns as objects."""
    if X.dtype.kind in "US":
        return X.astype(object)
    if X.dtype == object:
        # An object array whose columns all hold one kind of number becomes that numeric dtype
        # (thresholds keep their Python type: int stays int).  Anything else stays an object
        # array and takes the ordinal path: strings (also numeric-looking ones, which the
        # reference orders as text), mixed int / float columns, ints beyond 64 bits.
        kinds = set()
        for f in range(X.shape[1]):
            try:
                kinds.add(np.array(X[:, f].tolist()).dtype.kind)
            except (TypeError, ValueError):
                return X
        if len(kinds) == 1 and next(iter(kinds)) in "biuf":
            out = np.array(X.tolist())
            if out.shape == X.shape and out.dtype.kind in "biuf":
                return out
        return X
    return X


class DecisionTreeClassifier(ClassifierMixin, BaseEstimator):
    """Exact-candidate / histogram decision tree classifier (entropy criterion).

    Parameters
    ----------
    max_depth : int, optional
        Maximum depth; ``None`` grows until every leaf is pure, constant or
        smaller than ``min_samples_split``.
    min_samples_split : int, default=2
        A node with fewer rows is not split.
    criterion : {"entropy", "gini"}, default="entropy"
        ``"entropy"`` is the reference's criterion; ``"gini"`` is an extension.
    max_bins : int, default=256
        Features with at most this many distinct values are searched exactly
        (every distinct value is a candidate, like the reference); others are
        quantile-binned.  Values above 256 (up to 65,536; numpy inputs) keep the
        search exact for such features at the price of extra one-byte columns
        (``BinMapper.expand_wide``).
    device : {"auto", "cpu", "cuda"}, default="auto"
        ``auto`` uses the CUDA engine whenever a GPU is present.
    """

    def __init__(
        self,
        *,
        max_depth: Optional[int] = None,
        min_samples_split: int = 2,
        criterion: str = "entropy",
        max_bins: int = 256,
        device: str = "auto",
    ):
        self.max_depth = max_depth
        self.min_samples_split = min_samples_split
        self.criterion = criterion
        self.max_bins = max_bins
        self.device = device

    # ``tree_`` (the reference's Node graph, ``:189``) is materialised from the flat arrays the
    # first time it is read: fitting a 2000-node tree should not pay for 2000 Python objects
    # unless somebody looks at them.
    def __getattr__(self, name):
        if name == "tree_":
            flat = self.__dict__.get("flat_tree_")
            if flat is not None:
                root = flat.to_nodes()
                self.__dict__["tree_"] = root
                return root
        raise AttributeError(f"{type(self).__name__!r} object has no attribute {name!r}")

    # ------------------------------------------------------------------ util
    def _params(self, depth_offset: int = 0) -> SplitParams:
        if self.criterion not in ("entropy", "gini"):
            raise ValueError("criterion must be 'entropy' or 'gini'")
        md = self.max_depth
        if md is not None:
            if isinstance(md, bool) or not isinstance(md, numbers.Real):
                raise ValueError("max_depth must be a number or None")
            # the reference's only use of max_depth is ``max_depth == depth`` (``:121``): 2.0 acts like 2,
            # a negative or fractional value never equals a depth and therefore never stops the growth
            if md < 0 or md != int(md):
                md = None
        if md is not None:
            # the reference stops on ``max_depth == depth`` (``:121``): a subtree started below
            # max_depth never meets the test and grows without a depth limit
            md = int(md) - depth_offset if int(md) >= depth_offset else None
        # ``n_samples < min_samples_split`` (``:122``) with a fractional value equals the test against its ceiling
        mss = int(np.ceil(self.min_samples_split))
        return SplitParams(max_depth=md, min_samples_split=mss, criterion=self.criterion)

    def _communicator(self):
        return None

    def _shard(self, X, y, comm):
        return X, y

    def _builder_kwargs(self):
        """Extra arguments for the CUDA ``TreeBuilder`` (``builder_kwargs`` attribute, hand-off size)."""
        kw = dict(getattr(self, "builder_kwargs", None) or {})
        handoff = getattr(self, "handoff_rows", None)
        if handoff is not None:
            kw.setdefault("small_rows", int(handoff))
        return kw

    # --------------------------------------------------- reference internals
    def _compute_entropy(self, x):
        """Shannon entropy (base 2) of a label array (reference ``:37-51``)."""
        x = np.asarray(x)
        if x.size == 0:
            return -0.0
        _, counts = np.unique(x, return_counts=True)
        return exact_entropy(counts)

    def _compute_information_gain(self, X, y, feature_idx):
        """``(gain, threshold)`` of the best split on one feature (reference ``:53-91``).

        The reference also admits the column maximum as a candidate (gain 0, empty
        right side); it is returned only when the column is constant.
        """
        X = np.asarray(X)
        y = _validate_labels(np.asarray(y))
        col = _as_feature_matrix(X[:, [feature_idx]])
        mapper = BinMapper.fit(col, max_bins=256)
        codes = mapper.transform(col)[:, 0]
        classes, y_codes = np.unique(y, return_inverse=True)
        B, C = int(mapper.n_bins[0]), len(classes)
        hist = np.zeros((1, B, C), dtype=np.int64)
        np.add.at(hist[0], (codes, y_codes), 1)
        res = best_split(hist, mapper.n_bins, self.criterion)
        as_input = (lambda t: X.dtype.type(t)) if X.dtype != object and X.dtype.kind in "biuf" else (lambda t: t)
        if res is None:
            return 0.0, as_input(mapper.threshold_of(0, B - 1))
        _, b, gain, _ = res
        return max(gain, 0.0) if gain < 0 else gain, as_input(mapper.threshold_of(0, b))

    def _make_tree(self, X, y, *, parent=None, depth=0):
        """Build the subtree for ``(X, y)`` and return its root ``Node`` (reference ``:93-166``)."""
        X = np.asarray(X)
        flat = self._fit_flat(X, np.asarray(y), depth_offset=depth, comm=None, local=True)
        root = flat.to_nodes()
        if parent is not None or depth:
            _reparent(root, parent, depth)
        _match_threshold_types(root, X)
        return root

    # ------------------------------------------------------------------- fit
    def _fit_flat(self, X, y, *, depth_offset=0, comm=None, local=False) -> FlatTree:
        y = _validate_labels(y)
        X = _as_feature_matrix(X)
        classes = self.classes_ if hasattr(self, "classes_") else np.unique(y)
        # class code = index into classes_
        y_codes = np.searchsorted(classes, y)
        numeric = X.dtype != object
        engine = resolve_engine(self.device, numeric=numeric, n_classes=len(classes))
        params = self._params(depth_offset)

        if numeric and self.max_bins > 256:
            # Exact search beyond 256 distinct values per feature (the reference examines every distinct
            # value, ``:73``): features with more bins than a one-byte code can name are presented to the
            # engines as several virtual columns (``BinMapper.expand_wide``); the chosen (column, bin) pairs
            # are folded back into (feature, global bin) afterwards.
            summaries = None
            if comm is not None and comm.size > 1:
                summaries = comm.allgather_object(BinMapper.local_summary(X, cap=max(4096, int(self.max_bins))))
            mapper = BinMapper.fit(X, max_bins=self.max_bins, summaries=summaries, wide=True)
            vcodes, v_feature, v_chunk, v_nbins = mapper.expand_wide(X)
            if engine == "cuda":
                from ..ops import cuda_engine

                flat = cuda_engine.fit_from_codes(vcodes, y_codes, len(classes), int(v_nbins.max()), params, comm=comm,
                                                  builder_kwargs=self._builder_kwargs())
            else:
                flat = fit_flat_tree_host(vcodes, y_codes, v_nbins, len(classes), params, comm,
                                          handoff_rows=getattr(self, "handoff_rows", None),
                                          group_rows=getattr(self, "group_rows", None))
            # (new arrays: thread-ranks of one process can share the tree object the scheduler returns)
            inner = flat.feature >= 0
            vf = flat.feature[inner]
            feature, bins = flat.feature.copy(), flat.bin.copy()
            bins[inner] = 255 * v_chunk[vf] + flat.bin[inner]
            feature[inner] = v_feature[vf]
            flat = dataclasses.replace(flat, feature=feature, bin=bins)
        elif engine == "cuda" and numeric:
            from ..ops import cuda_engine

            flat, mapper = cuda_engine.fit_from_host(
                X, y_codes, len(classes), params, max_bins=self.max_bins, comm=comm,
                builder_kwargs=self._builder_kwargs(),
            )
        elif engine == "cuda":
            # object / string columns: ordinal codes in sorted-unique order on the host (the order Python's
            # ``<=`` gives the reference, ``:73-77``), the tree from the codes on the device
            from ..ops import cuda_engine

            summaries = None
            if comm is not None and comm.size > 1:
                summaries = comm.allgather_object(BinMapper.local_summary(X))
            mapper = BinMapper.fit(X, max_bins=self.max_bins, summaries=summaries)
            flat = cuda_engine.fit_from_codes(
                mapper.transform(X), y_codes, len(classes), int(mapper.n_bins.max()), params, comm=comm,
                builder_kwargs=self._builder_kwargs(),
            )
        else:
            summaries = None
            if comm is not None and comm.size > 1:
                mine = BinMapper.local_summary(X)
                summaries = comm.allgather_object(mine)  # every rank, own one included, in rank order
            mapper = BinMapper.fit(X, max_bins=self.max_bins, summaries=summaries)
            codes = mapper.transform(X)
            split_comm = getattr(self, "_split_comm", None)
            if split_comm is not None:  # replicated rows: redundant trunk, subtrees dealt out (see ``fit``)
                flat = fit_flat_tree_replicated(codes, y_codes, mapper.n_bins, len(classes), params, split_comm)
            else:
                flat = fit_flat_tree_host(
                    codes, y_codes, mapper.n_bins, len(classes), params, comm,
                    handoff_rows=getattr(self, "handoff_rows", None), group_rows=getattr(self, "group_rows", None),
                )
        flat.class_labels = np.asarray(classes, dtype=np.int64)
        flat.thresholds = mapper.thresholds_of(flat.feature, flat.bin)
        if engine == "cuda" and X.dtype.kind in "iub":
            # the CUDA engine bins integer columns through a float edge table: hand thresholds back in
            # the column's own type, as the host engine and the reference do (int stays int)
            flat.thresholds = [None if t is None else X.dtype.type(t).item() for t in flat.thresholds]
        _restore_zero_signs(flat, X, comm)
        if not local:
            self.bin_mapper_ = mapper
            self.engine_ = engine
        return flat

    def fit(self, X, y):
        """Train the classifier on ``X`` (n_samples, n_features) and integer labels ``y``."""
        comm = self._communicator()
        if _is_torch_tensor(X):
            return self._fit_torch(X, y, comm)
        X, y = check_X_y(X, y, dtype=None, ensure_all_finite=True)
        y = _validate_labels(y)
        self.n_features_ = X.shape[1]
        self.n_features_in_ = X.shape[1]
        self.classes_ = np.unique(y)
        X, y = self._shard(X, y, comm)
        self.__dict__.pop("tree_", None)
        self.flat_tree_ = self._fit_flat(X, y, comm=comm)
        return self

    def _fit_torch(self, X, y, comm):
        """``fit`` for torch tensors (host-pinned or CUDA): no numpy round trip."""
        from ..ops import cuda_engine

        if X.dim() != 2:
            raise ValueError("X must be 2-dimensional")
        resolve_engine(
            "cuda" if self.device == "auto" else self.device,
            numeric=True,
            n_classes=2,
            is_cuda_input=bool(X.is_cuda),
        )
        flat, mapper, classes = cuda_engine.fit_from_torch(
            X, y, self._params(), max_bins=self.max_bins, comm=comm,
            replicated=getattr(self, "data_layout", "sharded") == "replicated",
            builder_kwargs=self._builder_kwargs(),
        )
        self.n_features_ = int(X.shape[1])
        self.n_features_in_ = int(X.shape[1])
        self.classes_ = classes
        flat.class_labels = np.asarray(classes, dtype=np.int64)
        flat.thresholds = mapper.thresholds_of(flat.feature, flat.bin)
        # (decided from the largest shard, a value every rank shares: the restore runs collectives)
        if not X.is_cuda and getattr(flat, "max_shard_rows", X.shape[0]) <= ZERO_SIGN_ROWS_MAX:
            # rows that live on the host can be looked at for free (see _restore_zero_signs); device
            # inputs and very large host inputs keep the sign of the global bin edge.  In the replicated
            # layout every rank holds all rows, so nothing has to be gathered.
            replicated = getattr(self, "data_layout", "sharded") == "replicated"
            _restore_zero_signs(flat, X.numpy(), None if replicated else comm)
        self.bin_mapper_ = mapper
        self.engine_ = "cuda"
        self.__dict__.pop("tree_", None)
        self.flat_tree_ = flat
        return self

    # --------------------------------------------------------------- predict
    def _leaf_ids(self, X) -> np.ndarray:
        flat = self._flat()
        if _is_torch_tensor(X):
            from ..ops import cuda_engine

            self._check_n_features(X)
            return cuda_engine.predict_leaf_ids(flat, X)
        X = _as_feature_matrix(check_array(X, dtype=None, ensure_all_finite=True))
        self._check_n_features(X)
        if X.dtype != object and getattr(self, "engine_", "cpu") == "cuda" and self.device != "cpu":
            from ..ops import cuda_engine

            if cuda_engine.available():
                return cuda_engine.predict_leaf_ids(flat, X)
        return flat.apply_raw(X)

    def _check_n_features(self, X) -> None:
        # (the reference indexes whatever it is given; a wrong width is always a caller bug)
        want = getattr(self, "n_features_in_", None)
        if want is not None and X.ndim == 2 and X.shape[1] != want:
            raise ValueError(
                f"X has {X.shape[1]} features, but {type(self).__name__} is expecting {want} features as input."
            )

    def _flat(self) -> FlatTree:
        check_is_fitted(self)
        flat = self.__dict__.get("flat_tree_")
        if flat is None:  # e.g. tree_ assigned by hand or unpickled from an older object
            flat = FlatTree.from_nodes(self.tree_)
            self.flat_tree_ = flat
        return flat

    def apply(self, X) -> np.ndarray:
        """Index (breadth-first node id) of the leaf each row ends in."""
        ids = self._leaf_ids(X)
        return ids.cpu().numpy() if _is_torch_tensor(ids) else ids

    def predict_proba(self, X):
        """Raw per-class training counts of the leaf each row reaches (int64).

        Like the reference these are counts, not normalised probabilities.
        """
        ids = self._leaf_ids(X)
        counts = self._flat().expanded_counts()
        if _is_torch_tensor(ids):
            import torch

            return torch.as_tensor(counts, device=ids.device)[ids]
        return counts[ids]

    def predict(self, X):
        """Predicted class label per row (arg-max of the leaf counts, ties -> lowest label)."""
        if _is_torch_tensor(X):
            from ..ops import cuda_engine

            flat = self._flat()
            self._check_n_features(X)
            return cuda_engine.predict_leaf_ids(flat, X, want="label")  # label looked up in-kernel
        ids = self._leaf_ids(X)
        labels = self._flat().expanded_counts().argmax(axis=1)
        return labels[ids]

    # ---------------------------------------------------------------- export
    def export_text(self, *, feature_names=None, class_names=None, precision=2):
        """Text rendering of the fitted tree (grammar in :mod:`mpitree_b200.models.export`)."""
        check_is_fitted(self, "tree_")
        return render_text(
            self.tree_, feature_names=feature_names, class_names=class_names, precision=precision
        )

    # ------------------------------------------------------------ inspection
    def get_depth(self) -> int:
        return self._flat().max_depth

    def get_n_leaves(self) -> int:
        return self._flat().n_leaves

    def __sklearn_tags__(self):
        tags = super().__sklearn_tags__()
        tags.input_tags.string = True
        return tags


def _match_threshold_types(root: Node, X: np.ndarray) -> None:
    """Thresholds are elements of ``np.unique(X[:, f])`` in the reference (``:73``, ``:91``): numpy
    scalars when ``_make_tree`` is handed a numeric array, Python objects after ``fit``'s object cast."""
    if X.dtype == object or X.dtype.kind not in "biuf":
        return
    for n in root.iter_preorder():
        if not n.is_leaf:
            n.threshold = X.dtype.type(n.threshold)


def _reparent(root: Node, parent: Optional[Node], depth: int) -> None:
    root.parent = parent
    # Node.depth follows the parent chain only (reference ``_base.py:59-61``); the ``depth`` argument
    # of ``_make_tree`` feeds the max_depth stop rule and nothing else (``:121``)
    base = parent.depth + 1 if parent is not None else 0
    for n in root.iter_preorder():
        n.depth = base if n is root else n.parent.depth + 1


class _WorldAttr:
    """Class attribute resolved lazily from the live process group.

    The reference evaluates ``MPI.COMM_WORLD`` in the class body, i.e. at import
    time (``:313-317``); resolving on access keeps ``clf.WORLD_RANK`` working in
    user scripts without making ``import`` depend on a launcher.
    """

    def __init__(self, what: str):
        self.what = what

    def __get__(self, obj, owner):
        w = _comm.world()
        return {"comm": w, "rank": w.rank, "size": w.size}[self.what]


class ParallelDecisionTreeClassifier(DecisionTreeClassifier):
    """Multi-rank decision tree classifier (one rank per GPU, rows sharded).

    Run the same script on every rank (``torchrun --nproc-per-node N``).  All
    ranks end ``fit`` holding the identical tree, as in the reference.

    Parameters (in addition to :class:`DecisionTreeClassifier`)
    ----------
    data_layout : {"replicated", "sharded"}, default="replicated"
        ``replicated``: every rank passes the full ``(X, y)`` like the reference
        (``:356-361``) and a row shard is taken internally.  ``sharded``: every
        rank passes only its own rows.
    handoff_rows : int, optional
        Global row count below which a subtree is handed to a single owner rank
        (``None`` = automatic, ``0`` = never).
    group_rows : int, optional
        Global row count below which a node leaves the all-ranks build for a rank
        *subgroup*: the communicator is halved the way the reference halves it
        (``:336-338``, ``:446-466``), the node's rows move to its half and the half goes
        on cooperatively, halving again further down (host engine, four ranks or
        more; ``None`` = four times ``handoff_rows``, ``0`` = never).
    comm : communicator, optional
        Defaults to the world group at ``fit`` time.
    """

    WORLD_COMM = _WorldAttr("comm")
    WORLD_RANK = _WorldAttr("rank")
    WORLD_SIZE = _WorldAttr("size")

    def __init__(
        self,
        *,
        max_depth: Optional[int] = None,
        min_samples_split: int = 2,
        criterion: str = "entropy",
        max_bins: int = 256,
        device: str = "auto",
        data_layout: str = "replicated",
        handoff_rows: Optional[int] = None,
        group_rows: Optional[int] = None,
        comm=None,
    ):
        super().__init__(
            max_depth=max_depth,
            min_samples_split=min_samples_split,
            criterion=criterion,
            max_bins=max_bins,
            device=device,
        )
        self.data_layout = data_layout
        self.handoff_rows = handoff_rows
        self.group_rows = group_rows
        self.comm = comm

    def _communicator(self):
        if getattr(self, "_solo", False):
            return None  # tiny replicated problem: no communication (see ``fit``)
        return self.comm if self.comm is not None else _comm.world()

    def _get_cyclic_dist(self, comm=None, *, n_blocks: int = 1):
        """Round-robin communicator split (reference ``:319-338``): colour = ``rank % n_blocks``."""
        comm = comm if comm is not None else self._communicator()
        key, color = divmod(comm.rank, n_blocks)
        return comm.split(color, key)

    def _shard(self, X, y, comm):
        if self.data_layout not in ("replicated", "sharded"):
            raise ValueError("data_layout must be 'replicated' or 'sharded'")
        if self.data_layout == "sharded" or comm is None or comm.size <= 1:
            return X, y
        n = X.shape[0]
        lo = (n * comm.rank) // comm.size
        hi = (n * (comm.rank + 1)) // comm.size
        return X[lo:hi], y[lo:hi]

    def fit(self, X, y):
        """Train on all ranks; see :class:`ParallelDecisionTreeClassifier` for data layouts."""
        comm = self._communicator()
        if self.data_layout == "sharded" and comm.size > 1 and not _is_torch_tensor(X):
            # classes must be agreed on globally before encoding labels
            # a rank may legitimately hold no rows of a small sharded data set
            X, y = check_X_y(X, y, dtype=None, ensure_all_finite=True, ensure_min_samples=0)
            y = _validate_labels(y)
            parts = comm.allgather_object(np.unique(y))
            self.n_features_ = X.shape[1]
            self.n_features_in_ = X.shape[1]
            self.classes_ = np.unique(np.concatenate(parts))
            self.__dict__.pop("tree_", None)
            self.flat_tree_ = self._fit_flat(X, y, comm=comm)
            return self
        if (
            self.data_layout == "replicated"
            and comm.size > 1
            and not _is_torch_tensor(X)
            and self.handoff_rows is None  # an explicit hand-off size asks for the cooperative path
            and np.shape(X)[0] <= REPLICATED_SERIAL_ROWS
        ):
            # Small replicated problem: every rank already holds all rows, so nothing has to be
            # sharded or reduced.  The top of the tree is built redundantly on every rank (the
            # reference does the same at the top of its recursion, ``:408-419``), the subtrees below
            # are dealt out and finished independently, and one all-gather distributes them
            # (``parallel/scheduler.py::fit_flat_tree_replicated``).  Inputs too small to be worth a
            # collective are simply fitted by every rank.
            serial = DecisionTreeClassifier.fit
            self._solo = True
            self._split_comm = comm if np.shape(X)[0] >= REPLICATED_SPLIT_MIN_ROWS else None
            try:
                return serial(self, X, y)
            finally:
                self._solo = False
                self._split_comm = None
        return super().fit(X, y)

    def _make_tree(self, X, y, *, parent=None, depth=0, comm=None):
        """Subtree for ``(X, y)`` built cooperatively over ``comm`` (reference ``:364-479``)."""
        comm = comm if comm is not None else self._communicator()
        Xs, ys = self._shard(np.asarray(X), np.asarray(y), comm)
        flat = self._fit_flat(Xs, ys, depth_offset=depth, comm=comm, local=True)
        root = flat.to_nodes()
        if parent is not None or depth:
            _reparent(root, parent, depth)
        return root

    def __getstate__(self):
        state = super().__getstate__()
        state = dict(state)
        state["comm"] = None  # communicators are process-local
        return state
