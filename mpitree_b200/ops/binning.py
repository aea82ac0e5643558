"""K1 — feature binning / encoding (host implementation + shared metadata).

The reference discovers split candidates with ``np.unique(X[:, f])`` at every
node (``mpitree/tree/decision_tree.py:73``).  Here candidates are discovered
**once per fit**: every feature is mapped to small integer *bin codes* whose
order equals the order of the raw values, so ``x <= t`` on raw values is
``code <= bin`` on codes.  The code->value table is kept so a split threshold is
still reported as a data value (``decision_tree.py:88-91`` returns
``possible_thresholds[argmin]``).

* ``<= max_bins`` distinct values  -> lossless: one bin per distinct value, the
  histogram search examines exactly the reference's candidate set.
* more distinct values             -> quantile bins; the bin's upper edge (a
  data value) is the threshold.
* string / object columns          -> ordinal codes in sorted-unique order, the
  same order Python's ``<=`` gives the reference on object arrays.

The CUDA counterpart (``mpitree_b200/csrc/encode.cu``) implements the encode
pass; edges are always computed from (a sample of) the data with the code here
or with its device twin in :mod:`mpitree_b200.ops.cuda_engine`.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np

__all__ = ["BinMapper", "FEATURE_GROUP", "MAX_WIDE_BINS", "padded_features"]

#: features per device tile row chunk (16 one-byte codes = one 128-bit load)
FEATURE_GROUP = 16
#: most bins per feature in the wide (exact beyond 256 distinct values) mode, see BinMapper.expand_wide
MAX_WIDE_BINS = 1 << 16


def padded_features(n_features: int) -> int:
    return ((n_features + FEATURE_GROUP - 1) // FEATURE_GROUP) * FEATURE_GROUP


def _column_uniques(col: np.ndarray) -> np.ndarray:
    if col.dtype == object:
        try:
            return np.array(sorted(set(col.tolist())), dtype=object)
        except TypeError as exc:  # mixed str/int in one column, like the reference
            raise TypeError(f"cannot order the values of a mixed-type column: {exc}") from None
    return np.unique(col)


@dataclass
class BinMapper:
    """Per-feature ascending upper edges; ``code = first edge >= value``.

    Attributes
    ----------
    edges : list of 1-D arrays (one per feature), ascending data values
    lossless : bool array, True where each bin holds exactly one distinct value
    max_bins : int (<= 256 so codes fit one byte)
    """

    edges: List[np.ndarray]
    lossless: np.ndarray
    max_bins: int = 256
    #: True where the edges list every distinct value that was seen when they were fitted (on a sample
    #: of the rows this is a claim about the sample only: see ``cuda_engine._repair_sampled_edges``)
    enumerated: Optional[np.ndarray] = None

    # ----------------------------------------------------------------- build
    @classmethod
    def fit(
        cls,
        X: np.ndarray,
        max_bins: int = 256,
        *,
        summaries: Optional[Sequence[Sequence[np.ndarray]]] = None,
        wide: bool = False,
    ) -> "BinMapper":
        """Compute edges from ``X`` (n_rows, n_features).

        ``summaries`` (multi-rank fits) is the all-gathered list of every rank's
        :meth:`local_summary`, own rank included, in rank order.  The edges are then derived from
        the summaries alone, with the same operations in the same order on every rank, so all
        ranks end up with identical bins (the cross-rank "edge agreement" step of K1) whatever
        their local rows are.

        ``wide``: allow up to ``MAX_WIDE_BINS`` bins per feature (codes then need :meth:`expand_wide`
        before they can be searched with one-byte bin codes).
        """
        if not 2 <= max_bins <= (MAX_WIDE_BINS if wide else 256):
            raise ValueError(f"max_bins must be in [2, {MAX_WIDE_BINS if wide else 256}]")
        n_features = X.shape[1]
        edges, lossless = [], np.zeros(n_features, dtype=bool)
        for f in range(n_features):
            if summaries:
                parts = [np.asarray(s[f]) for s in summaries]
                parts = [p for p in parts if len(p)]
                values = _column_uniques(np.concatenate(parts)) if parts else np.zeros(0, X.dtype)
                weighted = values  # the summaries carry no multiplicities: quantiles over distinct values
            else:
                values = _column_uniques(X[:, f]) if X.shape[0] else np.zeros(0, X.dtype)
                weighted = X[:, f]
            if len(values) <= max_bins:
                edges.append(values)
                lossless[f] = True
            else:
                edges.append(_quantile_edges(values, weighted, max_bins))
        return cls(edges=edges, lossless=lossless, max_bins=max_bins, enumerated=lossless.copy())

    # ------------------------------------------------------------- wide bins
    def expand_wide(self, X: np.ndarray):
        """Codes for features with more than 256 bins, as several one-byte *virtual* columns.

        The kernels (and the slabs) are built around one-byte bin codes.  A feature with ``U > 256`` edges
        is therefore presented as ``ceil(U / 255)`` columns: column ``p`` holds ``clip(code - 255 p, 0, 255)``.
        A split of that column at local bin ``b <= 254`` is exactly the split ``code <= 255 p + b`` of the
        real feature (smaller codes clip to 0 and go left, larger ones clip to 255 and go right; local bin
        255 has an empty right side and is never chosen), every global threshold appears in exactly one
        column, and the columns' order (feature, then chunk, then bin) is the reference's tie-break order
        (first feature, first threshold).  So the histogram search over the virtual columns is the exact
        search over all ``U`` candidates; it only costs more columns.

        Returns ``(vcodes uint8 (n, Fv), v_feature int (Fv,), v_chunk int (Fv,), v_nbins int32 (Fv,))``.
        """
        n, F = X.shape
        cols, v_feature, v_chunk, v_nbins = [], [], [], []
        for f in range(F):
            e = self.edges[f]
            U = max(len(e), 1)
            code = np.zeros(n, dtype=np.int64) if len(e) == 0 else np.minimum(np.searchsorted(e, X[:, f], side="left"), U - 1)
            for p in range((U + 254) // 255) if U > 256 else range(1):
                lo = 255 * p
                cols.append(np.clip(code - lo, 0, 255).astype(np.uint8))
                v_feature.append(f)
                v_chunk.append(p)
                v_nbins.append(min(256, U - lo) if U > 256 else U)
        return (np.stack(cols, axis=1) if cols else np.zeros((n, 0), np.uint8), np.asarray(v_feature), np.asarray(v_chunk),
                np.asarray(v_nbins, dtype=np.int32))

    @staticmethod
    def local_summary(X: np.ndarray, cap: int = 4096) -> List[np.ndarray]:
        """Per-feature value summary small enough to all-gather between ranks."""
        out = []
        for f in range(X.shape[1]):
            u = _column_uniques(X[:, f]) if X.shape[0] else np.zeros(0, X.dtype)
            if len(u) > cap:
                idx = np.unique(np.linspace(0, len(u) - 1, cap).round().astype(np.int64))
                u = u[idx]
            out.append(u)
        return out

    # ------------------------------------------------------------- transform
    @property
    def n_features(self) -> int:
        return len(self.edges)

    @property
    def n_bins(self) -> np.ndarray:
        return np.array([max(len(e), 1) for e in self.edges], dtype=np.int32)

    @property
    def is_numeric(self) -> bool:
        return all(e.dtype != object for e in self.edges)

    def transform(self, X: np.ndarray) -> np.ndarray:
        """Raw values -> ``uint8`` codes ``(n_rows, n_features)`` (row-major)."""
        n, F = X.shape
        if F != self.n_features:
            raise ValueError(f"X has {F} features, the bin mapper was fitted on {self.n_features}")
        codes = np.empty((n, F), dtype=np.uint8)
        for f in range(F):
            e = self.edges[f]
            if len(e) == 0:
                codes[:, f] = 0
                continue
            col = X[:, f]
            if e.dtype == object:
                lut = {v: i for i, v in enumerate(e.tolist())}
                try:
                    codes[:, f] = [lut[v] for v in col.tolist()]
                except KeyError:
                    codes[:, f] = [_bisect_obj(e, v) for v in col.tolist()]
            elif col.dtype == np.uint8:
                # 256 possible inputs: encode the value range once and look rows up
                lut = np.minimum(np.searchsorted(e, np.arange(256), side="left"), len(e) - 1).astype(np.uint8)
                codes[:, f] = lut[col]
            else:
                c = np.searchsorted(e, col, side="left")
                np.minimum(c, len(e) - 1, out=c)
                codes[:, f] = c
        return codes

    def threshold_of(self, feature: int, bin_code: int):
        """Raw-value threshold for ``code <= bin_code`` as a plain Python object."""
        v = self.edges[feature][bin_code]
        return v.item() if isinstance(v, np.generic) else v

    def thresholds_of(self, feature: np.ndarray, bin_code: np.ndarray) -> list:
        """:meth:`threshold_of` for whole node arrays (``feature < 0`` marks a leaf -> ``None``).

        One gather (one per feature when the features' value types differ) instead of one Python-level lookup per node: for a
        2,047-node tree this is tens of microseconds instead of more than a millisecond, which is
        visible next to a 10 ms multi-GPU fit.
        """
        feature = np.asarray(feature)
        bin_code = np.asarray(bin_code)
        out = np.full(feature.shape[0], None, dtype=object)
        inner = np.flatnonzero(feature >= 0)
        dtypes = {e.dtype for e in self.edges}
        if len(dtypes) == 1 and next(iter(dtypes)) != object:
            # one dtype for every feature: a single gather from the padded (F, max bins) edge matrix
            table = self.__dict__.get("_edge_table")
            if table is None:
                table = np.zeros((len(self.edges), max(max(len(e) for e in self.edges), 1)), dtype=next(iter(dtypes)))
                for f, e in enumerate(self.edges):
                    table[f, : len(e)] = e
                self.__dict__["_edge_table"] = table
            out[inner] = table[feature[inner], bin_code[inner]].tolist()  # tolist: plain Python objects, like .item()
            return out.tolist()
        for f in np.unique(feature[inner]):
            idx = np.flatnonzero(feature == f)
            out[idx] = self.edges[int(f)][bin_code[idx]].tolist()
        return out.tolist()

    def edges_matrix(self, dtype=np.float64) -> np.ndarray:
        """(n_features, 256) array of edges padded with +inf (device encode table)."""
        out = np.full((self.n_features, 256), np.inf, dtype=dtype)
        for f, e in enumerate(self.edges):
            if e.dtype == object:
                raise TypeError("object/string features are encoded on the host")
            out[f, : len(e)] = e
        return out


def _quantile_edges(uniques: np.ndarray, col: np.ndarray, max_bins: int) -> np.ndarray:
    """``max_bins`` ascending data values; the last edge is the column maximum."""
    if col.dtype == object:
        idx = np.linspace(0, len(uniques) - 1, max_bins).round().astype(np.int64)
        return uniques[np.unique(idx)]
    # "higher" order statistics (every edge is an actual data value): element ceil(q (n-1)) of the
    # sorted column, which is what np.quantile(..., method="higher") returns, without its overhead
    qs = np.linspace(0.0, 1.0, max_bins + 1)[1:]
    pos = np.ceil(qs * (len(col) - 1)).astype(np.int64)
    e = np.unique(np.partition(col, np.unique(pos))[pos])
    if e[-1] < uniques[-1]:
        e[-1] = uniques[-1]
    return e.astype(col.dtype, copy=False)


def _bisect_obj(edges: np.ndarray, v) -> int:
    lo, hi = 0, len(edges) - 1
    while lo < hi:
        mid = (lo + hi) // 2
        if edges[mid] >= v:
            hi = mid
        else:
            lo = mid + 1
    return lo
