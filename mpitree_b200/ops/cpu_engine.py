"""Host (numpy) engine: exact split search and the level-wise builder.

This is the correctness scaffold and the CPU product path (small inputs, the
``iris / 2 ranks on CPU`` parity configuration, the CPU half of subtree
hand-off).  It runs the *same algorithm* as the CUDA engine — per-node
feature x bin x class count histograms, cross-rank sum, prefix over bins,
entropy gain, argmax — and reproduces the reference's choices bit for bit:

* criterion: Shannon entropy in base 2, ``-sum(p * log2 p)`` over the classes
  present, float64 (reference ``mpitree/tree/decision_tree.py:37-51``);
* split cost: ``np.dot([nL/n, nR/n], [H(yL), H(yR)])`` (``:79-86``); gain is
  ``H(y) - min(cost)`` and the threshold is the *first* arg-min over ascending
  candidates (``:88-91``); the feature is the *first* arg-max gain (``:140``);
* candidates: every distinct value present in the node (``:73``), except the
  largest one, whose right side is empty — the reference can only crash after
  choosing it (``:150-164``), so dropping it changes no successful fit;
* stop rules (``:118-123``): one class, all rows identical (== no feature has
  two non-empty bins), ``depth == max_depth``, ``n < min_samples_split``.

Bit-exactness strategy: costs are first evaluated vectorised, then every
candidate within a tiny tolerance of the best is re-evaluated with the same
scalar numpy calls the reference makes, so rounding-level ties resolve the same
way without paying a Python loop over all candidates.
"""

from __future__ import annotations

import functools
from dataclasses import dataclass
from fractions import Fraction
from typing import List, Optional, Tuple

import numpy as np

from ..models.flat_tree import FlatTree

__all__ = [
    "exact_entropy",
    "exact_cost",
    "best_split",
    "best_split_many",
    "best_split_rows",
    "best_split_rows_many",
    "build_tree_cpu",
    "klogk",
    "klogk_constants",
    "probe_dot_mode",
    "SplitParams",
]

_TOL = 1e-9
_BATCH_CELLS = 1 << 20  # histogram cells evaluated per vectorised batch of nodes (8 MB of int64 temporaries)


@dataclass(frozen=True)
class SplitParams:
    max_depth: Optional[int] = None
    min_samples_split: int = 2
    criterion: str = "entropy"  # "entropy" (reference) or "gini" (extension)


# --------------------------------------------------------------------------
# float model probes (which rounding does this host's 2-element np.dot use?)
# --------------------------------------------------------------------------
@functools.lru_cache(maxsize=1)
def probe_dot_mode() -> int:
    """0: a*c + b*d, 1: fma(b, d, a*c), 2: fma(a, c, b*d) for a 2-element ``np.dot``.

    The device gain kernel is told which one the host BLAS uses so that GPU and
    reference costs round identically.
    """
    rng = np.random.default_rng(12345)
    score = [0, 0, 0]

    def fma(x, y, z):
        return float(Fraction(x) * Fraction(y) + Fraction(z))

    for _ in range(64):
        a, b, c, d = (float(v) for v in rng.random(4))
        r = float(np.dot(np.array([a, b]), np.array([c, d])))
        score[0] += r == a * c + b * d
        score[1] += r == fma(b, d, a * c)
        score[2] += r == fma(a, c, b * d)
    return int(np.argmax(score))


# --------------------------------------------------------------------------
# exact scalar formulas (same numpy calls as the reference)
# --------------------------------------------------------------------------
def exact_entropy(counts: np.ndarray) -> float:
    """Entropy of a class-count vector, computed like ``_compute_entropy``."""
    present = counts[counts > 0]
    n = int(present.sum())
    proba = present / n
    return -np.sum(proba * np.log2(proba))


def exact_gini(counts: np.ndarray) -> float:
    present = counts[counts > 0]
    n = int(present.sum())
    proba = present / n
    return 1.0 - np.sum(proba * proba)


def exact_cost(left: np.ndarray, right: np.ndarray, criterion: str = "entropy") -> float:
    """Weighted child impurity, computed like ``_compute_information_gain``."""
    imp = exact_entropy if criterion == "entropy" else exact_gini
    n_left, n_right = int(left.sum()), int(right.sum())
    n = n_left + n_right
    weights = np.array([n_left / n, n_right / n])
    impurity = np.array([imp(left), imp(right)])
    return np.dot(weights, impurity)


def _exact_cost_cached(cache: dict, left: np.ndarray, right: np.ndarray, criterion: str) -> float:
    """``exact_cost`` memoised on the left count vector (the right one is ``total - left``).

    Near-tied candidates of a small node very often share their class counts (the same rows end up
    on the left for several features or thresholds), and equal counts give the same float64 cost.
    """
    key = left.tobytes()
    hit = cache.get(key)
    if hit is None:
        hit = cache[key] = exact_cost(left, right, criterion)
    return hit


# --------------------------------------------------------------------------
# vectorised approximate impurity
# --------------------------------------------------------------------------
def _impurity_rows(counts: np.ndarray, totals: np.ndarray, criterion: str) -> np.ndarray:
    """Impurity of each row of ``counts`` (..., C) with row sums ``totals`` (...)."""
    safe = np.where(totals > 0, totals, 1).astype(np.float64)
    p = counts / safe[..., None]
    if criterion == "entropy":
        with np.errstate(divide="ignore", invalid="ignore"):
            t = np.where(counts > 0, p * np.log2(np.where(counts > 0, p, 1.0)), 0.0)
        return -t.sum(axis=-1)
    return 1.0 - (p * p).sum(axis=-1)


_KLOGK_IDX_BITS = 7


@functools.lru_cache(maxsize=1)
def klogk_constants() -> np.ndarray:
    """Constants of the device ``k * log2(k)`` evaluation (``csrc/impurity.cuh::KlogK``), device layout:
    ``inv[128] = 1/m0`` | ``el[32][128] = e + log2 m0`` | ``c1..c6`` for interval centres
    ``m0 = 1 + (i + 1/2)/128`` and exponents ``e = 0..31``."""
    n = 1 << _KLOGK_IDX_BITS
    m0 = 1.0 + (np.arange(n, dtype=np.float64) + 0.5) / n  # exact
    ln2 = float(np.log(2.0))
    el = np.arange(32, dtype=np.float64)[:, None] + np.log2(m0)[None, :]
    coef = [(1.0 if j % 2 else -1.0) / (j * ln2) for j in range(1, 7)]
    return np.concatenate([1.0 / m0, el.ravel(), np.asarray(coef)])


def klogk(k: np.ndarray) -> np.ndarray:
    """``k * log2(k)`` for unsigned 32-bit counts by the device's formula (numpy twin of
    ``csrc/impurity.cuh::KlogK``; the device fuses the Horner steps, so results agree to a few ulp).

    ``floor(log2 k)`` comes from the exponent of ``float(k)``; the top 7 mantissa bits pick an interval
    centre ``m0``; ``r = (m - m0) * fl(1/m0)``, ``|r| <= 2**-8``; ``log2(1 + r)`` is a degree-6 Horner
    polynomial.  ``g(0) = g(1) = 0``.
    """
    n = 1 << _KLOGK_IDX_BITS
    tab = klogk_constants()
    inv, el, c = tab[:n], tab[n : n + 32 * n], tab[n + 32 * n :]
    k = np.asarray(k, dtype=np.uint32)
    kd = k.astype(np.float64)
    bits = kd.view(np.int64)
    hi = bits >> 32
    top = (hi >> (20 - _KLOGK_IDX_BITS)) - (1023 << _KLOGK_IDX_BITS)
    idx = top & (n - 1)
    m = ((bits & 0x000FFFFFFFFFFFFF) | 0x3FF0000000000000).view(np.float64)
    m0 = ((0x3FF00000 | (idx << (20 - _KLOGK_IDX_BITS)) | (1 << (19 - _KLOGK_IDX_BITS))) << 32).view(np.float64)
    r = (m - m0) * inv[idx]
    p = c[5] * r + c[4]
    for j in (3, 2, 1, 0):
        p = p * r + c[j]
    out = kd * (p * r + el[np.clip(top, 0, 32 * n - 1)])
    return np.where(k <= 1, 0.0, out)


_KLOGK_MAX = 1 << 23  # table of k*log2(k) up to 8M rows per node (64 MB); larger nodes use log2 directly
_klogk_cache = {"table": np.zeros(1)}


def _klogk_table(n: int) -> Optional[np.ndarray]:
    """``g[k] = k * log2(k)`` for ``k = 0 .. n`` (``g[0] = 0``), grown on demand; None when ``n`` is too large."""
    if n > _KLOGK_MAX:
        return None
    t = _klogk_cache["table"]
    if t.shape[0] <= n:
        size = min(_KLOGK_MAX, max(n, 2 * t.shape[0])) + 1
        k = np.arange(size, dtype=np.float64)
        t = np.zeros(size)
        np.multiply(k[1:], np.log2(k[1:]), out=t[1:])
        _klogk_cache["table"] = t
    return t


def _approx_cost(left: np.ndarray, right: np.ndarray, n_left: np.ndarray, n_right: np.ndarray, n: int,
                 criterion: str) -> np.ndarray:
    """Weighted child impurity of every candidate, good to ~1e-13 (ties are settled by exact_cost).

    Entropy uses ``n*cost = g(nL) + g(nR) - sum g(kL) - sum g(kR)`` with ``g(k) = k log2 k`` (the same
    algebra as the device kernel's large-node path, ``csrc/impurity.cuh``): table gathers and sums, no
    division, logarithm or masking per cell.
    """
    if criterion == "entropy":
        g = _klogk_table(int(np.max(n)))  # ``n`` may be one count or an array broadcast against the candidates
        if g is not None:
            return (g[n_left] + g[n_right] - g[left].sum(axis=-1) - g[right].sum(axis=-1)) / n
    return (n_left / n) * _impurity_rows(left, n_left, criterion) + (n_right / n) * _impurity_rows(right, n_right, criterion)


def best_split(
    hist: np.ndarray, n_bins: np.ndarray, criterion: str = "entropy"
) -> Optional[Tuple[int, int, float, np.ndarray]]:
    """Best ``(feature, bin, gain, left_counts)`` of one node, or ``None``.

    ``hist`` is the node's (F, B, C) count histogram.  ``None`` means no feature
    has two non-empty bins, i.e. all rows are identical -> the node is a leaf.
    """
    return best_split_many(hist[None], n_bins, criterion)[0]


def best_split_many(hist: np.ndarray, n_bins: np.ndarray, criterion: str = "entropy") -> list:
    """:func:`best_split` for a stack of nodes ``(S, F, B, C)``: the prefix sums and the candidate
    costs of all nodes are formed by the same array operations, only the tie resolution is per node."""
    S = hist.shape[0]
    group = max(1, _BATCH_CELLS // max(int(np.prod(hist.shape[1:])), 1))
    if S > group:  # keep the temporaries of one batch cache-resident
        out: list = []
        for s0 in range(0, S, group):
            out.extend(best_split_many(hist[s0 : s0 + group], n_bins, criterion))
        return out
    total = hist[:, 0].sum(axis=1)  # (S, C)
    n = total.sum(axis=1)  # (S,)
    left = np.cumsum(hist, axis=2)  # (S, F, B, C)
    per_bin = hist.sum(axis=-1)  # (S, F, B)
    n_left = np.cumsum(per_bin, axis=2)
    n_right = n[:, None, None] - n_left
    cand = (per_bin > 0) & (n_right > 0)
    right = total[:, None, None, :] - left
    safe_n = np.maximum(n, 1)[:, None, None]
    cost = np.where(cand, _approx_cost(left, right, n_left, n_right, safe_n, criterion), np.inf)
    fmin_all = cost.min(axis=2)  # (S, F)
    return [
        _resolve_ties(cost[s], fmin_all[s], left[s], right[s], total[s], criterion) if cand[s].any() else None
        for s in range(S)
    ]


def _resolve_ties(cost, fmin, left, right, total, criterion):
    """Exact re-evaluation of the candidates within rounding distance of the best approximate cost."""
    best_approx = fmin.min()
    tol = _TOL * (1.0 + abs(best_approx))

    imp = exact_entropy if criterion == "entropy" else exact_gini
    h_parent = imp(total)
    best = None  # (gain, feature, bin)
    cache: dict = {}
    for f in np.flatnonzero(fmin <= best_approx + tol):
        bins = np.flatnonzero(cost[f] <= fmin[f] + tol)
        costs = [_exact_cost_cached(cache, left[f, b], right[f, b], criterion) for b in bins]
        k = int(np.argmin(costs))  # first arg-min over ascending thresholds
        gain = h_parent - costs[k]
        if best is None or gain > best[0]:  # strict: first arg-max over features
            best = (gain, int(f), int(bins[k]))
    gain, f, b = best
    return f, b, float(gain), left[f, b].copy()


def best_split_rows(
    codes: np.ndarray, y: np.ndarray, n_classes: int, criterion: str = "entropy"
) -> Optional[Tuple[int, int, float, np.ndarray]]:
    """Same result as :func:`best_split`, computed from the node's rows instead of a dense histogram.

    For a node with few rows the (F, B, C) histogram is mostly zeros; sorting the node's rows per
    feature and taking class prefix counts at the distinct values costs ``O(F m log m)`` instead of
    ``O(F B C)``.  All features are handled by the same few array operations (one stable argsort of
    the ``(m, F)`` code block, one cumulative sum).  Candidates, tie-breaks and the exact
    re-evaluation are identical.
    """
    m, F = codes.shape
    if m < 2:
        return None
    order = np.argsort(codes, axis=0, kind="stable")  # (m, F)
    cs = np.take_along_axis(codes, order, axis=0)
    boundary = cs[1:] != cs[:-1]  # (m-1, F): sorted position i is the last row of its value, a larger value follows
    if not boundary.any():
        return None  # every feature is constant in this node
    present, yi = np.unique(y, return_inverse=True)
    Cp = len(present)
    total = np.bincount(yi, minlength=Cp).astype(np.int64)
    onehot = np.zeros((m, Cp), dtype=np.int64)
    onehot[np.arange(m), yi] = 1
    left = np.cumsum(onehot[order], axis=0)[:-1]  # (m-1, F, Cp) class counts of sorted positions <= i
    right = total - left
    n_left = np.broadcast_to(np.arange(1, m, dtype=np.int64)[:, None], (m - 1, F))
    cost = np.where(boundary, _approx_cost(left, right, n_left, m - n_left, m, criterion), np.inf)
    fmin = cost.min(axis=0)
    best_approx = float(fmin.min())
    tol = _TOL * (1.0 + abs(best_approx))
    imp = exact_entropy if criterion == "entropy" else exact_gini
    h_parent = imp(total)
    best = None
    cache: dict = {}
    for f in np.flatnonzero(fmin <= best_approx + tol):  # ascending feature index
        near = np.flatnonzero(cost[:, f] <= fmin[f] + tol)  # ascending position == ascending threshold
        costs = [_exact_cost_cached(cache, left[i, f], right[i, f], criterion) for i in near]
        k = int(np.argmin(costs))
        gain = h_parent - costs[k]
        if best is None or gain > best[0]:
            best = (gain, int(f), int(cs[near[k], f]), left[near[k], f])
    gain, f, b, left_present = best
    left_counts = np.zeros(n_classes, dtype=np.int64)
    left_counts[present] = left_present
    return f, b, float(gain), left_counts


def best_split_rows_many(codes: np.ndarray, y: np.ndarray, n_classes: int, criterion: str = "entropy") -> list:
    """:func:`best_split_rows` for ``S`` nodes that all hold ``m`` rows: ``codes`` (S, m, F), ``y`` (S, m).

    Deep levels consist of very many small nodes; stacking the nodes of equal size turns their sorts,
    prefix counts and candidate costs into one set of array operations.  Tie resolution stays per node.
    """
    S, m, F = codes.shape
    if m < 2:
        return [None] * S
    group = max(1, _BATCH_CELLS // max(m * F * n_classes, 1))
    if S > group:
        out: list = []
        for s0 in range(0, S, group):
            out.extend(best_split_rows_many(codes[s0 : s0 + group], y[s0 : s0 + group], n_classes, criterion))
        return out
    order = np.argsort(codes, axis=1, kind="stable")  # (S, m, F)
    cs = np.take_along_axis(codes, order, axis=1)
    boundary = cs[:, 1:] != cs[:, :-1]  # (S, m-1, F)
    onehot = np.zeros((S, m, n_classes), dtype=np.int64)
    onehot[np.arange(S)[:, None], np.arange(m)[None, :], y] = 1
    total = onehot.sum(axis=1)  # (S, C)
    left = np.cumsum(onehot[np.arange(S)[:, None, None], order], axis=1)[:, :-1]  # (S, m-1, F, C)
    right = total[:, None, None, :] - left
    n_left = np.broadcast_to(np.arange(1, m, dtype=np.int64)[None, :, None], (S, m - 1, F))
    cost = np.where(boundary, _approx_cost(left, right, n_left, m - n_left, m, criterion), np.inf)
    fmin_all = cost.min(axis=1)  # (S, F)
    imp = exact_entropy if criterion == "entropy" else exact_gini
    results = []
    for s in range(S):
        if not boundary[s].any():
            results.append(None)  # every feature is constant in this node
            continue
        fmin = fmin_all[s]
        best_approx = float(fmin.min())
        tol = _TOL * (1.0 + abs(best_approx))
        h_parent = imp(total[s])
        best = None
        cache: dict = {}
        for f in np.flatnonzero(fmin <= best_approx + tol):  # ascending feature index
            near = np.flatnonzero(cost[s, :, f] <= fmin[f] + tol)  # ascending position == ascending threshold
            costs = [_exact_cost_cached(cache, left[s, i, f], right[s, i, f], criterion) for i in near]
            k = int(np.argmin(costs))
            gain = h_parent - costs[k]
            if best is None or gain > best[0]:
                best = (gain, int(f), int(cs[s, near[k], f]), left[s, near[k], f].copy())
        results.append((best[1], best[2], float(best[0]), best[3]))
    return results


# --------------------------------------------------------------------------
# level-wise builder
# --------------------------------------------------------------------------
def _is_terminal(counts: np.ndarray, depth: int, params: SplitParams) -> bool:
    n = int(counts.sum())
    return bool(
        (counts > 0).sum() <= 1
        or (params.max_depth is not None and depth >= params.max_depth)
        or n < params.min_samples_split
    )


def _level_histograms(codes, y, slot, n_slots, B, C, comm) -> np.ndarray:
    """(n_slots, F, B, C) int64 histogram of the rows with ``slot >= 0``."""
    F = codes.shape[1]
    hist = np.zeros((n_slots, F, B, C), dtype=np.int64)
    rows = np.flatnonzero(slot >= 0)
    if rows.size:
        # one bincount per feature over that feature's (slot, bin, class) cells only
        base = slot[rows].astype(np.int64) * (B * C) + y[rows]
        sub = codes[rows]
        cells = n_slots * B * C
        for f in range(F):
            idx = base + sub[:, f].astype(np.int64) * C
            hist[:, f] = np.bincount(idx, minlength=cells).reshape(n_slots, B, C)
    if comm is not None and comm.size > 1:
        hist = comm.allreduce_sum(hist)
    return hist


def build_tree_cpu(
    codes: np.ndarray,
    y: np.ndarray,
    n_bins: np.ndarray,
    n_classes: int,
    params: SplitParams,
    comm=None,
    *,
    max_level_entries: int = 1 << 24,
    handoff_rows: int = 0,
    pending_out: Optional[dict] = None,
) -> FlatTree:
    """Grow a tree level by level from this rank's row shard.

    ``codes`` (n_local, F) uint8 bin codes, ``y`` (n_local,) class codes.  With
    a communicator of size > 1 every rank ends with the identical ``FlatTree``
    (counts are global).  Node ids are breadth-first.

    With ``handoff_rows > 0`` a non-terminal child whose global row count is
    ``<= handoff_rows`` is not expanded here: it stays a leaf of the returned
    trunk and ``pending_out[node_id]`` receives this rank's local row indices
    for it (phase 2 of the scheduler finishes those subtrees on owner ranks).
    """
    codes = np.ascontiguousarray(codes)
    y = np.asarray(y, dtype=np.int64)
    n_local, F = codes.shape
    B = int(max(int(np.max(n_bins)) if len(n_bins) else 1, 1))
    C = int(n_classes)

    root_counts = np.bincount(y, minlength=C).astype(np.int64)
    if comm is not None and comm.size > 1:
        root_counts = comm.allreduce_sum(root_counts)

    feature, bin_, left_, right_, parent, depth, counts, gain = [], [], [], [], [], [], [], []

    def new_node(par: int, dep: int, cnt: np.ndarray) -> int:
        feature.append(-1)
        bin_.append(-1)
        left_.append(-1)
        right_.append(-1)
        parent.append(par)
        depth.append(dep)
        counts.append(cnt)
        gain.append(np.nan)
        return len(feature) - 1

    new_node(-1, 0, root_counts)
    slot = np.zeros(n_local, dtype=np.int32)
    pending_ids: List[int] = []  # node ids handed off; rows carry slot = -2 - index
    frontier: List[int] = [] if _is_terminal(root_counts, 0, params) else [0]
    if not frontier:
        slot[:] = -1

    per_node = max(F * B * C, 1)
    chunk = max(1, max_level_entries // per_node)

    while frontier:
        # (node id -> decision) for this level, histogram built in bounded chunks
        decisions = {}
        dense = list(range(len(frontier)))
        if comm is None or comm.size <= 1:
            # nodes much smaller than a dense (F, B, C) histogram are searched from their rows
            active = np.flatnonzero(slot >= 0)
            order = active[np.argsort(slot[active], kind="stable")]
            bounds = np.searchsorted(slot[order], np.arange(len(frontier) + 1))
            dense = []
            by_size: dict = {}  # row count -> slots of that size (searched together)
            for s_ in range(len(frontier)):
                m_ = int(bounds[s_ + 1] - bounds[s_])
                if m_ * 8 < B * C:
                    by_size.setdefault(m_, []).append(s_)
                else:
                    dense.append(s_)
            for m_, slots_ in by_size.items():
                rows_ = order[(bounds[slots_][:, None] + np.arange(m_)[None, :]).reshape(-1)]
                found = best_split_rows_many(
                    codes[rows_].reshape(len(slots_), m_, F), y[rows_].reshape(len(slots_), m_), C, params.criterion
                )
                for s_, dec in zip(slots_, found):
                    decisions[frontier[s_]] = dec
        for c0 in range(0, len(dense), chunk):
            part = dense[c0 : c0 + chunk]
            remap = np.full(len(frontier) + 1, -1, dtype=np.int64)
            remap[part] = np.arange(len(part))
            local_slot = remap[np.where(slot >= 0, slot, len(frontier))]
            hist = _level_histograms(codes, y, local_slot, len(part), B, C, comm)
            for s_, dec in zip(part, best_split_many(hist, n_bins, params.criterion)):
                decisions[frontier[s_]] = dec

        next_frontier: List[int] = []
        # per current slot: (feature, bin, left slot or -1, right slot or -1)
        route_f = np.full(len(frontier), -1, dtype=np.int64)
        route_b = np.zeros(len(frontier), dtype=np.int64)
        route_l = np.full(len(frontier), -1, dtype=np.int32)
        route_r = np.full(len(frontier), -1, dtype=np.int32)
        for s, nid in enumerate(frontier):
            dec = decisions[nid]
            if dec is None:
                continue  # all rows identical -> stays a leaf
            f, b, g, left_counts = dec
            right_counts = counts[nid] - left_counts
            feature[nid], bin_[nid], gain[nid] = f, b, g
            d = depth[nid] + 1
            lid = new_node(nid, d, left_counts)
            rid = new_node(nid, d, right_counts)
            left_[nid], right_[nid] = lid, rid
            route_f[s], route_b[s] = f, b
            for cid, cnt, route in ((lid, left_counts, route_l), (rid, right_counts, route_r)):
                if _is_terminal(cnt, d, params):
                    continue
                if handoff_rows and int(cnt.sum()) <= handoff_rows:
                    route[s] = -2 - len(pending_ids)
                    pending_ids.append(cid)
                else:
                    route[s] = len(next_frontier)
                    next_frontier.append(cid)

        rows = np.flatnonzero(slot >= 0)
        if rows.size:
            s = slot[rows]
            f = route_f[s]
            split = f >= 0
            go_left = np.zeros(rows.size, dtype=bool)
            go_left[split] = codes[rows[split], f[split]] <= route_b[s[split]]
            new_slot = np.where(go_left, route_l[s], route_r[s])
            new_slot[~split] = -1
            slot[rows] = new_slot
        frontier = next_frontier

    if pending_out is not None:
        for k, nid in enumerate(pending_ids):
            pending_out[nid] = np.flatnonzero(slot == -2 - k)

    m = len(feature)
    return FlatTree(
        feature=np.asarray(feature, np.int32),
        bin=np.asarray(bin_, np.int32),
        left=np.asarray(left_, np.int32),
        right=np.asarray(right_, np.int32),
        parent=np.asarray(parent, np.int32),
        depth=np.asarray(depth, np.int32),
        counts=np.asarray(counts, np.int64).reshape(m, C),
        gain=np.asarray(gain, np.float64),
        thresholds=[None] * m,
    )
