"""Subtree hand-off: finish small subtrees on rank subgroups (host logic).

The reference's only parallel strategy is task parallelism: the communicator is
halved by rank parity at every internal node, even ranks recurse into the
``<=`` child and odd ranks into the ``>`` child, and finished subtrees come back
through a pickled ``allgather`` (``mpitree/tree/decision_tree.py:447-477``,
parity split ``:336-338``).  In this framework that idea is phase 2 of the
level-wise scheduler: while nodes are large, all ranks cooperate on every node
(row-sharded histograms); once a frontier node's *global* row count drops below
a threshold, its rows are gathered onto one owner and the whole subtree is
finished there without further communication.  Owners are chosen by sized
bin-packing rather than by parity so load follows work, and the finished
subtrees return as flat arrays (:class:`FlatTree` rows), not pickled graphs.

This module holds the policy and the stitching, which are engine independent
and unit-testable on CPU; the GPU row movement itself is kernel K6
(``k_gather_rows`` in ``mpitree_b200/csrc/subtree.cu``, driven by ``ops/subtree_phase.py``).
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

import numpy as np

from ..models.flat_tree import FlatTree

__all__ = ["assign_owners", "halving_groups", "stitch_subtrees", "default_handoff_rows", "Graft"]


def default_handoff_rows(n_features: int, n_bins: int, n_classes: int, world: int) -> int:
    """Row threshold below which a node is cheaper to move than to all-reduce.

    Moving a node costs ``rows * (F + 1)`` bytes once; keeping it distributed
    costs one ``F * B * C * 4``-byte histogram exchange per remaining level.
    Rule of thumb from the design notes: hand off when ``rows < B * C * 4``.
    """
    if world <= 1:
        return 0
    return int(n_bins) * int(n_classes) * 4


def assign_owners(node_rows: Sequence[int], world: int) -> List[int]:
    """Greedy longest-processing-time packing of nodes onto ``world`` owners.

    Deterministic (ties broken by node order then by lowest rank) so every rank
    computes the same assignment without communicating.
    """
    load = [0] * world
    owners = [0] * len(node_rows)
    order = sorted(range(len(node_rows)), key=lambda i: (-int(node_rows[i]), i))
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        owners[i] = r
        load[r] += max(int(node_rows[i]), 1)
    return owners


def halving_groups(ranks: Sequence[int]) -> Tuple[List[int], List[int]]:
    """The reference's parity split of a rank group (``decision_tree.py:336-338``).

    ``Split(color=rank % 2, key=rank // 2)``: even positions take the ``<=``
    child, odd positions the ``>`` child.  The host scheduler halves its communicator this
    way when nodes become too small to be worth sharing among all ranks
    (``scheduler.fit_flat_tree_host``, phase 2); which nodes go to which half is decided by
    size (:func:`assign_owners` over two bins), not by the side of the split.
    """
    ranks = list(ranks)
    return ranks[0::2], ranks[1::2]


@dataclass
class Graft:
    """A finished subtree to be attached at leaf ``node_id`` of the trunk."""

    node_id: int
    subtree: FlatTree


def stitch_subtrees(trunk: FlatTree, grafts: Sequence[Graft]) -> FlatTree:
    """Attach finished subtrees to the trunk and renumber breadth-first.

    Every rank calls this with the same inputs and obtains the identical tree
    (deterministic numbering replaces the reference's "last writer wins" stitch
    at ``decision_tree.py:468-475``).
    """
    if not grafts:
        return trunk
    graft_at: Dict[int, FlatTree] = {g.node_id: g.subtree for g in grafts}
    C = trunk.counts.shape[1]

    # walk the combined tree breadth-first; entries are (source, local id)
    order: List[Tuple[int, int]] = []
    parent_new: List[int] = []
    queue: List[Tuple[int, int, int]] = [(-1, 0, -1)]  # (source, local id, new parent)
    src_tree = lambda s: trunk if s < 0 else graft_at[s]  # noqa: E731
    child_slots: List[Tuple[int, int]] = []
    while queue:
        nxt: List[Tuple[int, int, int]] = []
        for s, i, p in queue:
            if s < 0 and i in graft_at and trunk.feature[i] < 0:
                s, i = i, 0  # jump into the grafted subtree's root
            new_id = len(order)
            order.append((s, i))
            parent_new.append(p)
            t = src_tree(s)
            if t.feature[i] >= 0:
                nxt.append((s, int(t.left[i]), new_id))
                nxt.append((s, int(t.right[i]), new_id))
        queue = nxt

    m = len(order)
    out = FlatTree(
        feature=np.full(m, -1, np.int32),
        bin=np.full(m, -1, np.int32),
        left=np.full(m, -1, np.int32),
        right=np.full(m, -1, np.int32),
        parent=np.asarray(parent_new, np.int32),
        depth=np.zeros(m, np.int32),
        counts=np.zeros((m, C), np.int64),
        gain=np.full(m, np.nan),
        thresholds=[None] * m,
        class_labels=trunk.class_labels,
    )
    seen_children: Dict[int, int] = {}
    for new_id, (s, i) in enumerate(order):
        t = src_tree(s)
        out.feature[new_id] = t.feature[i]
        out.bin[new_id] = t.bin[i]
        out.counts[new_id] = t.counts[i]
        out.gain[new_id] = t.gain[i]
        out.thresholds[new_id] = t.thresholds[i] if t.thresholds else None
        p = parent_new[new_id]
        if p >= 0:
            out.depth[new_id] = out.depth[p] + 1
            if seen_children.get(p, 0) == 0:
                out.left[p] = new_id
            else:
                out.right[p] = new_id
            seen_children[p] = seen_children.get(p, 0) + 1
    del child_slots
    return out
