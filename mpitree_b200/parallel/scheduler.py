"""Level-wise scheduler: engine selection, phase 1 (cooperative) + phase 2 (hand-off).

The reference builds depth-first with Python recursion and parallelises only by
giving subtrees to halves of the communicator
(``mpitree/tree/decision_tree.py:93-166`` serial, ``:364-479`` parallel).  This
scheduler is breadth-first:

1. **phase 1 — cooperative levels.**  All frontier nodes of a depth are
   processed together; every rank contributes the count histogram of its own
   row shard and the cross-rank sum gives every rank the same split decisions.
   On GPUs one fused kernel per level does histogram + peer reduction + gain +
   arg-max, one kernel partitions the rows (see :mod:`mpitree_b200.ops.cuda_engine`).
2. **phase 2 — rank subgroups** (host engine, four ranks or more).  Frontier nodes
   whose global row count fell below ``group_rows`` leave the all-ranks build: the
   communicator is halved like the reference's (``halving_groups``), the nodes are
   dealt to the halves by size, their rows move there, and each half continues
   cooperatively on its own sub-communicator, halving again further down.
3. **phase 3 — hand-off.**  Nodes below ``handoff_rows`` are assigned to single
   owner ranks (sized bin-packing), their rows are gathered there, owners finish
   those subtrees alone, and the finished subtrees are exchanged as flat arrays
   and stitched with deterministic ids.  (On GPUs phase 1 goes straight to this
   phase: K6, ``ops/subtree_phase.py``.)
"""

from __future__ import annotations

from typing import Optional

import numpy as np

from ..models.flat_tree import FlatTree
from ..ops.cpu_engine import SplitParams, build_tree_cpu
from .handoff import Graft, assign_owners, default_handoff_rows, halving_groups, stitch_subtrees

__all__ = ["fit_flat_tree_host", "fit_flat_tree_replicated", "resolve_engine", "default_group_rows"]


def resolve_engine(device: Optional[str], *, numeric: bool, n_classes: int, is_cuda_input: bool = False) -> str:
    """Pick ``"cuda"`` or ``"cpu"``.  ``auto`` never silently falls back on a GPU box."""
    device = (device or "auto").lower()
    if device not in ("auto", "cpu", "cuda"):
        raise ValueError("device must be 'auto', 'cpu' or 'cuda'")
    if device == "cpu":
        if is_cuda_input:
            raise ValueError("device='cpu' was requested but X lives on a CUDA device")
        return "cpu"
    import torch

    have = torch.cuda.is_available()
    if device == "cuda" and not have:
        raise RuntimeError("device='cuda' requested but no CUDA device is available")
    if not have:
        return "cpu"
    from ..ops import cuda_engine

    ok, why = cuda_engine.supports(numeric=numeric, n_classes=n_classes)
    if not ok:
        if device == "cuda":
            raise RuntimeError(f"the CUDA engine cannot run this problem: {why}")
        return "cpu"
    return "cuda"


def default_group_rows(handoff_rows: int, world: int) -> int:
    """Global row count below which a node leaves the all-ranks build for a rank *subgroup* (0 = never).

    Between ``handoff_rows`` (one owner finishes the subtree alone) and this size a node is still worth
    sharing, but not among all ranks: every level of a cooperative build costs one histogram exchange over
    the whole communicator.  Subgroups only make sense from four ranks on (halves of two)."""
    return 0 if world < 4 or handoff_rows <= 0 else 4 * int(handoff_rows)


def fit_flat_tree_host(
    codes: np.ndarray,
    y: np.ndarray,
    n_bins: np.ndarray,
    n_classes: int,
    params: SplitParams,
    comm=None,
    *,
    handoff_rows: Optional[int] = None,
    group_rows: Optional[int] = None,
) -> FlatTree:
    """Host engine entry over ``comm`` (row-sharded input): all ranks, then rank subgroups, then single owners.

    1. while nodes are large, all ranks of ``comm`` cooperate on every node (row-sharded histograms);
    2. nodes with at most ``group_rows`` global rows leave that build: the communicator is halved the way
       the reference halves it (``halving_groups``: even positions / odd positions,
       ``decision_tree.py:336-338``), the nodes are dealt to the two halves by size, their rows move to
       the half that owns them (sharded over its members) and each half carries on *cooperatively* with
       the same procedure on its own sub-communicator (the reference's recursion ``:446-466``, with rows
       that really are divided instead of replicated);
    3. nodes with at most ``handoff_rows`` rows go to a single owner rank, which finishes the subtree alone.

    Finished subtrees come back as flat arrays and are stitched with deterministic numbering, so every rank
    ends with the identical tree.
    """
    world = 1 if comm is None else comm.size
    if handoff_rows is None:
        handoff_rows = default_handoff_rows(codes.shape[1], int(np.max(n_bins, initial=1)), n_classes, world)
    if world <= 1 or handoff_rows <= 0:
        return build_tree_cpu(codes, y, n_bins, n_classes, params, comm)
    if group_rows is None:
        group_rows = default_group_rows(handoff_rows, world)
    if world < 4 or group_rows <= handoff_rows:
        group_rows = 0

    # the whole problem is below the hand-off size: one owner builds it, everyone receives the tree
    n_global = int(comm.allreduce_sum(np.array([codes.shape[0]], dtype=np.int64))[0])
    if n_global <= handoff_rows:
        shards = comm.allgather_object((codes, y))
        blob = None
        if comm.rank == 0:
            all_codes = np.concatenate([s[0] for s in shards])
            all_y = np.concatenate([s[1] for s in shards])
            blob = build_tree_cpu(all_codes, all_y, n_bins, n_classes, params, None).to_wire()
        return FlatTree.from_wire(comm.broadcast_object(blob, src=0))

    pending: dict = {}
    trunk = build_tree_cpu(
        codes, y, n_bins, n_classes, params, comm, handoff_rows=max(handoff_rows, group_rows), pending_out=pending
    )
    node_ids = sorted(pending)
    if not node_ids:
        return trunk
    node_rows = [int(trunk.counts[nid].sum()) for nid in node_ids]

    def sub_params_of(nid: int) -> SplitParams:
        depth0 = int(trunk.depth[nid])
        return SplitParams(
            max_depth=None if params.max_depth is None else params.max_depth - depth0,
            min_samples_split=params.min_samples_split,
            criterion=params.criterion,
        )

    if group_rows:
        # phase 2 (subgroups): halve the communicator, deal the nodes to the halves by size, move the rows
        halves = halving_groups(range(world))
        half_of_node = assign_owners(node_rows, 2)
        my_half = 0 if comm.rank in halves[0] else 1
        outbox = {}
        for k, nid in enumerate(node_ids):
            if len(pending[nid]):
                members = halves[half_of_node[k]]
                # a source sends all its rows of a node to one member: the members' shares follow the sources'
                outbox[nid] = (members[comm.rank % len(members)], codes[pending[nid]], y[pending[nid]])
        inbox = comm.allgather_object(outbox)
        sub = comm.split(my_half, comm.rank)
        mine = []
        for k, nid in enumerate(node_ids):
            if half_of_node[k] != my_half:
                continue
            parts = [box[nid] for box in inbox if nid in box and box[nid][0] == comm.rank]
            sub_codes = np.concatenate([p[1] for p in parts]) if parts else codes[:0]
            sub_y = np.concatenate([p[2] for p in parts]) if parts else y[:0]
            # the half continues cooperatively (and halves again further down while it has four ranks or more)
            tree = fit_flat_tree_host(sub_codes, sub_y, n_bins, n_classes, sub_params_of(nid), sub,
                                      handoff_rows=handoff_rows, group_rows=group_rows // 2)
            if sub.rank == 0:  # every member holds the same subtree: one of them reports it
                mine.append((nid, tree))
        sub.free()
        return _exchange_and_stitch(trunk, mine, comm)

    # phase 3a: owners, decided identically on every rank from global counts
    owners = assign_owners(node_rows, world)

    # phase 3b: gather each pending node's rows on its owner
    outbox = {
        nid: (owners[k], codes[pending[nid]], y[pending[nid]])
        for k, nid in enumerate(node_ids)
        if len(pending[nid])
    }
    inbox = comm.allgather_object({nid: v for nid, v in outbox.items()})

    # phase 3c: owners finish their subtrees with no communication
    mine = []
    for k, nid in enumerate(node_ids):
        if owners[k] != comm.rank:
            continue
        parts = [box[nid] for box in inbox if nid in box]
        sub_codes = np.concatenate([p[1] for p in parts]) if parts else codes[:0]
        sub_y = np.concatenate([p[2] for p in parts]) if parts else y[:0]
        mine.append((nid, build_tree_cpu(sub_codes, sub_y, n_bins, n_classes, sub_params_of(nid), None)))

    # phase 3d: exchange finished subtrees (flat arrays) and stitch
    return _exchange_and_stitch(trunk, mine, comm)


def _exchange_and_stitch(trunk: FlatTree, mine, comm) -> FlatTree:
    everyone = comm.allgather_object([(nid, sub.to_wire()) for nid, sub in mine])
    grafts = [Graft(nid, FlatTree.from_wire(blob)) for box in everyone for nid, blob in box]
    grafts.sort(key=lambda g: g.node_id)
    return stitch_subtrees(trunk, grafts)


def fit_flat_tree_replicated(
    codes: np.ndarray, y: np.ndarray, n_bins: np.ndarray, n_classes: int, params: SplitParams, comm
) -> FlatTree:
    """Every rank holds *all* rows (the reference's data layout, ``:356``): no row ever moves.

    The top of the tree is grown redundantly on every rank, with no messages, until nodes are down
    to ``n / world`` rows (the reference also searches the top splits redundantly, ``:408-419``);
    those subtrees are then dealt out by size, finished independently from the local copy of the
    data, and exchanged with a single all-gather of flat arrays (the reference: pickled ``Node``
    graphs, ``:456-466``).  One collective per fit, and the deep, node-rich part of the tree — where
    the host engine spends most of its time — is divided among the ranks.
    """
    world = comm.size
    n = int(codes.shape[0])
    pending: dict = {}
    trunk = build_tree_cpu(
        codes, y, n_bins, n_classes, params, None, handoff_rows=max(n // world, 1), pending_out=pending
    )
    node_ids = sorted(pending)
    node_rows = [int(trunk.counts[nid].sum()) for nid in node_ids]
    owners = assign_owners(node_rows, world) if node_ids else []
    mine = []
    for k, nid in enumerate(node_ids):
        if owners[k] != comm.rank:
            continue
        rows = pending[nid]
        depth0 = int(trunk.depth[nid])
        sub_params = SplitParams(
            max_depth=None if params.max_depth is None else params.max_depth - depth0,
            min_samples_split=params.min_samples_split,
            criterion=params.criterion,
        )
        mine.append((nid, build_tree_cpu(codes[rows], y[rows], n_bins, n_classes, sub_params, None)))
    return _exchange_and_stitch(trunk, mine, comm)
