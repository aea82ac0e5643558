"""Object form of a fitted tree: ``BranchType`` and ``Node``.

This is the user-visible tree format (``clf.tree_``) and the pickle format.  It
keeps the reference's field names, defaults, ``repr`` and pickle behaviour
(reference: ``mpitree/tree/_base.py:16-19`` for the glyph enum, ``:22-57`` for
the dataclass fields, ``:59-61`` depth-from-parent, ``:63-75`` the ordering
comparator used by the text export, ``:77-101`` ``is_leaf``/``children``).

The training engines never build ``Node`` graphs while fitting: they produce a
structure-of-arrays :class:`mpitree_b200.models.flat_tree.FlatTree` on the
device and the graph is materialised once, on the host, at the end of ``fit``.
"""

from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import Any, List, Optional

__all__ = ["BranchType", "Node"]


class BranchType(enum.Enum):
    """Box-drawing connector printed in front of a node by ``export_text``."""

    ROOT = "┌──"  # ┌──
    INTERIOR_LIKE = "├──"  # ├──
    LEAF_LIKE = "└──"  # └──


@dataclass(kw_only=True)
class Node:
    """One node of a fitted decision tree.

    Parameters
    ----------
    value : int
        Feature index for an interior node, class label for a leaf.
    threshold : object, optional
        Split point of an interior node.  Always one of the training values of
        the split feature (rows with ``x[value] <= threshold`` go left).
    depth : int
        Distance from the root.  Derived from ``parent`` when one is given.
    count : array-like
        Per-class number of training rows that reached the node.
    parent, left, right : Node, optional
        Links.  A node without children is a leaf.
    """

    value: int
    threshold: Optional[Any] = None
    depth: int = field(default_factory=int)
    count: Any = field(default_factory=list)
    parent: Optional["Node"] = field(default=None, repr=False)
    left: Optional["Node"] = field(default=None, repr=False)
    right: Optional["Node"] = field(default=None, repr=False)
    _btype: BranchType = field(default=BranchType.ROOT, repr=False)

    def __post_init__(self) -> None:
        if self.parent is not None:
            self.depth = self.parent.depth + 1

    # ``sorted([left, right])`` evaluates ``right < left`` exactly once; the text
    # export relies on that call both for the print order and for tagging the
    # two siblings with their connector glyphs.
    def __lt__(self, other: "Node") -> bool:
        right_first = not self.is_leaf
        if right_first:
            self._btype, other._btype = BranchType.INTERIOR_LIKE, BranchType.LEAF_LIKE
        else:
            other._btype, self._btype = BranchType.INTERIOR_LIKE, BranchType.LEAF_LIKE
        return right_first

    @property
    def is_leaf(self) -> bool:
        """True when the node has neither a left nor a right child."""
        return self.left is None and self.right is None

    @property
    def children(self) -> List["Node"]:
        """``[left, right]`` for an interior node, ``[]`` for a leaf."""
        if self.is_leaf:
            return []
        return [self.left, self.right]

    # -- helpers that the reference does not have -------------------------
    def iter_preorder(self):
        """Yield the subtree's nodes root-first, left before right (iterative)."""
        stack = [self]
        while stack:
            node = stack.pop()
            yield node
            if not node.is_leaf:
                stack.append(node.right)
                stack.append(node.left)

    @property
    def n_nodes(self) -> int:
        return sum(1 for _ in self.iter_preorder())

    @property
    def max_depth(self) -> int:
        return max(n.depth for n in self.iter_preorder())
