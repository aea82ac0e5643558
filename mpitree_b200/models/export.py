"""Text serialisation of a tree (the human-readable format).

Grammar, reproduced from the reference's ``export_text``
(``mpitree/tree/decision_tree.py:250-307``) and its sibling comparator
(``mpitree/tree/_base.py:63-75``):

* one line per node, joined by ``"\\n"``, no trailing newline;
* ``line = prefix + glyph + " " + label`` and, below the root,
  ``+ " [<= t]"`` / ``" [> t]"`` where ``t`` is the *parent's* threshold printed
  with ``precision`` decimals;
* label: leaf -> ``class_names[value]`` or ``"class: {value}"``; interior ->
  ``feature_names[value]`` or ``"feature_{value}"``;
* sibling order: when the right (``>``) child is interior it is printed first
  with ``├──`` and the left child second with ``└──``; otherwise left first
  (``├──``) and right second (``└──``);
* the children of a ``└──`` node are indented by three spaces, all others by
  ``"│  "``.

The renderer is iterative (explicit stack), so deep trees do not hit Python's
recursion limit, and it works on either a :class:`Node` graph or a
:class:`FlatTree`.
"""

from __future__ import annotations

from typing import Optional, Sequence

from ..tree._base import BranchType, Node

__all__ = ["render_text"]


def _format_threshold(threshold, precision: int) -> str:
    try:
        return f"{threshold:.{precision}f}"
    except (TypeError, ValueError):
        # Ordinal string/categorical thresholds: the reference raises here
        # (format code 'f' on str); printing the raw value is the fix.
        return str(threshold)


def render_text(
    root: Node,
    *,
    feature_names: Optional[Sequence] = None,
    class_names: Optional[Sequence] = None,
    precision: int = 2,
) -> str:
    """Render ``root`` with the mpitree text grammar and return the string."""
    lines = []
    root._btype = BranchType.ROOT if root.parent is None else root._btype
    stack = [(root, "")]
    while stack:
        node, prefix = stack.pop()
        if node.is_leaf:
            label = class_names[node.value] if class_names is not None else f"class: {node.value}"
        else:
            label = (
                feature_names[node.value] if feature_names is not None else f"feature_{node.value}"
            )
        text = f"{node._btype.value} {label}"
        if node.depth:
            sign = "<=" if node is node.parent.left else ">"
            text += f" [{sign} {_format_threshold(node.parent.threshold, precision)}]"
        lines.append(prefix + text)

        if node.is_leaf:
            continue
        left, right = node.left, node.right
        if right.is_leaf:
            first, second = left, right
        else:
            first, second = right, left
        first._btype, second._btype = BranchType.INTERIOR_LIKE, BranchType.LEAF_LIKE
        child_prefix = prefix + ("   " if node._btype is BranchType.LEAF_LIKE else "│  ")
        stack.append((second, child_prefix))
        stack.append((first, child_prefix))
    return "\n".join(lines)
