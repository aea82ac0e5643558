"""``mpitree.tree._base`` alias (reference ``mpitree/tree/_base.py``)."""

from mpitree_b200.tree._base import BranchType, Node  # noqa: F401
