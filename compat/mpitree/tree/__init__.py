"""``mpitree.tree`` alias of :mod:`mpitree_b200.tree` (reference ``mpitree/tree/__init__.py:1-3``)."""

from mpitree_b200.tree import BranchType, DecisionTreeClassifier, Node, ParallelDecisionTreeClassifier

__all__ = ["DecisionTreeClassifier", "ParallelDecisionTreeClassifier", "Node", "BranchType"]
