"""Public estimator API (mirrors ``mpitree.tree``, reference ``mpitree/tree/__init__.py:1-3``)."""

from ._base import BranchType, Node
from .decision_tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

__all__ = ["DecisionTreeClassifier", "ParallelDecisionTreeClassifier", "Node", "BranchType"]
