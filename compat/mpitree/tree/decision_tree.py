"""``mpitree.tree.decision_tree`` alias (reference ``mpitree/tree/decision_tree.py``)."""

from mpitree_b200.tree.decision_tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier  # noqa: F401
