"""mpitree_b200 — a Blackwell-native parallel decision-tree framework.

Same estimator surface as ``mpitree`` (``from mpitree_b200.tree import
DecisionTreeClassifier, ParallelDecisionTreeClassifier``), rebuilt around
row-sharded count histograms, hand-written sm_100a kernels and NVLink peer
memory.  See DESIGN.md for the architecture.
"""

__version__ = "0.1.0"

from .tree import BranchType, DecisionTreeClassifier, Node, ParallelDecisionTreeClassifier  # noqa: E402,F401

__all__ = [
    "DecisionTreeClassifier",
    "ParallelDecisionTreeClassifier",
    "Node",
    "BranchType",
    "__version__",
]
