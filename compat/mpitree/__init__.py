"""Drop-in import path for code written against the reference: put ``compat/`` on ``PYTHONPATH``
and ``from mpitree.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier`` resolves
to this framework (reference layout: ``mpitree/__init__.py`` is empty, ``mpitree/tree/__init__.py:1-3``
exports the two estimators)."""
