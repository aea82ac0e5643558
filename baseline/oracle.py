"""Import the unmodified reference (``mpitree``) with the mpi4py shim on the path."""

from __future__ import annotations

import importlib
import pathlib
import sys

HERE = pathlib.Path(__file__).resolve().parent


def reference_root():
    for cand in (HERE / "_ref", pathlib.Path("/root/reference")):
        if (cand / "mpitree" / "tree" / "decision_tree.py").exists():
            return cand
    return None


def load_reference():
    """Return the reference's ``mpitree.tree`` module or ``None`` when unavailable."""
    root = reference_root()
    if root is None:
        return None
    try:
        import mpi4py  # noqa: F401  (a real one wins if present)
    except ImportError:
        shim = str(HERE / "shim")
        if shim not in sys.path:
            sys.path.insert(0, shim)
    if str(root) not in sys.path:
        sys.path.insert(0, str(root))
    return importlib.import_module("mpitree.tree")
