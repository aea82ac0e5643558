"""Tree model formats: flat device arrays, object graph conversion, text export."""

from .export import render_text
from .flat_tree import FlatTree

__all__ = ["FlatTree", "render_text"]
