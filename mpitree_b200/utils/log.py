"""Rank-aware logging.  The reference prints from user code behind ``if not clf.WORLD_RANK``
(``experiments.ipynb:67-69``); here a logger does that guard for the library."""

from __future__ import annotations

import logging
import os

__all__ = ["get_logger"]


def get_logger(name: str = "mpitree_b200") -> logging.Logger:
    """Logger that is silent on ranks > 0 unless ``MTB_LOG_ALL_RANKS=1`` (level via ``MTB_LOG``)."""
    log = logging.getLogger(name)
    if not log.handlers:
        h = logging.StreamHandler()
        rank = os.environ.get("RANK", "0")
        h.setFormatter(logging.Formatter(f"[%(asctime)s mtb r{rank}] %(levelname)s %(message)s", "%H:%M:%S"))
        log.addHandler(h)
        log.setLevel(os.environ.get("MTB_LOG", "WARNING").upper())
        if rank != "0" and os.environ.get("MTB_LOG_ALL_RANKS") != "1":
            log.setLevel(logging.ERROR)
        log.propagate = False
    return log
