"""Utilities: synthetic data, timing, logging, checkpoints, clocks."""
