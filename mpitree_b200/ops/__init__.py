"""Operators: binning (K1), split search (K2), partition (K3), predict (K4), all-reduce (K5)."""
