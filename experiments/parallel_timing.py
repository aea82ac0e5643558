"""Parallel rows of the timing experiment with REAL processes (gloo), one torchrun launch per k.

    python -m torch.distributed.run --nproc-per-node K --master-addr 127.0.0.1 experiments/parallel_timing.py

Same workload as ``experiments.ipynb:198-209`` (n distinct classes, one feature, unlimited depth);
every rank passes the full data like the reference does, best of 3, max over ranks, milliseconds.
"""
import pathlib
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
from mpitree_b200.parallel.comm import init_process_group_from_env  # noqa: E402
from mpitree_b200.tree import ParallelDecisionTreeClassifier  # noqa: E402

assert init_process_group_from_env("gloo")
rank, k = dist.get_rank(), dist.get_world_size()
row = []
for n in range(1, 250, 10):
    X, y = np.arange(n).reshape(-1, 1), np.arange(n)
    best = float("inf")
    for _ in range(3):
        dist.barrier()
        t0 = time.perf_counter()
        ParallelDecisionTreeClassifier(device="cpu").fit(X, y)
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t) * 1000)
    row.append(round(best, 2))
if rank == 0:
    print(f"k={k}: " + ", ".join(str(v) for v in row))
    out = pathlib.Path(__file__).resolve().parent / f"time_data_k{k}_processes.csv"
    out.write_text(", ".join(str(v) for v in row) + "\n")

# A larger replicated problem (3000 x 8 continuous, unlimited depth): serial fit on every rank vs the
# parallel class, same processes, max over ranks, best of 3.
import json  # noqa: E402

from mpitree_b200.tree import DecisionTreeClassifier  # noqa: E402

rng = np.random.default_rng(0)
XL = rng.normal(size=(3000, 8)).round(2)
yL = (XL[:, 0] + XL[:, 1] * XL[:, 2] + rng.normal(scale=0.7, size=3000) > 0).astype(int) + (XL[:, 3] > 1)


def timed(make):
    best = float("inf")
    for _ in range(3):
        dist.barrier()
        t0 = time.perf_counter()
        make().fit(XL, yL)
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t) * 1000)
    return round(best, 1)


large = {"k": k, "rows": 3000, "features": 8, "serial_ms": timed(lambda: DecisionTreeClassifier(device="cpu")),
         "parallel_ms": timed(lambda: ParallelDecisionTreeClassifier(device="cpu"))}
if rank == 0:
    print("large:", json.dumps(large))
    (pathlib.Path(__file__).resolve().parent / f"timing_large_k{k}.json").write_text(json.dumps(large) + "\n")
dist.barrier()
dist.destroy_process_group()
