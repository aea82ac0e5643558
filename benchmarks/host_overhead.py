"""Host-side cost of one public-API fit: a small device-resident data set (GPU time is short), wall clock per
fit and a cProfile listing of where the host time goes.  Run alone or under torchrun."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import synth_rows  # noqa: E402
from mpitree_b200.parallel.comm import init_process_group_from_env  # noqa: E402
from mpitree_b200.tree import ParallelDecisionTreeClassifier  # noqa: E402

world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    init_process_group_from_env("nccl")
rows = int(os.environ.get("ROWS", "400000"))
lo, hi = rows * rank // world, rows * (rank + 1) // world
X, y = synth_rows(torch, dev, lo, hi, 64, 10)
clf = ParallelDecisionTreeClassifier(max_depth=10, data_layout="sharded", device="cuda")
for _ in range(5):
    clf.fit(X, y)
torch.cuda.synchronize()
t0 = time.perf_counter()
n = 30
for _ in range(n):
    clf.fit(X, y)
torch.cuda.synchronize()
ms = (time.perf_counter() - t0) / n * 1e3
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(n):
    clf.fit(X, y)
e.record()
torch.cuda.synchronize()
if rank == 0:
    print(f"world={world} rows={rows}: wall {ms:.3f} ms per fit, device-event {s.elapsed_time(e) / n:.3f} ms per fit")
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    clf.fit(X, y)
torch.cuda.synchronize()
pr.disable()
if rank == 0:
    st = pstats.Stats(pr)
    st.sort_stats("cumulative").print_stats(45)
if world > 1:
    import torch.distributed as dist

    dist.barrier()
    dist.destroy_process_group()
