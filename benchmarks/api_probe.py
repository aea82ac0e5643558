"""Where does the public-API fit spend its time?  (development probe, 1 GPU)"""
import sys
import time

import torch

sys.path.insert(0, ".")
from bench import synth_shard  # noqa: E402
from mpitree_b200.ops import cuda_engine  # noqa: E402
from mpitree_b200.ops.cpu_engine import SplitParams  # noqa: E402
from mpitree_b200.tree import ParallelDecisionTreeClassifier  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
dev = torch.device("cuda")
X, y = synth_shard(torch, dev, rows, 64, 10, 1234)
clf = ParallelDecisionTreeClassifier(max_depth=10, data_layout="sharded", device="cuda")


def t(fn, name, reps=3):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    torch.cuda.synchronize()
    print(f"{name:28s} {(time.perf_counter() - t0) / reps * 1e3:8.2f} ms")
    return out


t(lambda: clf.fit(X, y), "clf.fit (API)")
t(lambda: (int(y.min()), int(y.max())), "label min/max")
t(lambda: torch.bincount(y.view(-1).int(), minlength=10), "bincount")
data = cuda_engine.DeviceDataset.allocate(rows, 64, 10, dev)
t(lambda: cuda_engine.DeviceDataset.allocate(rows, 64, 10, dev), "DeviceDataset.allocate")
t(lambda: data.encode_u8(X), "encode_u8")
t(lambda: data.set_labels(y), "set_labels")
b = t(lambda: cuda_engine.TreeBuilder(data, 256, SplitParams(max_depth=10)), "TreeBuilder()")


def build():
    data.cur = 0
    data.encode_u8(X)
    data.set_labels(y)
    bb = cuda_engine.TreeBuilder(data, 256, SplitParams(max_depth=10))
    return bb.build()


flat = t(build, "encode+builder+build")
t(lambda: flat.to_nodes(), "to_nodes")
