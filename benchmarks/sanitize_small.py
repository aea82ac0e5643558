"""Small single-GPU workload for compute-sanitizer (one tool per run):

    compute-sanitizer --tool memcheck  python benchmarks/sanitize_small.py
    compute-sanitizer --tool racecheck python benchmarks/sanitize_small.py
    compute-sanitizer --tool synccheck python benchmarks/sanitize_small.py

Touches every kernel family on small shapes: float + byte encode, the fused level kernels (root + fused
partition levels), the unfused phase kernels, the subtree kernel (hand-off), the three predict kernels.
"""
import numpy as np
import torch

from mpitree_b200.ops import cuda_engine
from mpitree_b200.ops.cpu_engine import SplitParams
from mpitree_b200.tree import DecisionTreeClassifier

rng = np.random.default_rng(0)
n, F = 6000, 20
X8 = rng.integers(0, 256, size=(n, F), dtype=np.uint8)
y = ((X8[:, 0] > 100).astype(np.int64) + (X8[:, 3] > 50) + (X8[:, 7] > 200))
Xf = rng.normal(size=(n, F)).astype(np.float32)
yf = (Xf[:, 0] + Xf[:, 1] * Xf[:, 2] > 0).astype(np.int64)

a = DecisionTreeClassifier(max_depth=5, device="cuda").fit(X8, y)               # byte encode + fused kernels
b = DecisionTreeClassifier(max_depth=5, device="cuda").fit(Xf, yf)              # quad float encode
c = DecisionTreeClassifier(max_depth=4, device="cuda").fit(Xf.astype(np.float64), yf)  # scalar float encode
flat, _ = cuda_engine.fit_from_host(X8, y, 4, SplitParams(max_depth=4), builder_kwargs=dict(fused=False))  # phase kernels
flat2, _ = cuda_engine.fit_from_host(X8, y, 4, SplitParams(max_depth=4), builder_kwargs=dict(fuse_partition=False))
deep = DecisionTreeClassifier(max_depth=12, device="cuda")
deep.builder_kwargs = {"small_rows": 600}
deep.fit(X8, y)                                                                  # hand-off: subtree kernel
for clf, Z in ((a, X8), (b, Xf), (c, Xf.astype(np.float64))):
    p = clf.predict(Z)
    assert p.shape[0] == n
pt = a.predict(torch.from_numpy(X8).cuda())
torch.cuda.synchronize()
print("sanitize workload ok", a.flat_tree_.n_nodes, b.flat_tree_.n_nodes, flat.n_nodes, flat2.n_nodes, deep.flat_tree_.n_nodes)
