"""Exercise every kernel once on moderate shapes (used under ncu to capture one profile per kernel)."""
import sys

import torch

sys.path.insert(0, ".")
from bench import synth_shard  # noqa: E402
from mpitree_b200.ops import cuda_engine  # noqa: E402
from mpitree_b200.ops.cpu_engine import SplitParams  # noqa: E402
from mpitree_b200.tree import DecisionTreeClassifier  # noqa: E402

dev = torch.device("cuda")
X, y = synth_shard(torch, dev, 8_000_000, 64, 10, 1)
# separate phases (plan / hist / gain / finalize / partition / commit), depth 4
data = cuda_engine.DeviceDataset.allocate(X.shape[0], 64, 10, dev)
data.encode_u8(X)
data.set_labels(y)
b = cuda_engine.TreeBuilder(data, 256, SplitParams(max_depth=2), fused=False, async_levels=False)
flat = b.build()
print("unfused nodes", flat.n_nodes)
# subtree kernel (hand-off) on a fully grown tree
clf = DecisionTreeClassifier(device="cuda")
clf.builder_kwargs = {"small_rows": 1024}
clf.fit(X[:200_000], y[:200_000])
print("grown nodes", clf.flat_tree_.n_nodes, cuda_engine.fit_from_torch.last_stats.handed_off)
# predict (staged u8) and float encode + float predict
deep = DecisionTreeClassifier(max_depth=12, device="cuda").fit(X[:2_000_000], y[:2_000_000])
print("pred", int(deep.predict(X).sum()))
Xf = torch.randn(2_000_000, 32, device=dev)
yf = (Xf[:, 0] > 0).to(torch.int64)
cf = DecisionTreeClassifier(max_depth=6, device="cuda").fit(Xf, yf)
print("float", cf.flat_tree_.n_nodes, int(cf.predict(Xf).sum()))
