"""Multi-rank GPU data path: virtual ranks on one GPU (lock-step) and real ranks over NVLink."""
import os
import subprocess
import sys

import numpy as np
import pytest

from mpitree_b200.ops.cpu_engine import SplitParams

pytestmark = pytest.mark.gpu


def _data(n, F, k, seed):
    rng = np.random.default_rng(seed)
    X = rng.integers(0, 64, size=(n, F), dtype=np.uint8)
    w = rng.normal(size=F)
    s = X.astype(np.float64) @ w + rng.normal(scale=20, size=n)
    y = np.digitize(s, np.quantile(s, np.linspace(0, 1, k + 1)[1:-1])).astype(np.uint8)
    return X, y


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("shape", [(6000, 7, 3, 6), (20000, 40, 10, 5)])
def test_virtual_ranks_equal_single_rank(world, shape):
    import torch

    from mpitree_b200.ops import cuda_engine

    n, F, k, depth = shape
    X, y = _data(n, F, k, seed=world * 100 + F)
    params = SplitParams(max_depth=depth)
    want, _ = cuda_engine.fit_from_host(X, y.astype(np.int64), k, params)
    cuts = [n * r // world for r in range(world + 1)]
    # uneven shards on purpose: the last rank also gets an empty-ish tail
    shards = [(torch.from_numpy(X[a:b]).cuda(), torch.from_numpy(y[a:b]).cuda()) for a, b in zip(cuts, cuts[1:])]
    trees = cuda_engine.fit_virtual_ranks(shards, k, 256, params)
    for t in trees:
        assert t.n_nodes == want.n_nodes
        assert np.array_equal(t.feature, want.feature) and np.array_equal(t.bin, want.bin)
        assert np.array_equal(t.counts, want.counts) and np.array_equal(t.left, want.left)


_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["REPO"])
import numpy as np, torch, torch.distributed as dist
from mpitree_b200.parallel.comm import init_process_group_from_env
from mpitree_b200.tree import ParallelDecisionTreeClassifier, DecisionTreeClassifier
assert init_process_group_from_env("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
rng = np.random.default_rng(0)
n = 200_000
X = rng.integers(0, 256, size=(n, 48), dtype=np.uint8)
s = X[:, :6].astype(np.float64) @ np.array([1, -1, .5, .5, -.25, 2.0]) + rng.normal(scale=60, size=n)
y = np.digitize(s, np.quantile(s, np.linspace(0, 1, 6)[1:-1]))
want = DecisionTreeClassifier(max_depth=7, device="cuda").fit(X, y)
for reduce in ("p2p", "nccl"):
    for layout in ("replicated", "sharded"):
        clf = ParallelDecisionTreeClassifier(max_depth=7, device="cuda", data_layout=layout)
        clf.builder_kwargs = {"reduce": reduce}
        lo, hi = n * rank // world, n * (rank + 1) // world
        Xs, ys = (X, y) if layout == "replicated" else (X[lo:hi], y[lo:hi])
        for _ in range(2):  # second fit reuses the mapped region with later epochs
            clf.fit(Xs, ys)
            assert clf.export_text() == want.export_text(), (reduce, layout, rank)
            assert np.array_equal(clf.flat_tree_.counts, want.flat_tree_.counts)
# phase 2: hand-off of small subtrees to owner ranks (row gather over NVLink + subtree kernel)
deep = DecisionTreeClassifier(max_depth=14, device="cuda").fit(X[:30000], y[:30000])
lo2, hi2 = 30000 * rank // world, 30000 * (rank + 1) // world
clf = ParallelDecisionTreeClassifier(max_depth=14, device="cuda", data_layout="sharded", handoff_rows=700)
clf.fit(X[lo2:hi2], y[lo2:hi2])
assert clf.flat_tree_.structurally_equal(deep.flat_tree_), ("handoff", rank)
# torch tensor API, sharded
Xt, yt = torch.from_numpy(X[lo:hi]).cuda(), torch.from_numpy(y[lo:hi]).cuda()
clf = ParallelDecisionTreeClassifier(max_depth=7, data_layout="sharded").fit(Xt, yt)
assert clf.export_text() == want.export_text()
# K5: peer all-reduce kernels vs torch
from mpitree_b200.parallel.allreduce import PeerAllReduce
from mpitree_b200.parallel.comm import TorchComm
ar = PeerAllReduce(TorchComm(), 1 << 20)
for count in (1, 5, 1024, 100_003, 1 << 20):
    x = torch.randint(0, 1 << 20, (count,), dtype=torch.int32, device="cuda")
    want = x.clone(); dist.all_reduce(want)
    for algo in (0, 1):
        got = ar.all_reduce(x, algo)
        assert torch.equal(got, want), (count, algo)
ar.close()
dist.barrier()
if rank == 0:
    print("MULTIGPU_OK", world)
dist.destroy_process_group()
"""


@pytest.mark.multigpu
def test_real_ranks_p2p_and_nccl(tmp_path):
    import torch

    world = min(torch.cuda.device_count(), 8)
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29633", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert f"MULTIGPU_OK {world}" in res.stdout


_FAILURE_WORKER = r"""
import os, sys, time
sys.path.insert(0, os.environ["REPO"])
import numpy as np, torch, torch.distributed as dist
from mpitree_b200.ops import cuda_engine
from mpitree_b200.ops.cpu_engine import SplitParams
from mpitree_b200.parallel.comm import TorchComm, init_process_group_from_env
assert init_process_group_from_env("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", torch.cuda.current_device())
g = torch.Generator(device=dev).manual_seed(rank)
n = 200_000
X = torch.randint(0, 256, (n, 32), dtype=torch.uint8, device=dev, generator=g)
y = ((X[:, 0] > 100) ^ (X[:, 5] > 30)).to(torch.uint8)
data = cuda_engine.DeviceDataset.allocate(n, 32, 2, dev)
data.encode_u8(X); data.set_labels(y)
b = cuda_engine.TreeBuilder(data, 256, SplitParams(max_depth=6), comm=TorchComm(), n_global=n * world)
first = b.build()                       # every rank takes part: works
assert first.n_nodes > 3
dist.barrier()
torch.cuda.synchronize()
if rank == world - 1:
    # this rank "fails": it never launches the next fit's kernels
    time.sleep(float(os.environ["MTB_BARRIER_TIMEOUT_S"]) + 4.0)
    print("SILENT_RANK_DONE", flush=True)
    os._exit(0)
data.cur = 0
data.encode_u8(X); data.set_labels(y)
t0 = time.time()
try:
    b.build()
    print("NO_ERROR", rank, flush=True)
except RuntimeError as exc:
    took = time.time() - t0
    assert "did not reach an in-kernel barrier" in str(exc), exc
    print(f"RAISED rank={rank} after {took:.2f}s", flush=True)
    assert took < float(os.environ["MTB_BARRIER_TIMEOUT_S"]) + 3.0, took
os._exit(0)
"""


@pytest.mark.multigpu
def test_a_missing_rank_makes_the_others_raise_instead_of_hanging(tmp_path):
    """The reference deadlocks in ``allgather`` when one rank fails (decision_tree.py:456).  Here the level
    kernels of the surviving ranks give up at their peer barrier after ``MTB_BARRIER_TIMEOUT_S`` and the
    host raises on every one of them."""
    import torch

    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    script = tmp_path / "failure_worker.py"
    script.write_text(_FAILURE_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))), MTB_BARRIER_TIMEOUT_S="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29637", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=120)
    out = res.stdout + res.stderr
    assert out.count("RAISED rank=") == world - 1, out[-3000:]
    assert "NO_ERROR" not in out
