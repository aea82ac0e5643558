"""GPU tests that compare alternative kernel variants with each other.

Quad float encode vs the scalar kernel, interleaved vs planar shared-memory tile, in-switch (NVLS
``multimem``) vs peer-push histogram reduction, object vs device-to-device record exchange of the
hand-off phase.  All of them were validated on B200s in round 2 (``profiles/r2/staged_*gpu.txt``); the
multi-rank ones need at least two GPUs and skip otherwise.
"""
import os

import numpy as np
import pytest

from mpitree_b200.ops.binning import BinMapper

pytestmark = [pytest.mark.gpu]


@pytest.mark.parametrize("F,max_bins", [(4, 256), (32, 256), (36, 64), (64, 17)])
def test_encode_f32_quads_equals_scalar_kernel(monkeypatch, F, max_bins):
    import torch

    from mpitree_b200.ops.cuda_engine import DeviceDataset

    rng = np.random.default_rng(F)
    n = 70_001
    X = rng.normal(size=(n, F)).astype(np.float32)
    X[:, 0] = rng.integers(0, 7, size=n)  # lossless column
    X[::97, 1] = np.inf
    X[::89, 2] = -np.inf
    X[::83, 3] = np.nan
    mapper = BinMapper.fit(np.nan_to_num(X[: n // 2], nan=0.0), max_bins=max_bins)  # values beyond the edges exist
    Xd = torch.from_numpy(X).cuda()
    out = []
    for mode in ("scalar", "quads", "slots"):
        monkeypatch.setenv("MTB_ENCODE", mode)
        data = DeviceDataset.allocate(n, F, 2, torch.device("cuda"))
        data.codes[0].zero_()
        data.encode_float(Xd, mapper, 0)
        out.append(data.codes_rowmajor().cpu().numpy().copy())
        data.release()
    assert np.array_equal(out[0], out[1])
    assert np.array_equal(out[0], out[2])
    finite = np.isfinite(X).all(axis=1)
    assert np.array_equal(out[1][finite], mapper.transform(X[finite]))


_NVLS_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["REPO"])
import torch, torch.distributed as dist
from mpitree_b200.parallel.comm import TorchComm, init_process_group_from_env
from mpitree_b200.parallel.allreduce import PeerAllReduce
from mpitree_b200.parallel.nvls import nvls_available
assert init_process_group_from_env("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
if not nvls_available(torch.cuda.current_device()):
    if rank == 0:
        print("NVLS_UNSUPPORTED")
    dist.destroy_process_group()
    sys.exit(0)
ar = PeerAllReduce(TorchComm(), 1 << 20, nvls=True)
for count in (1, 2, 5, 1024, 100_003, 1 << 20):
    x = torch.randint(0, 1 << 27, (count,), dtype=torch.int32, device="cuda")
    want = x.clone(); dist.all_reduce(want)
    for algo in (0, 1, 2, 2):  # the peer-mapped algorithms must also work on a VMM region
        got = ar.all_reduce(x, algo)
        assert torch.equal(got, want), (count, algo, rank)
ar.close()
dist.barrier()
if rank == 0:
    print("NVLS_OK", world)
dist.destroy_process_group()
"""


@pytest.mark.multigpu
def test_nvls_allreduce_equals_nccl(tmp_path):
    import subprocess
    import sys

    import torch

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    script = tmp_path / "nvls_worker.py"
    script.write_text(_NVLS_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29653", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert f"NVLS_OK {world}" in res.stdout or "NVLS_UNSUPPORTED" in res.stdout


def _same_tree(a, b):
    assert a.n_nodes == b.n_nodes
    for name in ("feature", "bin", "counts", "left", "right", "depth"):
        assert np.array_equal(getattr(a, name), getattr(b, name)), name
    assert np.array_equal(np.nan_to_num(a.gain, nan=-1.0), np.nan_to_num(b.gain, nan=-1.0))


@pytest.mark.parametrize("shape", [(5000, 7, 3, 6), (300_000, 40, 10, 9), (1_200_000, 64, 10, 8), (60_000, 33, 2, 12)])
@pytest.mark.parametrize("mode", [dict(fused=True), dict(fused=True, fuse_partition=False), dict(fused=False)])
def test_interleaved_tile_equals_planar(shape, mode):
    """Same tree, bit for bit, from the [B*C][16] shared-memory tile as from the [f][B][C] one."""
    from mpitree_b200.ops import cuda_engine
    from mpitree_b200.ops.cpu_engine import SplitParams

    n, F, k, depth = shape
    rng = np.random.default_rng(n + F)
    X = rng.integers(0, 256, size=(n, F), dtype=np.uint8)
    X[:, 1] = rng.integers(0, 3, size=n)  # a few heavily repeated codes: same-address increments
    w = rng.normal(size=F)
    s = X.astype(np.float64) @ w + rng.normal(scale=200, size=n)
    y = np.digitize(s, np.quantile(s, np.linspace(0, 1, k + 1)[1:-1])).astype(np.int64)
    params = SplitParams(max_depth=depth)
    want, _ = cuda_engine.fit_from_host(X, y, k, params, builder_kwargs=dict(mode, tile_layout="planar"))
    got, _ = cuda_engine.fit_from_host(X, y, k, params, builder_kwargs=dict(mode, tile_layout="interleaved"))
    assert cuda_engine.fit_from_host.last_stats.tile_layout == "interleaved"
    _same_tree(got, want)


_NVLS_TREE_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["REPO"])
import numpy as np, torch, torch.distributed as dist
from mpitree_b200.ops import cuda_engine
from mpitree_b200.parallel.comm import init_process_group_from_env
from mpitree_b200.parallel.nvls import nvls_available
from mpitree_b200.tree import ParallelDecisionTreeClassifier
assert init_process_group_from_env("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
if not nvls_available(torch.cuda.current_device()):
    if rank == 0:
        print("NVLS_UNSUPPORTED")
    dist.destroy_process_group()
    sys.exit(0)
rng = np.random.default_rng(0)
n = 400_000
X = rng.integers(0, 256, size=(n, 48), dtype=np.uint8)
s = X[:, :6].astype(np.float64) @ np.array([1, -1, .5, .5, -.25, 2.0]) + rng.normal(scale=60, size=n)
y = np.digitize(s, np.quantile(s, np.linspace(0, 1, 6)[1:-1]))
lo, hi = n * rank // world, n * (rank + 1) // world
trees = {}
for reduce in ("p2p", "nvls"):
    for tile in ("planar", "interleaved"):
        clf = ParallelDecisionTreeClassifier(max_depth=9, device="cuda", data_layout="sharded")
        clf.builder_kwargs = {"reduce": reduce, "tile_layout": tile}
        for _ in range(2):  # the second fit reuses the cached regions with later epochs
            clf.fit(X[lo:hi], y[lo:hi])
        st = cuda_engine.fit_from_host.last_stats
        assert st.reduce == reduce and st.tile_layout == tile, (st.reduce, st.tile_layout)
        trees[(reduce, tile)] = clf.flat_tree_
want = trees[("p2p", "planar")]
for key, t in trees.items():
    assert t.structurally_equal(want), key
    assert np.array_equal(t.counts, want.counts), key
dist.barrier()
if rank == 0:
    print("NVLS_TREE_OK", world)
dist.destroy_process_group()
"""


@pytest.mark.multigpu
def test_nvls_tree_equals_p2p(tmp_path):
    import subprocess
    import sys

    import torch

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    script = tmp_path / "nvls_tree_worker.py"
    script.write_text(_NVLS_TREE_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29657", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert f"NVLS_TREE_OK {world}" in res.stdout or "NVLS_UNSUPPORTED" in res.stdout


_FLOAT_SHARD_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["REPO"])
import numpy as np, torch, torch.distributed as dist
from mpitree_b200.parallel.comm import TorchComm, init_process_group_from_env
from mpitree_b200.tree import ParallelDecisionTreeClassifier
assert init_process_group_from_env("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
rng = np.random.default_rng(0)
n = 300_000
X = rng.normal(size=(n, 12)).astype(np.float32)            # far more distinct values than bins
y = (X[:, 0] + X[:, 3] * X[:, 5] + rng.normal(scale=0.5, size=n) > 0).astype(np.int64)
lo, hi = n * rank // world, n * (rank + 1) // world
clf = ParallelDecisionTreeClassifier(max_depth=8, device="cuda", data_layout="sharded").fit(X[lo:hi], y[lo:hi])
texts = TorchComm().allgather_object(clf.export_text())
assert all(t == texts[0] for t in texts), "ranks derived different trees (bin edges must be agreed on)"
acc = float((clf.predict(X[:20000]) == y[:20000]).mean())
assert acc > 0.8, acc
dist.barrier()
if rank == 0:
    print("FLOAT_SHARD_OK", world, round(acc, 3))
dist.destroy_process_group()
"""


@pytest.mark.multigpu
def test_sharded_float_numpy_ranks_agree(tmp_path):
    """numpy float input, row-sharded: every rank must bin with the same edges (fixed on the host path
    after a CPU sweep; this is the same code feeding the CUDA engine)."""
    import subprocess
    import sys

    import torch

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    script = tmp_path / "float_shard_worker.py"
    script.write_text(_FLOAT_SHARD_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29659", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert f"FLOAT_SHARD_OK {world}" in res.stdout


_RECORD_EXCHANGE_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["REPO"])
import numpy as np, torch, torch.distributed as dist
from mpitree_b200.ops import subtree_phase
from mpitree_b200.parallel.comm import init_process_group_from_env
from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier
assert init_process_group_from_env("nccl")
rank, world = dist.get_rank(), dist.get_world_size()
rng = np.random.default_rng(0)
n = 60_000
X = rng.integers(0, 256, size=(n, 24), dtype=np.uint8)
s = X[:, :6].astype(np.float64) @ np.array([1, -1, .5, .5, -.25, 2.0]) + rng.normal(scale=60, size=n)
y = np.digitize(s, np.quantile(s, np.linspace(0, 1, 5)[1:-1]))
want = DecisionTreeClassifier(max_depth=14, device="cuda").fit(X, y)
lo, hi = n * rank // world, n * (rank + 1) // world
for threshold in (1 << 40, 0):  # records as objects, then device to device through a peer-mapped region
    subtree_phase.DEVICE_EXCHANGE_MIN_BYTES = threshold
    clf = ParallelDecisionTreeClassifier(max_depth=14, device="cuda", data_layout="sharded", handoff_rows=900)
    clf.fit(X[lo:hi], y[lo:hi])
    assert clf.flat_tree_.structurally_equal(want.flat_tree_), (threshold, rank)
dist.barrier()
if rank == 0:
    print("RECORD_EXCHANGE_OK", world)
dist.destroy_process_group()
"""


@pytest.mark.multigpu
def test_handoff_record_exchange_paths(tmp_path):
    """Phase 2 on real ranks with the finished-subtree records exchanged both ways (objects / device to device)."""
    import subprocess
    import sys

    import torch

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    script = tmp_path / "record_exchange_worker.py"
    script.write_text(_RECORD_EXCHANGE_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29661", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert f"RECORD_EXCHANGE_OK {world}" in res.stdout


@pytest.mark.parametrize("F", [16, 32, 48, 64, 128])
def test_root_pass_variants_equal_separate_passes(monkeypatch, F):
    """First level of uint8 rows that already are bin codes, three ways: ``direct`` (the default: planes + labels
    + root histogram in one pass of 128-bit loads from the row-major matrix), ``cluster`` (opt-in: thread-block
    clusters of F/16 CTAs share each row tile through a multicast TMA copy; F = 32 / 64 / 128 only) and ``off``
    (retile + label copy, histogram by the level kernel).  Same tree."""
    import torch

    from mpitree_b200.tree import DecisionTreeClassifier

    monkeypatch.delenv("MTB_ROOT_CLUSTER", raising=False)
    rng = np.random.default_rng(F)
    for n in (999, 70_001, 300_000):
        X = torch.from_numpy(rng.integers(0, 256, size=(n, F), dtype=np.uint8)).cuda()
        y = ((X[:, 1] > 90).to(torch.uint8) + (X[:, F - 1] > 200).to(torch.uint8) + (X[:, F // 2 + 1] > 30).to(torch.uint8))
        trees = {}
        for mode in ("off", "direct", "cluster"):
            monkeypatch.setenv("MTB_ROOT", mode)
            trees[mode] = DecisionTreeClassifier(max_depth=6, device="cuda").fit(X, y).flat_tree_
        _same_tree(trees["direct"], trees["off"])
        _same_tree(trees["cluster"], trees["off"])
