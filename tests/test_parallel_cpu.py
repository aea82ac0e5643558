"""Multi-rank host logic on CPU: thread-ranks and real gloo processes (world_size 2..5)."""
import os
import sys

import numpy as np
import pytest

from mpitree_b200.parallel.comm import ThreadComm, run_threaded
from mpitree_b200.parallel.handoff import Graft, assign_owners, halving_groups, stitch_subtrees
from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

from .goldens import IRIS2_DEPTH3


def _fit_rank(comm, X, y, kw):
    clf = ParallelDecisionTreeClassifier(comm=comm, device="cpu", **kw).fit(X, y)
    return clf.export_text(), clf.predict(X).tolist()


@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
def test_parallel_equals_serial_replicated(iris, world):
    X, y = iris.data, iris.target
    want = DecisionTreeClassifier(max_depth=3, device="cpu").fit(X, y)
    out = run_threaded(world, _fit_rank, X, y, dict(max_depth=3, handoff_rows=0))
    for text, pred in out:
        assert text == want.export_text()
        assert pred == want.predict(X).tolist()


def test_parallel_golden_four_ranks(iris):
    # the notebook's 4-rank run (experiments.ipynb:58-70)
    X, y = iris.data[:, :2], iris.target

    def run(comm):
        clf = ParallelDecisionTreeClassifier(max_depth=3, comm=comm, device="cpu").fit(X, y)
        return clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names)

    assert run_threaded(4, run) == [IRIS2_DEPTH3] * 4


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("handoff", [10, 60, 1000])
def test_handoff_gives_identical_tree(iris, world, handoff):
    X, y = iris.data, iris.target
    want = DecisionTreeClassifier(device="cpu").fit(X, y).export_text()
    out = run_threaded(world, _fit_rank, X, y, dict(handoff_rows=handoff))
    assert [t for t, _ in out] == [want] * world


def test_sharded_layout_with_disjoint_values():
    rng = np.random.default_rng(3)
    X = rng.integers(0, 50, size=(400, 5)).astype(float)
    y = (X[:, 0] + X[:, 3] > 50).astype(int)
    want = DecisionTreeClassifier(max_depth=5, device="cpu").fit(X, y).export_text()

    def run(comm):
        lo, hi = 400 * comm.rank // comm.size, 400 * (comm.rank + 1) // comm.size
        clf = ParallelDecisionTreeClassifier(max_depth=5, comm=comm, device="cpu", data_layout="sharded")
        return clf.fit(X[lo:hi], y[lo:hi]).export_text()

    assert run_threaded(3, run) == [want] * 3


def test_rank_failure_propagates_instead_of_deadlocking():
    def run(comm):
        if comm.rank == 1:
            raise RuntimeError("boom")
        comm.barrier()

    with pytest.raises(RuntimeError, match="boom"):
        run_threaded(3, run)


def test_comm_split_parity():
    def run(comm):
        sub = comm.split(comm.rank % 2, comm.rank // 2)
        return sub.rank, sub.size, sub.allgather_object(comm.rank)

    out = run_threaded(5, run)
    assert out[0] == (0, 3, [0, 2, 4]) and out[3] == (1, 2, [1, 3]) and out[4] == (2, 3, [0, 2, 4])
    assert halving_groups([0, 1, 2, 3, 4]) == ([0, 2, 4], [1, 3])


def test_assign_owners_is_balanced_and_deterministic():
    rows = [100, 90, 50, 40, 30, 5]
    owners = assign_owners(rows, 3)
    assert owners == assign_owners(rows, 3)
    load = [sum(r for r, o in zip(rows, owners) if o == k) for k in range(3)]
    assert max(load) - min(load) <= 30  # LPT bound: within the largest small item


def test_stitch_renumbers_breadth_first(iris):
    X, y = iris.data, iris.target
    full = DecisionTreeClassifier(device="cpu").fit(X, y).flat_tree_
    trunk = DecisionTreeClassifier(max_depth=1, device="cpu").fit(X, y).flat_tree_
    right_rows = X[:, int(trunk.feature[0])] > trunk.thresholds[0]
    sub = DecisionTreeClassifier(device="cpu").fit(X[right_rows], y[right_rows]).flat_tree_
    # subtree saw only classes {1,2}: widen counts to the trunk's class space
    wide = np.zeros((sub.n_nodes, 3), dtype=np.int64)
    wide[:, 1:] = sub.counts
    sub.counts = wide
    merged = stitch_subtrees(trunk, [Graft(int(trunk.right[0]), sub)])
    assert merged.structurally_equal(full)


def _fd_rank(comm):
    import tempfile

    from mpitree_b200.parallel.fdpass import allgather_fds

    files = []
    for k in range(comm.rank % 3):  # ranks contribute 0, 1 or 2 descriptors
        fh = tempfile.TemporaryFile()
        fh.write(f"rank{comm.rank}-{k}".encode())
        fh.flush()
        files.append(fh)
    got = allgather_fds(comm, [fh.fileno() for fh in files], "t")
    out = []
    for fds in got:
        out.append([os.pread(fd, 64, 0).decode() for fd in fds])
        for fd in fds:
            os.close(fd)
    comm.barrier()
    return out


@pytest.mark.parametrize("world", [1, 2, 5, 16])
def test_descriptor_allgather_thread_ranks(world):
    want = [[f"rank{q}-{k}" for k in range(q % 3)] for q in range(world)]
    assert run_threaded(world, _fd_rank) == [want] * world


_WORKER = r"""
import os, sys
sys.path.insert(0, os.environ["REPO"])
import numpy as np, torch.distributed as dist
from sklearn.datasets import load_iris
from mpitree_b200.parallel.comm import init_process_group_from_env
from mpitree_b200.tree import ParallelDecisionTreeClassifier, DecisionTreeClassifier
assert init_process_group_from_env("gloo")
iris = load_iris()
clf = ParallelDecisionTreeClassifier(max_depth=3, device="cpu").fit(iris.data, iris.target)
assert clf.WORLD_SIZE == 2 and clf.WORLD_RANK == dist.get_rank()
want = DecisionTreeClassifier(max_depth=3, device="cpu").fit(iris.data, iris.target).export_text()
assert clf.export_text() == want, clf.export_text()
coop = ParallelDecisionTreeClassifier(max_depth=3, device="cpu", handoff_rows=0).fit(iris.data, iris.target)
assert coop.export_text() == want  # row-sharded histograms + all-reduce on every level
deep = ParallelDecisionTreeClassifier(device="cpu", handoff_rows=60).fit(iris.data, iris.target)
assert deep.export_text() == DecisionTreeClassifier(device="cpu").fit(iris.data, iris.target).export_text()
sub = clf._get_cyclic_dist(clf.WORLD_COMM, n_blocks=2)
assert sub.size == 1
meta = clf.WORLD_COMM.allgather_ints([dist.get_rank(), 7])
assert meta.tolist() == [[0, 7], [1, 7]]
# row-sharded continuous data (more distinct values than bins): both processes must bin identically
rng = np.random.default_rng(0)
Xc, yc = rng.random((400, 3)) * 100, rng.integers(0, 3, size=400)
half = slice(0, 200) if dist.get_rank() == 0 else slice(200, 400)
sh = ParallelDecisionTreeClassifier(max_depth=4, device="cpu", data_layout="sharded").fit(Xc[half], yc[half])
texts = clf.WORLD_COMM.allgather_object(sh.export_text())
assert texts[0] == texts[1]
# descriptors cross the process boundary (what the VMM / multicast handle exchange relies on)
import tempfile
from mpitree_b200.parallel.fdpass import allgather_fds
with tempfile.TemporaryFile() as fh:
    fh.write(b"from rank %d" % dist.get_rank()); fh.flush()
    got = allgather_fds(clf.WORLD_COMM, [fh.fileno()], "test")
assert [os.pread(g[0], 32, 0) for g in got] == [b"from rank 0", b"from rank 1"]
for g in got:
    os.close(g[0])
if dist.get_rank() == 0:
    print("GLOO_OK", clf.score(iris.data, iris.target))
dist.destroy_process_group()
"""


def test_two_process_gloo_torchrun(tmp_path):
    """The BASELINE.json plumbing config: iris, entropy, max_depth=3, 2 ranks on CPU."""
    import subprocess

    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, REPO=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29613", str(script)]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "GLOO_OK 0.97" in res.stdout


def _fit_sharded(comm, X, y, kw):
    cuts = [X.shape[0] * r // comm.size for r in range(comm.size + 1)]
    lo, hi = cuts[comm.rank], cuts[comm.rank + 1]
    clf = ParallelDecisionTreeClassifier(comm=comm, device="cpu", data_layout="sharded", **kw).fit(X[lo:hi], y[lo:hi])
    return clf.export_text(), clf.predict(X).tolist()


def test_sharded_quantile_bins_agree_across_ranks():
    """More distinct values than bins: every rank must derive the same edges from the merged summaries
    (a random sweep found rank-local quantiles, i.e. different histogram shapes per rank)."""
    rng = np.random.default_rng(542)
    X = rng.random((293, 2)) * 1000
    y = rng.integers(0, 4, size=293)
    for world in (2, 4):
        out = run_threaded(world, _fit_sharded, X, y, dict(max_depth=3, handoff_rows=0))
        assert all(o == out[0] for o in out)
    out = run_threaded(3, _fit_rank, X, y, dict(max_depth=3, handoff_rows=1000))  # replicated layout, hand-off path
    assert all(o == out[0] for o in out)


def test_sharded_ranks_may_hold_no_rows():
    X = np.array([[0.0, 1.0], [1.0, 0.0], [2.0, 1.0], [3.0, 0.0]])
    y = np.array([0, 0, 1, 1])
    want = DecisionTreeClassifier(device="cpu").fit(X, y)
    out = run_threaded(7, _fit_sharded, X, y, {})
    assert all(o == (want.export_text(), want.predict(X).tolist()) for o in out)


def test_ranks_agree_on_the_printed_sign_of_zero():
    """-0.0 and 0.0 in one column: replicated and row-sharded fits both print the reference's choice."""
    rng = np.random.default_rng(540)
    X = rng.normal(size=(99, 2)).round(1)
    y = rng.integers(0, 2, size=99)
    assert np.signbit(X[X == 0]).any() and not np.signbit(X[X == 0]).all()
    want = DecisionTreeClassifier(device="cpu").fit(X, y).export_text()
    out = run_threaded(7, _fit_rank, X, y, dict(handoff_rows=0))
    assert [o[0] for o in out] == [want] * 7
    out = run_threaded(3, _fit_sharded, X, y, dict(handoff_rows=0))  # the node's column is assembled in rank order
    assert [o[0] for o in out] == [want] * 3


def test_device_edge_agreement_with_an_empty_rank():
    """``_fit_mapper_device`` (the torch-tensor front end's K1 agreement) on CPU tensors: three ranks, one
    of them without rows, must end with identical edges that cover all the data."""
    import torch

    from mpitree_b200.ops.cuda_engine import _fit_mapper_device

    rng = np.random.default_rng(0)
    X = np.c_[rng.normal(size=(900, 2)), rng.integers(0, 5, size=(900, 1))].astype(np.float32)
    shards = [X[:500], X[:0], X[500:]]

    def work(comm):
        m = _fit_mapper_device(torch.from_numpy(shards[comm.rank]), 64, comm)
        return [e.tolist() for e in m.edges]

    out = run_threaded(3, work)
    assert out[0] == out[1] == out[2]
    assert out[0][2] == [0.0, 1.0, 2.0, 3.0, 4.0]  # few distinct values: kept exactly
    assert all(2 <= len(e) <= 64 and e == sorted(e) for e in out[0])
    assert out[0][0][-1] >= np.quantile(X[:, 0], 0.98)


def test_numpy_front_end_edge_agreement_on_thread_ranks():
    """``_fit_mapper_numpy`` feeds the CUDA engine for numpy inputs: the merged summaries must give every
    rank the same edges (uneven shards, one empty), and the edges must cover every value."""
    from mpitree_b200.ops import cuda_engine as ce

    rng = np.random.default_rng(1)
    X = np.c_[rng.normal(size=(5000, 2)), rng.integers(0, 7, size=(5000, 1))].astype(np.float32)
    X[4999, 0] = 1e6
    cuts = [0, 700, 700, 3100, 5000]  # uneven shards, one of them empty

    def work(comm):
        m = ce._fit_mapper_numpy(X[cuts[comm.rank] : cuts[comm.rank + 1]], 32, comm)
        return [e.tolist() for e in m.edges]

    out = run_threaded(4, work)
    assert all(o == out[0] for o in out)
    assert out[0][2] == [0.0, 1.0, 2.0, 3.0, 4.0, 5.0, 6.0]
    assert out[0][0][-1] == 1e6 and all(len(e) <= 32 for e in out[0])


def test_numpy_front_end_sampling_decision_is_collective(monkeypatch):
    """Only one shard is large enough to be sampled: the maximum exchange must still involve every rank."""
    from mpitree_b200.ops import cuda_engine as ce

    monkeypatch.setattr(ce, "SAMPLE_ROWS", 256)
    rng = np.random.default_rng(2)
    X = rng.normal(size=(3000, 2)).astype(np.float32)
    X[2999, 1] = 77.0  # odd position in the big shard: a strided sample skips it
    cuts = [0, 100, 3000]

    def work(comm):
        lo, hi = cuts[comm.rank], cuts[comm.rank + 1]
        idx = ce._sample_rows(hi - lo, 256)
        m = ce._fit_mapper_numpy(X[lo:hi], 16, comm)
        return idx is not None, [e.tolist() for e in m.edges]

    out = run_threaded(2, work)
    assert [o[0] for o in out] == [False, True]
    assert out[0][1] == out[1][1] and out[0][1][1][-1] == 77.0


def test_torch_front_end_dry_run_on_thread_ranks(monkeypatch):
    """``fit_from_torch`` on three thread-ranks with every native call replaced by a recorder (CPU tensors):
    uneven shards (one empty, only one large enough to be sampled) must pass through the label range,
    class and bin-edge agreements without a rank waiting in a collective the others skip, and end with
    identical bin edges."""
    import ctypes

    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.ops import cuda_engine as ce
    from mpitree_b200.ops.cpu_engine import SplitParams

    try:
        L.load()
    except Exception as exc:  # noqa: BLE001
        pytest.skip(f"native library unavailable: {exc}")

    class Stream:
        def synchronize(self):
            pass

    class Region:
        def __init__(self, nbytes):
            self.nbytes, self.mem = nbytes, torch.zeros(nbytes, dtype=torch.uint8)
            self.ptrs = [self.mem.data_ptr()] * 3

    monkeypatch.setattr(ce, "SAMPLE_ROWS", 256)
    monkeypatch.setattr(ce, "_call", lambda name, *a: None)
    monkeypatch.setattr(ce, "_stream_ptr", lambda: ctypes.c_void_p(0))
    monkeypatch.setattr(ce, "_device", lambda: torch.device("cpu"))
    monkeypatch.setattr(ce, "_dist_group", lambda comm: None)
    monkeypatch.setattr(ce, "_shared_region", lambda n, c: Region(n))
    monkeypatch.setattr(ce.DeviceDataset, "encode_float", lambda self, *a, **k: None)
    monkeypatch.setattr(ce.DeviceDataset, "encode_u8", lambda self, *a, **k: None)
    monkeypatch.setattr(ce.TreeBuilder, "_launch", lambda self, name: None)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: Stream())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "mem_get_info", lambda *a, **k: (8 << 30, 16 << 30))
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self, *a, **k: self)

    rng = np.random.default_rng(0)
    X = rng.normal(size=(3000, 4)).astype(np.float32)
    X[2999, 2] = 55.0
    y = rng.integers(0, 3, size=3000) * 1000  # label range too wide for the lazy path: classes are gathered
    cuts = [0, 100, 100, 3000]

    def work(comm):
        lo, hi = cuts[comm.rank], cuts[comm.rank + 1]
        flat, mapper, classes = ce.fit_from_torch(torch.from_numpy(X[lo:hi]), torch.from_numpy(y[lo:hi]),
                                                  SplitParams(max_depth=3), comm=comm)
        return [e.tolist() for e in mapper.edges], classes.tolist(), flat.n_nodes

    out = run_threaded(3, work)
    assert out[0] == out[1] == out[2]
    # (the outlier 55.0 is not in every rank's sample: it is clamped into the last bin, whose upper edge is never a
    # threshold, so the last edge need not be the column maximum; what matters is that all ranks agree)
    assert out[0][1] == [0, 1000, 2000] and out[0][0][2][-1] <= 55.0


@pytest.mark.parametrize("world", [4, 5, 8])
def test_rank_subgroups_keep_cooperating_on_medium_nodes(world, ref):
    """Phase 2 of the host scheduler: nodes below ``group_rows`` leave the all-ranks build for one half of the
    communicator (the reference's parity halving, decision_tree.py:336-338, :446-466), the half continues
    cooperatively on its own sub-communicator and halves again; nodes below ``handoff_rows`` end on single
    owners.  Every rank must end with the serial tree, which is the reference's tree."""
    from mpitree_b200.parallel import scheduler
    from mpitree_b200.parallel.comm import run_threaded
    from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

    rng = np.random.default_rng(world)
    n = 1500
    X = rng.integers(0, 9, size=(n, 5)).astype(np.float64)
    y = ((X[:, 0] + X[:, 1] * X[:, 2] + rng.integers(0, 5, size=n)) % 3).astype(np.int64)
    want = DecisionTreeClassifier(device="cpu").fit(X, y).export_text()
    assert want == ref.DecisionTreeClassifier().fit(X, y).export_text()
    splits = []
    real_split = scheduler.halving_groups

    def spy(ranks):
        splits.append(len(list(ranks)))
        return real_split(ranks)

    scheduler.halving_groups = spy
    try:
        cuts = [n * r // world for r in range(world + 1)]

        def work(comm):
            lo, hi = cuts[comm.rank], cuts[comm.rank + 1]
            clf = ParallelDecisionTreeClassifier(device="cpu", data_layout="sharded", handoff_rows=40, group_rows=600, comm=comm)
            return clf.fit(X[lo:hi], y[lo:hi]).export_text()

        out = run_threaded(world, work)
    finally:
        scheduler.halving_groups = real_split
    assert set(out) == {want}
    assert world in splits  # the world was halved ...
    if world >= 8:
        assert any(s < world for s in splits)  # ... and a half of four ranks halved again
