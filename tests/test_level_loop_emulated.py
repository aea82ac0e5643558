"""The CUDA engine's whole Python driver on CPU tensors, with the kernels replaced by stand-ins.

``tests/emulated_kernels.py`` performs, per native call, the state transition the kernel performs on
the context's buffers.  ``TreeBuilder.build`` then runs unchanged — level loop (asynchronous and with
read-backs), ping-pong sides, context copies for the fused partition, tree growth, download, and the
hand-off phase — and must return the tree the host engine builds.
"""
import ctypes
import os
import types

import numpy as np
import pytest

from mpitree_b200.ops.cpu_engine import SplitParams, build_tree_cpu

from . import emulated_kernels as emu
from .test_phase2_dry_run import _FakeLib


@pytest.fixture
def cpu_driver(monkeypatch):
    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.ops import cuda_engine as ce

    try:
        L.load()
    except Exception as exc:  # noqa: BLE001
        pytest.skip(f"native library unavailable: {exc}")

    class Stream:
        def synchronize(self):
            pass

    call = emu.dispatch()
    monkeypatch.setattr(ce, "_call", call)
    monkeypatch.setattr(ce.TreeBuilder, "_launch", lambda self, name: call(name, ctypes.byref(self.ctx)))
    monkeypatch.setattr(ce, "_stream_ptr", lambda: ctypes.c_void_p(0))
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: Stream())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.cuda, "mem_get_info", lambda *a, **k: (1 << 30, 2 << 30))
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self, *a, **k: self)
    return ce, L, monkeypatch


def _dataset(codes, y, C):
    import torch

    n, F = codes.shape
    G, n_cap = (F + 15) // 16, ((n + 1023) // 1024) * 1024
    planes = [torch.zeros(G * n_cap * 16, dtype=torch.uint8) for _ in range(2)]
    labels = [torch.zeros(n_cap, dtype=torch.uint8) for _ in range(2)]
    view = planes[0].numpy().reshape(G, n_cap, 16)
    for f in range(F):
        view[f // 16, :n, f % 16] = codes[:, f]
    labels[0].numpy()[:n] = y
    return types.SimpleNamespace(device=torch.device("cpu"), n_features=F, n_classes=C, G=G, n_cap=n_cap, n_local=n, cur=0,
                                 codes=planes, labels=labels)


def _data(seed, n=2500, F=20, C=4, B=32):
    rng = np.random.default_rng(seed)
    codes = rng.integers(0, B, size=(n, F)).astype(np.uint8)
    s = codes.astype(float) @ rng.normal(size=F) + rng.normal(scale=25, size=n)
    y = np.digitize(s, np.quantile(s, np.linspace(0, 1, C + 1)[1:-1]))
    return codes, y


def _same(a, b):
    assert a.n_nodes == b.n_nodes
    for name in ("feature", "bin", "left", "right", "parent", "depth", "counts"):
        assert np.array_equal(getattr(a, name), getattr(b, name)), name
    assert np.array_equal(np.nan_to_num(a.gain, nan=-1.0), np.nan_to_num(b.gain, nan=-1.0))


@pytest.mark.parametrize("mode", [dict(), dict(fuse_partition=False), dict(fused=False), dict(async_levels=False)])
@pytest.mark.parametrize("max_depth,mss,crit", [(6, 2, "entropy"), (4, 40, "gini"), (1, 2, "entropy")])
def test_bounded_depth_fits(cpu_driver, mode, max_depth, mss, crit):
    ce, _, _ = cpu_driver
    codes, y = _data(1)
    params = SplitParams(max_depth=max_depth, min_samples_split=mss, criterion=crit)
    want = build_tree_cpu(codes, y, np.full(20, 32), 4, params)
    b = ce.TreeBuilder(_dataset(codes, y, 4), 32, params, hist_budget_bytes=256 << 20, **mode)
    got = b.build()
    assert b.async_levels == (mode.get("async_levels", True))
    _same(got, want)


@pytest.mark.parametrize("small_rows", [0, 150])
def test_unbounded_depth_with_and_without_hand_off(cpu_driver, small_rows):
    ce, L, monkeypatch = cpu_driver
    codes, y = _data(2, n=1800)
    params = SplitParams(max_depth=None)
    want = build_tree_cpu(codes, y, np.full(20, 32), 4, params)
    b = ce.TreeBuilder(_dataset(codes, y, 4), 32, params, hist_budget_bytes=512 << 20, small_rows=small_rows)
    fake = _FakeLib(np.full(20, 32), "entropy")
    monkeypatch.setattr(L, "load", lambda: fake)  # phase 2 asks the library for the subtree kernel
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    got = b.build()
    assert not b.async_levels and (b.stats.levels > 8 or small_rows)
    assert (b.stats.handed_off > 10) == bool(small_rows) and fake.launches == (1 if small_rows else 0)
    _same(got, want)


def _emulate_encode(ce, monkeypatch):
    """``encode_float`` / ``encode_u8`` stand-ins: host binning written into the tiled planes."""
    import torch

    def write(self, codes, row0):
        n, F = codes.shape
        view = self.codes[self.cur].numpy()[: self.G * self.n_cap * 16].reshape(self.G, self.n_cap, 16)
        for f in range(F):
            view[f // 16, row0 : row0 + n, f % 16] = codes[:, f]

    def encode_float(self, X, mapper, row0=0, mismatch=None):
        write(self, mapper.transform(X.numpy()), row0)

    def encode_u8(self, X, lut=None, row0=0):
        codes = X.numpy()
        if lut is not None:  # (F, 256) value -> code table
            codes = np.take_along_axis(lut.numpy().T, codes.astype(np.int64), axis=0)
        write(self, codes, row0)

    monkeypatch.setattr(ce.DeviceDataset, "encode_float", encode_float)
    monkeypatch.setattr(ce.DeviceDataset, "encode_u8", encode_u8)
    monkeypatch.setattr(ce, "_device", lambda: torch.device("cpu"))
    monkeypatch.setattr(ce, "_dist_group", lambda comm: None)
    monkeypatch.setattr(ce, "_WORKSPACE", {})


@pytest.mark.parametrize("kind", ["float32", "int64", "uint8", "negative zeros"])
def test_estimator_through_the_cuda_front_end(cpu_driver, kind):
    """``DecisionTreeClassifier(device="cuda").fit`` on numpy input with the native layer emulated: front end
    (binning, label codes, thresholds in the input's type, printed zero signs) + driver == host estimator."""
    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.tree import DecisionTreeClassifier

    ce, L, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    real = L.load()

    class Hybrid(_FakeLib):  # phase 2 entry points emulated, everything else (host-only queries) from the real library
        def __getattr__(self, name):
            return getattr(real, name)

    hybrid = Hybrid(np.full(5, 256), "entropy")
    monkeypatch.setattr(L, "load", lambda: hybrid)
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    rng = np.random.default_rng(3)
    n = 900
    if kind == "float32":
        X = rng.normal(size=(n, 5)).astype(np.float32).round(1)
    elif kind == "int64":
        X = rng.integers(-20, 20, size=(n, 5))
    elif kind == "uint8":
        X = rng.integers(0, 40, size=(n, 5)).astype(np.uint8)
    else:
        X = rng.normal(size=(n, 5)).round(1)
        X[rng.random((n, 5)) < 0.1] = -0.0
        X[rng.random((n, 5)) < 0.1] = 0.0
    y = (X[:, 0].astype(float) + X[:, 1] * 0.5 + rng.normal(scale=3 if kind in ("int64", "uint8") else 0.5, size=n) > X[:, 0].mean()) * 1
    y = y + 2 * (X[:, 2] > np.median(X[:, 2]))  # labels 0..3
    for max_depth in (3, None):
        want = DecisionTreeClassifier(max_depth=max_depth, device="cpu").fit(X, y)
        got = DecisionTreeClassifier(max_depth=max_depth, device="cuda").fit(X, y)
        assert got.engine_ == "cuda" and want.engine_ == "cpu"
        assert got.export_text() == want.export_text()
        assert [type(t) for t in got.flat_tree_.thresholds] == [type(t) for t in want.flat_tree_.thresholds]
        assert np.array_equal(got.classes_, want.classes_) and np.array_equal(got.flat_tree_.counts, want.flat_tree_.counts)


@pytest.mark.parametrize("labels", ["dense", "gappy", "wide"])
def test_estimator_from_torch_tensors(cpu_driver, labels):
    """The torch-tensor front end (labels used as class codes and read back from the root's counts, or
    gathered when their range is wide) with the native layer emulated."""
    import torch

    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.tree import DecisionTreeClassifier

    ce, _, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    rng = np.random.default_rng(4)
    X = rng.integers(0, 30, size=(1200, 6)).astype(np.uint8)
    y = (X[:, 0] > 14).astype(np.int64) + 2 * (X[:, 3] > 20)
    y = {"dense": y, "gappy": y * 3, "wide": y * 1000}[labels]
    want = DecisionTreeClassifier(max_depth=5, device="cpu").fit(X, y)
    got = DecisionTreeClassifier(max_depth=5, device="cuda").fit(torch.from_numpy(X), torch.from_numpy(y))
    assert got.export_text() == want.export_text()
    assert np.array_equal(got.classes_, want.classes_)
    assert np.array_equal(got.flat_tree_.counts, want.flat_tree_.counts)
    assert np.array_equal(got.predict(X), want.predict(X))  # numpy input: host walk over the same flat tree
    # fewer bins than byte values: the byte rows are encoded through a value -> code table
    want8 = DecisionTreeClassifier(max_depth=5, max_bins=8, device="cpu").fit(X, y)
    got8 = DecisionTreeClassifier(max_depth=5, max_bins=8, device="cuda").fit(torch.from_numpy(X), torch.from_numpy(y))
    assert got8.export_text() == want8.export_text() and int(got8.bin_mapper_.n_bins.max()) <= 8


def test_slot_capacity_is_enforced_or_avoided_by_hand_off(cpu_driver):
    """A frontier that outgrows the histogram slots raises with hand-off switched off; the automatic policy
    (hand-off whenever the depth is unbounded) keeps the same fit within capacity."""
    ce, L, monkeypatch = cpu_driver
    codes, y = _data(6, n=2000)
    params = SplitParams(max_depth=None)
    want = build_tree_cpu(codes, y, np.full(20, 32), 4, params)
    b = ce.TreeBuilder(_dataset(codes, y, 4), 32, params, h_cap=16, small_rows=0)
    with pytest.raises(RuntimeError, match="exceeds the histogram slot capacity"):
        b.build()
    real = L.load()

    class Hybrid(_FakeLib):
        def __getattr__(self, name):
            return getattr(real, name)

    monkeypatch.setattr(L, "load", lambda: Hybrid(np.full(20, 32), "entropy"))
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    monkeypatch.setenv("MTB_SMALL_ROWS", "300")
    b = ce.TreeBuilder(_dataset(codes, y, 4), 32, params, h_cap=16)
    assert b.small_rows == 300
    got = b.build()
    assert b.stats.handed_off > 4
    _same(got, want)


def test_tree_arrays_grow_with_the_tree(cpu_driver, monkeypatch):
    """Unbounded depth starts with a modest node capacity and reallocates the tree arrays as levels are added."""
    ce, _, _ = cpu_driver
    codes, y = _data(7, n=1500)
    params = SplitParams(max_depth=None)
    want = build_tree_cpu(codes, y, np.full(20, 32), 4, params)
    b = ce.TreeBuilder(_dataset(codes, y, 4), 32, params, hist_budget_bytes=512 << 20, small_rows=0)
    b._alloc_tree(8)  # force several reallocations
    grown = []
    orig = b._alloc_tree
    monkeypatch.setattr(b, "_alloc_tree", lambda cap, old=None: (grown.append(cap), orig(cap, old))[1])
    got = b.build()
    assert len(grown) >= 3 and b.node_cap >= want.n_nodes
    _same(got, want)


@pytest.mark.parametrize("layout,max_depth", [("sharded", 5), ("replicated", 5), ("sharded", None)])
def test_parallel_estimator_on_thread_ranks_through_the_cuda_engine(cpu_driver, layout, max_depth):
    """Three thread-ranks through ``ParallelDecisionTreeClassifier(device="cuda")``: numpy front end with its
    agreements, multi-rank builder set-up, level loop, and (unbounded depth) the multi-rank hand-off phase.
    The peer reduction inside the level kernel is replaced by a sum over the ranks' local histograms."""
    import threading

    import torch

    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.parallel import symm
    from mpitree_b200.parallel.comm import run_threaded
    from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

    from .test_phase2_dry_run import _fake_gather_rows, _ThreadRegion

    ce, L, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    reducers: dict = {}
    call = emu.dispatch(reducers)
    monkeypatch.setattr(ce, "_call", call)
    monkeypatch.setattr(ce.TreeBuilder, "_launch", lambda self, name: call(name, ctypes.byref(self.ctx)))
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    monkeypatch.setattr(dt, "REPLICATED_SERIAL_ROWS", 0)  # take the cooperative path for the replicated layout too
    monkeypatch.setattr(ce, "_shared_region", lambda nbytes, comm: _ThreadRegion(nbytes, comm))
    monkeypatch.setattr(symm, "IpcRegion", _ThreadRegion)
    real = L.load()

    class Hybrid(_FakeLib):
        def __getattr__(self, name):
            return getattr(real, name)

    hybrid = Hybrid(np.full(6, 256), "entropy")
    hybrid.mtb_gather_rows = _fake_gather_rows
    monkeypatch.setattr(L, "load", lambda: hybrid)
    monkeypatch.setattr(L, "check", lambda rc, what: None)

    rng = np.random.default_rng(8)
    n = 1500
    X = rng.integers(0, 25, size=(n, 6)).astype(np.float32)
    y = (X[:, 0] > 12).astype(np.int64) + 2 * (X[:, 2] + X[:, 4] > 24)
    want = DecisionTreeClassifier(max_depth=max_depth, device="cpu").fit(X, y)
    cuts = [0, 300, 1100, n]

    def work(comm):
        reducers[threading.get_ident()] = lambda hists: [sum(parts) for parts in zip(*comm.allgather_object(hists))]
        lo, hi = (0, n) if layout == "replicated" else (cuts[comm.rank], cuts[comm.rank + 1])
        clf = ParallelDecisionTreeClassifier(max_depth=max_depth, device="cuda", comm=comm, data_layout=layout)
        clf.fit(X[lo:hi], y[lo:hi])
        return clf.export_text(), clf.flat_tree_.counts.tolist(), clf.engine_

    out = run_threaded(3, work)
    assert all(o == (want.export_text(), want.flat_tree_.counts.tolist(), "cuda") for o in out)


@pytest.mark.parametrize("seed", range(int(os.environ.get("MTB_FUZZ_CASES", "24"))))
def test_random_shapes_and_modes(cpu_driver, seed):
    """Seeded fuzz of the driver: random shapes, stop rules, launch modes and hand-off sizes
    (``MTB_FUZZ_CASES=1000`` runs a longer sweep)."""
    ce, L, monkeypatch = cpu_driver
    rng = np.random.default_rng(1000 + seed)
    n, F, C, B = int(rng.integers(1, 900)), int(rng.integers(1, 40)), int(rng.integers(1, 6)), int(rng.integers(2, 64))
    codes = rng.integers(0, B, size=(n, F)).astype(np.uint8)
    if rng.random() < 0.3:
        codes[:, 0] = 3 % B  # a constant feature
    y = rng.integers(0, C, size=n)
    max_depth = [None, 0, 1, 3, 7][int(rng.integers(0, 5))]
    params = SplitParams(max_depth=max_depth, min_samples_split=int(rng.choice([2, 2, 5, 30])),
                         criterion=["entropy", "gini"][int(rng.integers(0, 2))])
    mode = [dict(), dict(fuse_partition=False), dict(fused=False), dict(async_levels=False)][int(rng.integers(0, 4))]
    small_rows = [0, 0, 16, 100][int(rng.integers(0, 4))]
    want = build_tree_cpu(codes, y, np.full(F, B), C, params)
    real = L.load()

    class Hybrid(_FakeLib):
        def __getattr__(self, name):
            return getattr(real, name)

    monkeypatch.setattr(L, "load", lambda: Hybrid(np.full(F, 64), params.criterion))
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    b = ce.TreeBuilder(_dataset(codes, y, C), B, params, hist_budget_bytes=1 << 30, small_rows=small_rows, **mode)
    _same(b.build(), want)


@pytest.mark.parametrize("signed_zeros", [False, True])
@pytest.mark.parametrize("layout", ["sharded", "replicated"])
def test_parallel_estimator_from_torch_tensors_on_thread_ranks(cpu_driver, layout, signed_zeros):
    """The torch-tensor front end on three thread-ranks (uneven shards, one of them empty in the sharded case)."""
    import threading

    import torch

    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.parallel.comm import run_threaded
    from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

    from .test_phase2_dry_run import _ThreadRegion

    ce, L, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    reducers: dict = {}
    call = emu.dispatch(reducers)
    monkeypatch.setattr(ce, "_call", call)
    monkeypatch.setattr(ce.TreeBuilder, "_launch", lambda self, name: call(name, ctypes.byref(self.ctx)))
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    monkeypatch.setattr(ce, "_shared_region", lambda nbytes, comm: _ThreadRegion(nbytes, comm))

    rng = np.random.default_rng(12)
    n = 1400
    X = rng.integers(0, 50, size=(n, 5)).astype(np.uint8)
    y = (X[:, 1] > 20).astype(np.int64) * 3 + (X[:, 4] > 35)  # labels 0, 1, 3, 4: read back from the root counts
    if signed_zeros:  # float rows holding both zeros: host tensors reproduce the printed sign in both layouts
        X = rng.normal(size=(n, 5)).round(1).astype(np.float32)
        X[rng.random((n, 5)) < 0.1] = -0.0
        X[rng.random((n, 5)) < 0.1] = 0.0
        y = (X[:, 1] > 0).astype(np.int64) * 3 + (X[:, 4] > 0.5)
    want = DecisionTreeClassifier(max_depth=6, device="cpu").fit(X, y)
    cuts = [0, 0, 600, n]

    def work(comm):
        reducers[threading.get_ident()] = lambda hists: [sum(parts) for parts in zip(*comm.allgather_object(hists))]
        lo, hi = (0, n) if layout == "replicated" else (cuts[comm.rank], cuts[comm.rank + 1])
        clf = ParallelDecisionTreeClassifier(max_depth=6, device="cuda", comm=comm, data_layout=layout)
        clf.fit(torch.from_numpy(X[lo:hi]), torch.from_numpy(y[lo:hi]))
        return clf.export_text(), clf.classes_.tolist(), clf.flat_tree_.counts.tolist()

    out = run_threaded(3, work)
    assert all(o == (want.export_text(), want.classes_.tolist(), want.flat_tree_.counts.tolist()) for o in out)


@pytest.mark.parametrize("kind", ["uint8", "float32", "float64", "int64"])
def test_predict_front_end(cpu_driver, kind):
    """``predict`` / ``predict_proba`` / ``apply`` through the CUDA front end (node packing, thresholds per
    input type, chunking, label look-up) with the predict kernel emulated, for numpy and torch inputs."""
    import torch

    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.tree import DecisionTreeClassifier

    ce, _, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    monkeypatch.setattr(ce, "available", lambda: True)
    monkeypatch.setattr(ce, "ENCODE_CHUNK_ROWS", 300)  # several chunks for host inputs
    rng = np.random.default_rng(21)
    n = 1000
    X = {"uint8": rng.integers(0, 60, size=(n, 6)).astype(np.uint8), "float32": rng.normal(size=(n, 6)).astype(np.float32),
         "float64": rng.normal(size=(n, 6)).round(3), "int64": rng.integers(-30, 30, size=(n, 6))}[kind]
    y = ((X[:, 0] > np.median(X[:, 0])) * 2 + (X[:, 3] > np.median(X[:, 3]))) * 2  # labels 0, 2, 4, 6
    want = DecisionTreeClassifier(max_depth=6, device="cpu").fit(X, y)
    got = DecisionTreeClassifier(max_depth=6, device="cuda").fit(X, y)
    Xn = np.concatenate([X, X[:50] + (1 if kind in ("uint8", "int64") else 0.37)]).astype(X.dtype)  # seen and unseen rows
    assert np.array_equal(got.predict(Xn), want.predict(Xn))
    assert np.array_equal(got.predict_proba(Xn), want.predict_proba(Xn))
    assert np.array_equal(got.apply(Xn), want.apply(Xn))
    Xt = torch.from_numpy(Xn)
    assert np.array_equal(got.predict(Xt).numpy(), want.predict(Xn))
    assert np.array_equal(got.predict_proba(Xt).numpy(), want.predict_proba(Xn))
    with pytest.raises(ValueError, match="expecting 6 features"):
        got.predict(Xt[:, :5])


@pytest.mark.parametrize("seed", range(int(os.environ.get("MTB_FUZZ_CASES", "16"))))
def test_estimator_fuzz_through_the_cuda_front_end(cpu_driver, seed):
    """Seeded fuzz at estimator level: numeric dtypes, gappy labels, quantile-binned columns, stop rules."""
    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.tree import DecisionTreeClassifier

    ce, L, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    monkeypatch.setattr(ce, "available", lambda: True)
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    real = L.load()

    class Hybrid(_FakeLib):
        def __getattr__(self, name):
            return getattr(real, name)

    rng = np.random.default_rng(5000 + seed)
    n, F, k = int(rng.integers(2, 700)), int(rng.integers(1, 9)), int(rng.integers(1, 6))
    kind = int(rng.integers(0, 6))
    X = [rng.normal(size=(n, F)).astype(np.float32), rng.normal(size=(n, F)).round(1), rng.integers(-9, 9, size=(n, F)),
         rng.integers(0, 200, size=(n, F)).astype(np.uint8), rng.random((n, F)) < 0.3, rng.random((n, F)) * 1e4][kind]
    y = rng.integers(0, k, size=n) * int(rng.integers(1, 4))
    crit = ["entropy", "gini"][int(rng.integers(0, 2))]
    kw = dict(max_depth=[None, 0, 2, 6][int(rng.integers(0, 4))], min_samples_split=int(rng.choice([2, 2, 7])),
              max_bins=int(rng.choice([256, 256, 32, 5])), criterion=crit)
    monkeypatch.setattr(L, "load", lambda: Hybrid(np.full(F, 256), crit))
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    want = DecisionTreeClassifier(device="cpu", **kw).fit(X, y)
    got = DecisionTreeClassifier(device="cuda", **kw).fit(X, y)
    assert got.engine_ == "cuda"
    assert got.export_text() == want.export_text()
    assert [type(t) for t in got.flat_tree_.thresholds] == [type(t) for t in want.flat_tree_.thresholds]
    assert np.array_equal(got.classes_, want.classes_)
    assert np.array_equal(got.predict(X), want.predict(X)) and np.array_equal(got.predict_proba(X), want.predict_proba(X))


@pytest.mark.parametrize("seed", range(max(4, int(os.environ.get("MTB_FUZZ_CASES", "8")) // 2)))
def test_parallel_estimator_fuzz_through_the_cuda_engine(cpu_driver, seed):
    """Seeded fuzz of the multi-rank flow on thread-ranks: world size, layout, input type, shard sizes
    (empty shards included), stop rules and hand-off sizes; every rank must end with the serial tree."""
    import threading

    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.parallel import symm
    from mpitree_b200.parallel.comm import run_threaded
    from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier

    from .test_phase2_dry_run import _fake_gather_rows, _fake_memcpy_async, _ThreadRegion

    ce, L, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    reducers: dict = {}
    call = emu.dispatch(reducers)
    monkeypatch.setattr(ce, "_call", call)
    monkeypatch.setattr(ce.TreeBuilder, "_launch", lambda self, name: call(name, ctypes.byref(self.ctx)))
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    monkeypatch.setattr(dt, "REPLICATED_SERIAL_ROWS", 0)
    monkeypatch.setattr(ce, "_shared_region", lambda nbytes, comm: _ThreadRegion(nbytes, comm))
    monkeypatch.setattr(symm, "IpcRegion", _ThreadRegion)
    real = L.load()

    class Hybrid(_FakeLib):
        def __getattr__(self, name):
            return getattr(real, name)

    rng = np.random.default_rng(9000 + seed)
    n, F, k = int(rng.integers(3, 900)), int(rng.integers(1, 7)), int(rng.integers(1, 5))
    kind = int(rng.integers(0, 3))
    X = [rng.integers(0, 40, size=(n, F)).astype(np.uint8), rng.normal(size=(n, F)).round(1).astype(np.float32),
         rng.integers(-5, 6, size=(n, F))][kind]
    y = rng.integers(0, k, size=n)
    world = int(rng.choice([2, 3, 4]))
    layout = ["sharded", "replicated"][int(rng.integers(0, 2))]
    crit = ["entropy", "gini"][int(rng.integers(0, 2))]
    kw = dict(max_depth=[None, 1, 4, 8][int(rng.integers(0, 4))], min_samples_split=int(rng.choice([2, 2, 9])), criterion=crit)
    handoff = [None, 0, 25, 200][int(rng.integers(0, 4))]
    hybrid = Hybrid(np.full(F, 256), crit)
    hybrid.mtb_gather_rows, hybrid.mtb_memcpy_async = _fake_gather_rows, _fake_memcpy_async
    monkeypatch.setattr(L, "load", lambda: hybrid)
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    from mpitree_b200.ops import subtree_phase

    monkeypatch.setattr(subtree_phase, "DEVICE_EXCHANGE_MIN_BYTES", [0, 1 << 40][seed % 2])
    want = DecisionTreeClassifier(device="cpu", **kw).fit(X, y)
    cuts = np.sort(np.concatenate([[0, n], rng.integers(0, n + 1, size=world - 1)]))

    def work(comm):
        reducers[threading.get_ident()] = lambda hists: [sum(parts) for parts in zip(*comm.allgather_object(hists))]
        lo, hi = (0, n) if layout == "replicated" else (int(cuts[comm.rank]), int(cuts[comm.rank + 1]))
        clf = ParallelDecisionTreeClassifier(device="cuda", comm=comm, data_layout=layout, handoff_rows=handoff, **kw)
        clf.fit(X[lo:hi], y[lo:hi])
        return clf.export_text(), clf.flat_tree_.counts.tolist()

    out = run_threaded(world, work)
    assert all(o == (want.export_text(), want.flat_tree_.counts.tolist()) for o in out)


@pytest.mark.parametrize("seed", range(int(os.environ.get("MTB_FUZZ_CASES", "12"))))
def test_torch_front_end_fuzz(cpu_driver, seed):
    """Seeded fuzz of fits from torch tensors: input and label dtypes, label ranges, stop rules, ``max_bins``."""
    import torch

    import mpitree_b200.tree.decision_tree as dt
    from mpitree_b200.tree import DecisionTreeClassifier

    ce, L, monkeypatch = cpu_driver
    _emulate_encode(ce, monkeypatch)
    monkeypatch.setattr(dt, "resolve_engine", lambda device, **kw: "cuda" if device == "cuda" else "cpu")
    real = L.load()

    class Hybrid(_FakeLib):
        def __getattr__(self, name):
            return getattr(real, name)

    rng = np.random.default_rng(7000 + seed)
    n, F, k = int(rng.integers(2, 600)), int(rng.integers(1, 8)), int(rng.integers(1, 6))
    kind = int(rng.integers(0, 5))
    X = [rng.integers(0, 90, size=(n, F)).astype(np.uint8), rng.normal(size=(n, F)).round(1).astype(np.float32),
         rng.normal(size=(n, F)).round(2), rng.integers(-40, 40, size=(n, F)).astype(np.int32),
         rng.integers(0, 1000, size=(n, F)).astype(np.int64)][kind]
    y = rng.integers(0, k, size=n) * int(rng.choice([1, 1, 2, 500]))
    y = y.astype([np.int64, np.int32, np.uint8][int(rng.integers(0, 3))] if y.max() < 256 else np.int64)
    crit = ["entropy", "gini"][int(rng.integers(0, 2))]
    kw = dict(max_depth=[None, 0, 3, 7][int(rng.integers(0, 4))], min_samples_split=int(rng.choice([2, 2, 6])),
              max_bins=int(rng.choice([256, 256, 16])), criterion=crit)
    monkeypatch.setattr(L, "load", lambda: Hybrid(np.full(F, 256), crit))
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    if X.dtype.kind == "f" and any(len(np.unique(X[:, f])) > kw["max_bins"] for f in range(F)):
        X = X + 0.0  # a quantile edge of 0 keeps the sign its mapper kept (host and device mappers may differ)
    want = DecisionTreeClassifier(device="cpu", **kw).fit(X, y)
    got = DecisionTreeClassifier(device="cuda", **kw).fit(torch.from_numpy(X), torch.from_numpy(y))
    assert got.export_text() == want.export_text()  # incl. the printed sign of zero thresholds (host tensors)
    assert np.array_equal(got.classes_, want.classes_) and np.array_equal(got.flat_tree_.counts, want.flat_tree_.counts)
