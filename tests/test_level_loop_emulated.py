"""The CUDA engine's whole Python driver on CPU tensors, with the kernels replaced by stand-ins.

``tests/emulated_kernels.py`` performs, per native call, the state transition the kernel performs on
the context's buffers.  ``TreeBuilder.build`` then runs unchanged — level loop (asynchronous and with
read-backs), ping-pong sides, context copies for the fused partition, tree growth, download, and the
hand-off phase — and must return the tree the host engine builds.
"""
import ctypes
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
