"""Phase 2 (hand-off) driver of the CUDA engine on CPU tensors with an emulated subtree kernel.

``ops/subtree_phase.py::finish_subtrees`` is host code around two kernels.  Here the subtree kernel
is replaced by a Python stand-in that honours the same argument block (``MtbSubtreeArgs``: item
lists in, flat node records out, worst-case sized output arrays of which only a prefix per item is
used) and grows each subtree with the host engine.  Everything else is the production driver:
item ordering, output sizing, compaction of the used records, stitching and renumbering.  The
stitched tree must equal the tree grown level-wise in one go.
"""
import ctypes
import types

import numpy as np
import pytest

from mpitree_b200.ops.cpu_engine import SplitParams, build_tree_cpu


def _view(ptr, dtype, count):
    ctype = {np.int32: ctypes.c_int32, np.uint32: ctypes.c_uint32, np.uint8: ctypes.c_uint8, np.float64: ctypes.c_double}[dtype]
    return np.ctypeslib.as_array(ctypes.cast(int(ptr), ctypes.POINTER(ctype)), shape=(int(count),))


class _FakeLib:
    """Stands in for libmpitree_b200.so: only what phase 2 calls on one rank."""

    def __init__(self, n_bins, criterion):
        self.n_bins, self.criterion, self.launches = n_bins, criterion, 0

    def mtb_subtree_config(self, B, C):
        return 8

    def mtb_sm_count(self, device):
        return 4

    def mtb_subtree_build(self, args_ref, grid, stream):
        a = args_ref._obj
        self.launches += 1
        n_items, C, F, G, n_cap = int(a.n_items), int(a.C), int(a.F), int(a.G), int(a.n_cap)
        side, begin, end = (_view(p, np.int32, n_items) for p in (a.it_side, a.it_begin, a.it_end))
        depth0, off = _view(a.it_depth, np.int32, n_items), _view(a.it_off, np.int32, n_items)
        total = int(off[-1] + 2 * (end[-1] - begin[-1]) - 1)
        o_feature, o_bin, o_left, o_depth = (_view(p, np.int32, total) for p in (a.o_feature, a.o_bin, a.o_left, a.o_depth))
        o_gain, o_counts, o_nn = _view(a.o_gain, np.float64, total), _view(a.o_counts, np.uint32, total * C), _view(a.o_nn, np.int32, n_items)
        planes = [_view(p, np.uint8, G * n_cap * 16).reshape(G, n_cap, 16) for p in (a.codes0, a.codes1)]
        labels = [_view(p, np.uint8, n_cap) for p in (a.labels0, a.labels1)]
        for i in range(n_items):
            rows = np.arange(begin[i], end[i])
            codes = np.stack([planes[side[i]][f // 16, rows, f % 16] for f in range(F)], axis=1)
            md = None if a.max_depth < 0 else int(a.max_depth) - int(depth0[i])
            sub = build_tree_cpu(codes, labels[side[i]][rows].astype(np.int64), self.n_bins, C,
                                 SplitParams(max_depth=md, min_samples_split=int(a.min_samples_split), criterion=self.criterion))
            k, o = sub.n_nodes, int(off[i])
            assert k <= 2 * len(rows) - 1
            o_feature[o : o + k], o_bin[o : o + k], o_left[o : o + k] = sub.feature, sub.bin, sub.left
            o_depth[o : o + k], o_gain[o : o + k] = sub.depth + depth0[i], sub.gain
            o_counts[o * C : (o + k) * C] = sub.counts.reshape(-1)
            o_nn[i] = k
        return 0


@pytest.mark.parametrize("max_depth", [None, 9])
def test_finish_subtrees_with_emulated_kernel(monkeypatch, max_depth):
    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.ops import cuda_engine as ce
    from mpitree_b200.ops import subtree_phase as sp

    rng = np.random.default_rng(5)
    n, F, C, B = 3000, 20, 4, 32
    codes = rng.integers(0, B, size=(n, F)).astype(np.uint8)
    w = rng.normal(size=F)
    s = codes.astype(float) @ w + rng.normal(scale=25, size=n)
    y = np.digitize(s, np.quantile(s, [0.25, 0.5, 0.75]))
    n_bins = np.full(F, B)
    params = SplitParams(max_depth=max_depth)
    want = build_tree_cpu(codes, y, n_bins, C, params)
    pending: dict = {}
    trunk = build_tree_cpu(codes, y, n_bins, C, params, handoff_rows=120, pending_out=pending)
    node_ids = sorted(pending)
    assert len(node_ids) > 10 and trunk.n_nodes < want.n_nodes

    # the level-wise kernels leave the rows of every handed-off node contiguous: lay them out like that,
    # alternating the ping-pong side and listing the nodes out of order (the commit phase appends them
    # in whatever order CTAs finish)
    G, n_cap = (F + 15) // 16, 4096
    planes = [torch.zeros(G * n_cap * 16, dtype=torch.uint8) for _ in range(2)]
    labels = [torch.zeros(n_cap, dtype=torch.uint8) for _ in range(2)]
    listed = list(rng.permutation(node_ids))
    ho = {k: torch.zeros(len(listed) + 5, dtype=torch.int32) for k in ("node", "begin", "end", "side")}
    cursor = 7
    for j, nid in enumerate(listed):
        rows = pending[nid]
        side = j % 2
        view = planes[side].numpy().reshape(G, n_cap, 16)
        for f in range(F):
            view[f // 16, cursor : cursor + len(rows), f % 16] = codes[rows, f]
        labels[side].numpy()[cursor : cursor + len(rows)] = y[rows]
        ho["node"][j], ho["begin"][j], ho["end"][j], ho["side"][j] = int(nid), cursor, cursor + len(rows), side
        cursor += len(rows) + 3

    lib = _FakeLib(n_bins, "entropy")
    monkeypatch.setattr(L, "load", lambda: lib)
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    monkeypatch.setattr(ce, "_stream_ptr", lambda: ctypes.c_void_p(0))
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    host_scal = torch.zeros(8, dtype=torch.int32)
    host_scal[5] = len(listed)
    ctx = types.SimpleNamespace(max_depth=-1 if max_depth is None else max_depth, min_samples_split=2, criterion=0,
                                dot_mode=0, plogp=0, plogp_n=0, klogk=0, klogk_n=0)
    data = types.SimpleNamespace(device=torch.device("cpu"), G=G, n_cap=n_cap, codes=planes, labels=labels)
    builder = types.SimpleNamespace(_host_scal=host_scal, ho=ho, ho_cap=len(listed) + 5, data=data, comm=None, C=C, F=F, G=G,
                                    B=B, small_rows=120, ctx=ctx, stats=types.SimpleNamespace(handed_off=0))
    got = sp.finish_subtrees(builder, trunk)
    assert lib.launches == 1 and builder.stats.handed_off == len(listed)
    assert got.n_nodes == want.n_nodes
    for name in ("feature", "bin", "left", "right", "parent", "depth", "counts"):
        assert np.array_equal(getattr(got, name), getattr(want, name)), name
    assert np.array_equal(np.nan_to_num(got.gain, nan=-1.0), np.nan_to_num(want.gain, nan=-1.0))


class _ThreadRegion:
    """``IpcRegion`` stand-in for thread-ranks: one process, so plain pointers already reach every 'peer'."""

    def __init__(self, nbytes, comm):
        import torch

        self.mem = torch.zeros(int(nbytes), dtype=torch.uint8)
        self.world, self.rank, self.nbytes = comm.size, comm.rank, int(nbytes)
        self.ptrs = comm.allgather_object(int(self.mem.data_ptr()))
        self._comm = comm
        comm.barrier()

    def close(self):
        self._comm.barrier()


def _fake_gather_rows(codes0, codes1, labels0, labels1, n_cap, G, side, begin, end, dst_codes, dst_labels, dst_cap,
                      dst_begin, n_ho, stream):
    """``mtb_gather_rows``: copy each listed row segment into its owner's receive planes."""
    n_cap, G, n_ho = n_cap.value, G.value, n_ho.value
    planes = [_view(p.value, np.uint8, G * n_cap * 16).reshape(G, n_cap, 16) for p in (codes0, codes1)]
    labels = [_view(p.value, np.uint8, n_cap) for p in (labels0, labels1)]
    side, begin, end, dst_begin = (_view(p.value, np.int32, n_ho) for p in (side, begin, end, dst_begin))
    dst_codes, dst_labels, dst_cap = (np.ctypeslib.as_array(ctypes.cast(p.value, ctypes.POINTER(ctypes.c_int64)), shape=(n_ho,))
                                      for p in (dst_codes, dst_labels, dst_cap))
    for i in range(n_ho):
        m = int(end[i] - begin[i])
        if m == 0:
            continue
        cap = int(dst_cap[i])
        out = _view(dst_codes[i], np.uint8, G * cap * 16).reshape(G, cap, 16)
        out[:, dst_begin[i] : dst_begin[i] + m] = planes[side[i]][:, begin[i] : end[i]]
        _view(dst_labels[i], np.uint8, cap)[dst_begin[i] : dst_begin[i] + m] = labels[side[i]][begin[i] : end[i]]
    return 0


def _fake_memcpy_async(dst, src, nbytes, stream):
    """``mtb_memcpy_async``: thread-ranks share one address space, so a peer mapping is just a pointer."""
    ctypes.memmove(dst.value, src.value, nbytes.value)
    return 0


@pytest.mark.parametrize("exchange", ["objects", "device"])
def test_finish_subtrees_on_thread_ranks_with_emulated_kernels(monkeypatch, exchange):
    """Three ranks, row-sharded: owners by size, rows written into the owners' receive planes, subtrees
    finished there, records exchanged (as objects when small, device to device through a peer-mapped
    region when large), every rank stitches the same tree as a single level-wise build."""
    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.ops import cuda_engine as ce
    from mpitree_b200.ops import subtree_phase as sp
    from mpitree_b200.parallel import symm
    from mpitree_b200.parallel.comm import run_threaded

    rng = np.random.default_rng(9)
    n, F, C, B, world = 2400, 18, 3, 16, 3
    codes = rng.integers(0, B, size=(n, F)).astype(np.uint8)
    s = codes.astype(float) @ rng.normal(size=F) + rng.normal(scale=12, size=n)
    y = np.digitize(s, np.quantile(s, [0.33, 0.66]))
    n_bins, params = np.full(F, B), SplitParams(max_depth=None)
    want = build_tree_cpu(codes, y, n_bins, C, params)
    cuts = [0, 500, 1700, 2400]

    lib = _FakeLib(n_bins, "entropy")
    lib.mtb_gather_rows = _fake_gather_rows
    lib.mtb_memcpy_async = _fake_memcpy_async
    monkeypatch.setattr(L, "load", lambda: lib)
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    monkeypatch.setattr(ce, "_stream_ptr", lambda: ctypes.c_void_p(0))
    monkeypatch.setattr(symm, "IpcRegion", _ThreadRegion)
    monkeypatch.setattr(sp, "DEVICE_EXCHANGE_MIN_BYTES", 0 if exchange == "device" else 1 << 40)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)

    def work(comm):
        lo, hi = cuts[comm.rank], cuts[comm.rank + 1]
        pending: dict = {}
        trunk = build_tree_cpu(codes[lo:hi], y[lo:hi], n_bins, C, params, comm, handoff_rows=100, pending_out=pending)
        node_ids = sorted(pending)
        G, n_cap = (F + 15) // 16, 2048
        planes = [torch.zeros(G * n_cap * 16, dtype=torch.uint8) for _ in range(2)]
        labels = [torch.zeros(n_cap, dtype=torch.uint8) for _ in range(2)]
        ho = {k: torch.zeros(len(node_ids) + 2, dtype=torch.int32) for k in ("node", "begin", "end", "side")}
        cursor = 0
        for j, nid in enumerate(node_ids[::-1]):  # listed in another order than the node ids
            rows = pending[nid] + lo
            side = (j + comm.rank) % 2
            view = planes[side].numpy().reshape(G, n_cap, 16)
            for f in range(F):
                view[f // 16, cursor : cursor + len(rows), f % 16] = codes[rows, f]
            labels[side].numpy()[cursor : cursor + len(rows)] = y[rows]
            ho["node"][j], ho["begin"][j], ho["end"][j], ho["side"][j] = int(nid), cursor, cursor + len(rows), side
            cursor += len(rows)
        host_scal = torch.zeros(8, dtype=torch.int32)
        host_scal[5] = len(node_ids)
        ctx = types.SimpleNamespace(max_depth=-1, min_samples_split=2, criterion=0, dot_mode=0, plogp=0, plogp_n=0, klogk=0,
                                    klogk_n=0)
        data = types.SimpleNamespace(device=torch.device("cpu"), G=G, n_cap=n_cap, codes=planes, labels=labels)
        builder = types.SimpleNamespace(_host_scal=host_scal, ho=ho, ho_cap=len(node_ids) + 2, data=data, comm=comm, C=C, F=F,
                                        G=G, B=B, small_rows=100, ctx=ctx, stats=types.SimpleNamespace(handed_off=0))
        got = sp.finish_subtrees(builder, trunk)
        return [getattr(got, k).tolist() for k in ("feature", "bin", "left", "right", "depth", "counts")]

    out = run_threaded(world, work)
    expect = [getattr(want, k).tolist() for k in ("feature", "bin", "left", "right", "depth", "counts")]
    assert out[0] == out[1] == out[2] == expect
