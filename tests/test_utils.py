"""Utilities: checkpoint format, synthetic data, device mapper (CPU tensors), stitch_arrays."""
import io

import numpy as np
import pytest

from mpitree_b200.tree import DecisionTreeClassifier
from mpitree_b200.utils.checkpoint import load_model, save_model
from mpitree_b200.utils.data import notebook_workload, synthetic_codes, synthetic_continuous, synthetic_mixed


def test_checkpoint_roundtrip(iris, tmp_path):
    clf = DecisionTreeClassifier(max_depth=4, device="cpu").fit(iris.data, iris.target)
    p = tmp_path / "model.npz"
    save_model(clf, str(p))
    again = load_model(str(p))
    assert again.export_text() == clf.export_text()
    assert np.array_equal(again.predict(iris.data), clf.predict(iris.data))
    assert again.get_params()["max_depth"] == 4
    buf = io.BytesIO()
    save_model(clf, buf)
    buf.seek(0)
    assert load_model(buf).score(iris.data, iris.target) == clf.score(iris.data, iris.target)


def test_synthetic_generators_are_seeded():
    X1, y1 = synthetic_codes(2000, 16, 4, seed=3)
    X2, y2 = synthetic_codes(2000, 16, 4, seed=3)
    assert X1.dtype == np.uint8 and np.array_equal(X1, X2) and np.array_equal(y1, y2)
    assert set(np.unique(y1)) == {0, 1, 2, 3}
    Xc, yc = synthetic_continuous(1000, 8)
    assert Xc.dtype == np.float32 and set(np.unique(yc)) == {0, 1}
    Xm, ym = synthetic_mixed(1000, 4, 40)
    assert Xm.shape == (1000, 44) and ym.max() <= 4
    clf = DecisionTreeClassifier(max_depth=6, device="cpu").fit(X1, y1)
    assert clf.score(X1, y1) > 0.4  # structure, not noise
    Xn, yn = notebook_workload(11)
    assert DecisionTreeClassifier(device="cpu").fit(Xn, yn).score(Xn, yn) == 1.0


def test_device_mapper_matches_host_mapper_on_cpu_tensors():
    import torch

    from mpitree_b200.ops.binning import BinMapper
    from mpitree_b200.ops.cuda_engine import _fit_mapper_device

    rng = np.random.default_rng(0)
    X = rng.normal(size=(4000, 5)).astype(np.float32)
    X[:, 1] = rng.integers(0, 9, 4000)
    X[:, 2] = np.round(X[:, 2], 1)
    for bins in (16, 64, 256):
        got = _fit_mapper_device(torch.from_numpy(X), bins)
        want = BinMapper.fit(X, bins)
        assert all(np.array_equal(a, b) for a, b in zip(got.edges, want.edges))


def test_stitch_arrays_equals_full_tree(iris):
    from mpitree_b200.ops.subtree_phase import stitch_arrays

    X, y = iris.data, iris.target
    full = DecisionTreeClassifier(device="cpu").fit(X, y).flat_tree_
    trunk = DecisionTreeClassifier(max_depth=2, device="cpu").fit(X, y).flat_tree_
    leaf_ids = trunk.apply_raw(X)
    roots, offs, nn, parts = [], [], [], {k: [] for k in ("feature", "bin", "left", "gain", "counts")}
    off = 0
    for leaf in np.unique(leaf_ids):
        if trunk.depth[leaf] < 2 or (trunk.counts[leaf] > 0).sum() <= 1:
            continue
        rows = leaf_ids == leaf
        sub_clf = DecisionTreeClassifier(device="cpu")
        sub_clf.classes_ = np.arange(3)
        sub = sub_clf._fit_flat(X[rows], y[rows], local=True)
        roots.append(leaf); offs.append(off); nn.append(sub.n_nodes)
        for k in parts:
            parts[k].append(getattr(sub, k))
        off += sub.n_nodes
    merged = stitch_arrays(trunk, np.array(roots), np.array(offs), np.array(nn), {k: np.concatenate(v) for k, v in parts.items()})
    assert merged.n_nodes == full.n_nodes
    for k in ("feature", "counts", "left", "right", "depth", "parent"):
        assert np.array_equal(getattr(merged, k), getattr(full, k)), k


def test_flat_tree_apply_codes_and_object_columns():
    from mpitree_b200.ops.binning import BinMapper

    X = np.array([["b", 3.0], ["a", 1.0], ["c", 2.0], ["a", 3.0], ["c", 1.0], ["b", 2.0]], dtype=object)
    y = np.array([1, 0, 2, 0, 2, 1])
    clf = DecisionTreeClassifier(device="cpu").fit(X, y)
    assert clf.engine_ == "cpu" and clf.predict(X).tolist() == y.tolist()
    mapper = clf.bin_mapper_
    assert mapper.edges[0].tolist() == ["a", "b", "c"]  # ordinal codes in sorted order
    codes = mapper.transform(X)
    assert np.array_equal(clf.flat_tree_.apply_codes(codes), clf.apply(X))
    # unseen category falls into the bin of the next larger known value
    assert BinMapper.transform(mapper, np.array([["bb", 2.0]], dtype=object))[0, 0] == 2


def test_assign_owners_edge_cases():
    from mpitree_b200.parallel.handoff import assign_owners, default_handoff_rows

    assert assign_owners([], 4) == []
    assert assign_owners([5], 1) == [0]
    assert sorted(assign_owners([1, 1, 1, 1], 4)) == [0, 1, 2, 3]
    assert default_handoff_rows(64, 256, 10, 1) == 0 and default_handoff_rows(64, 256, 10, 8) == 10240


def test_subtree_record_compaction_keeps_used_prefixes():
    """Phase 2 copies only the records a subtree kernel wrote (ops/subtree_phase.py::_compact_records)."""
    import torch

    from mpitree_b200.ops.subtree_phase import _compact_records

    rng = np.random.default_rng(0)
    rows = rng.integers(1, 50, size=20)
    cap = 2 * rows - 1  # worst-case records per item
    offs = np.concatenate([[0], np.cumsum(cap)[:-1]])
    total, C = int(cap.sum()), 5
    nn = np.array([rng.integers(1, c + 1) for c in cap])
    nn[3] = cap[3]
    f = torch.arange(total, dtype=torch.int32)
    cnt = torch.arange(total * C, dtype=torch.int32).view(total, C)
    out = _compact_records(torch.from_numpy(nn.astype(np.int32)), offs, f, f + 1, f + 2, f.double(), cnt)
    want = np.concatenate([np.arange(o, o + k) for o, k in zip(offs, nn)])
    assert np.array_equal(out["feature"], want) and np.array_equal(out["bin"], want + 1)
    assert np.array_equal(out["left"], want + 2) and np.array_equal(out["gain"], want.astype(float))
    assert out["counts"].dtype == np.uint32 and np.array_equal(out["counts"][:, 0], want * C)
    assert np.array_equal(out["offs"], np.concatenate([[0], np.cumsum(nn)[:-1]]))
    assert np.array_equal(out["nn"], nn)


def test_ctypes_mirror_matches_the_c_header(tmp_path):
    """``MtbCtx`` is filled from Python: every field of the mirror must sit at the C struct's offset."""
    import ctypes
    import os
    import shutil
    import subprocess

    from mpitree_b200.ops import _lib as L

    cxx = shutil.which("g++") or shutil.which("c++")
    if cxx is None:
        pytest.skip("no C++ compiler")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # codes0/codes1 etc. in the mirror are the elements of the C arrays codes[2] etc.
    c_name = {"codes0": "codes[0]", "codes1": "codes[1]", "labels0": "labels[0]", "labels1": "labels[1]",
              "hist0": "hist[0]", "hist1": "hist[1]"}
    names = [f[0] for f in L.MtbCtx._fields_]
    prints = "".join(f'  std::printf("%zu\\n", offsetof(MtbCtx, {c_name.get(n, n)}));\n' for n in names)
    src = tmp_path / "layout.cpp"
    src.write_text('#include "mpitree_b200.h"\n#include <cstdio>\n#include <cstddef>\n'
                   f'int main() {{\n  std::printf("%zu\\n", sizeof(MtbCtx));\n{prints}  return 0;\n}}\n')
    exe = tmp_path / "layout"
    subprocess.run([cxx, "-Wno-invalid-offsetof", "-I", os.path.join(root, "include"), str(src), "-o", str(exe)], check=True)
    out = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert ctypes.sizeof(L.MtbCtx) == out[0]
    assert [getattr(L.MtbCtx, n).offset for n in names] == out[1:]


def test_tree_builder_host_setup_runs_without_a_gpu():
    """The CUDA engine's host-side set-up (sizing, context struct, buffer binding) on CPU tensors.

    No kernel is launched; this catches Python-level mistakes in the GPU driver code on the CPU box.
    """
    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.ops import cuda_engine as ce
    from mpitree_b200.ops.cpu_engine import SplitParams

    try:
        L.load()
    except Exception as exc:  # noqa: BLE001  (no nvcc and no prebuilt library)
        pytest.skip(f"native library unavailable: {exc}")

    class Data:
        device = torch.device("cpu")
        n_features, n_classes, G, n_cap, n_local, cur = 64, 10, 4, 4096, 4000, 0
        codes = [torch.zeros(4 * 4096 * 16, dtype=torch.uint8) for _ in range(2)]
        labels = [torch.zeros(4096, dtype=torch.uint8) for _ in range(2)]

    for kw, layout in ((dict(), 1), (dict(tile_layout="planar"), 0), (dict(fused=False), 1)):
        b = ce.TreeBuilder(Data(), 256, SplitParams(max_depth=5), hist_budget_bytes=64 << 20, **kw)
        b._bind()
        assert b.stats.FT == 16 and b.ctx.tile_layout == layout and not b.nvls
        assert list(b.ctx.mc_hist) == [0, 0] and b.ctx.world == 1
        assert b.h_cap == 16
    with pytest.raises(ValueError):
        ce.TreeBuilder(Data(), 256, SplitParams(max_depth=5), hist_budget_bytes=64 << 20, tile_layout="diagonal")


def test_interleaved_tile_index_arithmetic_host_emulation():
    """Host emulation of ``tile_count_il`` + the interleaved ``flush_tile`` (csrc/build.cu).

    Mirrors the kernel's integer arithmetic step by step: the per-lane byte permutation (word swaps +
    ``__byte_perm``) of the exclusive-or feature order, the address formed by one ``LOP3`` (row term ^ 4k) and
    one byte dot product (``__dp4a`` with the multiplier 64 in byte k & 3), the ``[cell = class*B + bin][16]``
    tile, and the flush's lane -> (cell, feature quad) mapping.  Checks that every (feature, class, bin) count
    lands in the ``[f][C][B]`` slab exactly once.  It pins the index algebra, not the CUDA code itself (that is
    ``tests/test_gpu_variants.py``).
    """

    def byte_perm(a, b, sel):
        by = [(a >> (8 * i)) & 255 for i in range(4)] + [(b >> (8 * i)) & 255 for i in range(4)]
        return sum(by[(sel >> (4 * n)) & 7] << (8 * n) for n in range(4))

    def dp4a(a, b, c):
        return (sum(((a >> (8 * i)) & 255) * ((b >> (8 * i)) & 255) for i in range(4)) + c) & 0xFFFFFFFF

    rng = np.random.default_rng(0)
    B, C = 256, 10
    BC = B * C
    tile_base = 0x400  # shared-window byte address of the tile: 128-byte aligned
    tile = np.zeros(BC * 16, np.int64)
    want = np.zeros((16, C, B), np.int64)
    for row in range(1500):
        lane = row % 32
        codes = rng.integers(0, 256, 16)
        lab = int(rng.integers(0, C))
        w = [int(sum(int(codes[4 * i + j]) << (8 * j) for j in range(4))) for i in range(4)]
        r = lane & 15
        if r & 4:
            w = [w[1], w[0], w[3], w[2]]
        if r & 8:
            w = [w[2], w[3], w[0], w[1]]
        sel = 0x3210 ^ (0x1111 * (r & 3))
        y = [byte_perm(w[i], 0, sel) for i in range(4)]
        row_base = tile_base + lab * B * 64 + 4 * r
        for k in range(16):
            addr = dp4a(y[k >> 2], 0x40 << (8 * (k & 3)), row_base ^ (4 * k))
            word, rem = divmod(addr - tile_base, 4)
            assert rem == 0
            cell, f = divmod(word, 16)
            assert f == k ^ r and cell == lab * B + int(codes[f])
            tile[word] += 1
        want[np.arange(16), lab, codes] += 1
    slab = np.zeros(16 * BC, np.int64)
    hits = np.zeros(16 * BC, np.int64)
    for warp in range(16):
        for c0 in range(warp * 32, BC, 16 * 32):
            for lane in range(32):
                q, r = lane & 3, lane >> 2
                for i in range(4):
                    cell = c0 + r + 8 * i
                    if cell < BC:
                        for j in range(4):
                            slab[(4 * q + j) * BC + cell] += tile[cell * 16 + 4 * q + j]
                            hits[(4 * q + j) * BC + cell] += 1
    assert (hits == 1).all()
    assert np.array_equal(slab.reshape(16, C, B), want)
    # bank view: in one instruction (fixed k) the 32 lanes touch 16 features, two lanes each,
    # and feature f only ever maps to banks f and 16 + f
    for k in range(16):
        feats = [k ^ (lane & 15) for lane in range(32)]
        assert sorted(feats) == sorted(list(range(16)) * 2)
    cells = rng.integers(0, BC, 1000)
    for f in range(16):
        assert set(((cells * 16 + f) % 32).tolist()) <= {f, 16 + f}


def test_quad_encoder_search_host_emulation():
    """Host emulation of ``k_encode_f32_quads`` (csrc/encode.cu, staged path): a fixed 8-step branch-free
    lower bound in a float32 copy of the edge table (edges converted with round-toward-minus-infinity,
    padded with +inf), clamped to the last real bin, must give the codes of ``BinMapper.transform``."""
    from mpitree_b200.ops.binning import BinMapper

    rng = np.random.default_rng(0)
    X = rng.normal(size=(20000, 3)).astype(np.float32)
    X[:, 2] = rng.integers(0, 9, size=20000)
    mapper = BinMapper.fit(X[:5000].astype(np.float64) * (1 + 1e-9), max_bins=200)  # float64 edges, not float32 values
    edges64 = mapper.edges_matrix(np.float64)  # [F][256], +inf padded
    e32 = edges64.astype(np.float32)
    too_big = e32.astype(np.float64) > edges64  # round-to-nearest went up: step down to the float32 below
    e32[too_big] = np.nextafter(e32[too_big], np.float32(-np.inf))
    assert (e32.astype(np.float64) <= edges64).all()
    Xs = X.copy()
    Xs[::101, 0], Xs[::103, 1], Xs[::107, 0] = np.inf, -np.inf, np.nan
    for f in range(3):
        x = Xs[:, f]
        lo = np.zeros(len(x), np.int64)
        step = 128
        while step:
            with np.errstate(invalid="ignore"):
                lo += np.where(~(e32[f][lo + step - 1] >= x), step, 0)
            step >>= 1
        got = np.minimum(lo, mapper.n_bins[f] - 1)
        finite = np.isfinite(x)
        want = mapper.transform(Xs[finite].astype(np.float64))[:, f]
        assert np.array_equal(got[finite], want)
        assert (got[np.isnan(x)] == mapper.n_bins[f] - 1).all() and (got[x == -np.inf] == 0).all()
        assert (got[x == np.inf] == mapper.n_bins[f] - 1).all()


def test_level_loop_dry_run_launch_sequences(monkeypatch):
    """The CUDA engine's level loop with every native call replaced by a recorder (CPU tensors, no GPU).

    Exercises set-up, per-level binding, side swapping and the read-back code of each mode and pins the
    launch sequence a fit enqueues: one cooperative launch per level on the fused path.
    """
    import ctypes

    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.ops import cuda_engine as ce
    from mpitree_b200.ops.cpu_engine import SplitParams

    try:
        L.load()
    except Exception as exc:  # noqa: BLE001
        pytest.skip(f"native library unavailable: {exc}")
    calls = []

    class Stream:
        def synchronize(self):
            pass

    monkeypatch.setattr(ce, "_call", lambda name, *a: calls.append(name))
    monkeypatch.setattr(ce, "_stream_ptr", lambda: ctypes.c_void_p(0))
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: Stream())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self, *a, **k: self)

    class Data:
        device = torch.device("cpu")
        n_features, n_classes, G, n_cap, n_local, cur = 64, 10, 4, 4096, 4000, 0
        codes = [torch.zeros(4 * 4096 * 16, dtype=torch.uint8) for _ in range(2)]
        labels = [torch.zeros(4096, dtype=torch.uint8) for _ in range(2)]

    def run(**kw):
        calls.clear()
        b = ce.TreeBuilder(Data(), 256, SplitParams(max_depth=4), hist_budget_bytes=64 << 20, **kw)
        monkeypatch.setattr(b, "_launch", lambda name: calls.append(name))
        flat = b.build()
        assert flat.n_nodes == 0  # nothing ran: the level-0 kernel (not the host) writes the root
        return b.async_levels, list(calls)

    assert run() == (True, ["mtb_level_fused"] + ["mtb_level_fused2"] * 3)
    assert run(tile_layout="interleaved")[1] == ["mtb_level_fused"] + ["mtb_level_fused2"] * 3
    assert run(fuse_partition=False)[1] == ["mtb_level_fused", "mtb_level_partition", "mtb_level_commit"] * 4
    unfused = ["mtb_level_plan", "mtb_level_hist", "mtb_level_gain", "mtb_level_finalize", "mtb_level_partition",
               "mtb_level_commit"]
    assert run(fused=False)[1] == unfused * 4
    sync, seq = run(small_rows=64)  # hand-off on: level-by-level with a read-back, stops when the frontier is empty
    assert not sync and seq == ["mtb_level_fused"]


def test_multi_rank_builder_wiring_dry_run(monkeypatch):
    """Peer (``p2p``) and in-switch (``nvls``, staged) set-up of the CUDA engine with a fake communicator
    and fake symmetric regions: which buffers the context points at, and that the level loop is the same."""
    import ctypes

    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.ops import cuda_engine as ce
    from mpitree_b200.ops.cpu_engine import SplitParams

    try:
        L.load()
    except Exception as exc:  # noqa: BLE001
        pytest.skip(f"native library unavailable: {exc}")
    calls, regions = [], []

    class Stream:
        def synchronize(self):
            pass

    class Comm:
        size, rank = 4, 1

        def allgather_object(self, obj):
            return [obj] * 4

        def barrier(self):
            pass

    class Region:
        def __init__(self, nbytes):
            self.nbytes, self.mem = nbytes, torch.zeros(nbytes, dtype=torch.uint8)
            self.mc_ptr, self.ptrs = 0xABC000, [self.mem.data_ptr()] * 4

        def as_tensor(self):
            return self.mem

    monkeypatch.setattr(ce, "_call", lambda name, *a: calls.append(name))
    monkeypatch.setattr(ce, "_stream_ptr", lambda: ctypes.c_void_p(0))
    monkeypatch.setattr(ce, "_nvls_region", lambda n, c: regions.append(Region(n)) or regions[-1])
    monkeypatch.setattr(ce, "_shared_region", lambda n, c: Region(n))
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: Stream())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "mem_get_info", lambda *a, **k: (8 << 30, 16 << 30))
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self, *a, **k: self)

    class Data:
        device = torch.device("cpu")
        n_features, n_classes, G, n_cap, n_local, cur = 64, 10, 4, 4096, 4000, 0
        codes = [torch.zeros(4 * 4096 * 16, dtype=torch.uint8) for _ in range(2)]
        labels = [torch.zeros(4096, dtype=torch.uint8) for _ in range(2)]

    for reduce in ("p2p", "nvls"):
        calls.clear()
        b = ce.TreeBuilder(Data(), 256, SplitParams(max_depth=4), comm=Comm(), reduce=reduce, n_global=16000)
        monkeypatch.setattr(b, "_launch", lambda name: calls.append(name))
        b.build()
        assert calls == ["mtb_level_fused"] + ["mtb_level_fused2"] * 3
        assert (b.ctx.world, b.ctx.rank, b.ctx.fpo, b.stats.reduce) == (4, 1, 16, reduce)
        if reduce == "nvls":  # the slabs live in the multicast-bound region and the kernels get its alias
            assert b.nvls and list(b.ctx.mc_hist) == [0xABC000] * 2
            assert b.hist[0].data_ptr() == regions[-1].mem.data_ptr() and b.ctx.hist0 == b.ctx.hist1
            assert regions[-1].nbytes == b.h_cap * b.slab_words * 4
        else:
            assert not b.nvls and list(b.ctx.mc_hist) == [0, 0]


def test_sampled_edges_are_repaired_when_the_sample_missed_a_value():
    """Edges fitted on a strided sample: a rare value of a low-cardinality column that the sample did not see
    must get its own bin (otherwise it silently shares the next edge's bin and the search is no longer the
    exact search the documentation promises for columns with at most ``max_bins`` distinct values)."""
    import torch

    from mpitree_b200.ops import cuda_engine as ce
    from mpitree_b200.ops.binning import BinMapper

    rng = np.random.default_rng(0)
    n = ce.SAMPLE_ROWS * 2 + 11
    X = rng.normal(size=(n, 3)).astype(np.float32)
    X[:, 1] = rng.integers(0, 5, size=n)
    idx = ce._sample_rows(n)
    missed = int(np.setdiff1d(np.arange(n), idx)[1000])
    X[missed, 1] = 77.0
    X[:, 2] = rng.integers(0, 1000, size=n)  # more distinct values than bins: stays quantile-binned
    mapper = BinMapper.fit(X[idx], max_bins=256)
    assert list(mapper.edges[1]) == [0, 1, 2, 3, 4]
    changed = ce._repair_sampled_edges(torch.from_numpy(X), mapper, 256)
    assert changed == 1 and list(mapper.edges[1]) == [0, 1, 2, 3, 4, 77] and mapper.lossless[1]
    assert len(mapper.edges[2]) == 256 and not mapper.lossless[2]
    # the whole front end: the host mapper of a large input finds the stray value too
    m2 = ce._fit_mapper_numpy(X, 256)
    assert 77.0 in m2.edges[1]
    # thread ranks: only one shard holds the stray value, every rank must end with the same edges
    from mpitree_b200.parallel.comm import run_threaded

    cuts = [0, n // 3, n]

    def work(comm):
        lo, hi = cuts[comm.rank], cuts[comm.rank + 1]
        return [e.tolist() for e in ce._fit_mapper_numpy(X[lo:hi], 256, comm).edges]

    got = run_threaded(2, work)
    assert got[0] == got[1] and 77.0 in got[0][1]


def test_root_mode_switch(monkeypatch):
    """``MTB_ROOT`` picks how the first level reads uint8 code input; the older ``MTB_ROOT_CLUSTER=1`` still means
    the cluster pass; anything unknown falls back to the default."""
    from mpitree_b200.ops import cuda_engine

    monkeypatch.delenv("MTB_ROOT", raising=False)
    monkeypatch.delenv("MTB_ROOT_CLUSTER", raising=False)
    assert cuda_engine._root_mode() == "direct"
    for mode in ("direct", "cluster", "off"):
        monkeypatch.setenv("MTB_ROOT", mode)
        assert cuda_engine._root_mode() == mode
    monkeypatch.setenv("MTB_ROOT", "nonsense")
    assert cuda_engine._root_mode() == "direct"
    monkeypatch.setenv("MTB_ROOT", "off")
    monkeypatch.setenv("MTB_ROOT_CLUSTER", "1")
    assert cuda_engine._root_mode() == "cluster"


def test_root_direct_row_ranges_cover_every_row_once():
    """Row assignment of ``k_root_direct`` (csrc/build.cu): CTA b takes feature group b mod G of row range b / G;
    a range is a run of 32-row blocks dealt to the CTA's warps four blocks per pass.  Every (row, group) pair
    must be visited exactly once for any n and G."""
    nw, unroll = 32, 4
    for n, G, sms in ((999, 4, 148), (70_001, 3, 148), (5, 1, 148), (300_000, 8, 148), (4097, 4, 6)):
        n_ranges = sms // G
        seen = np.zeros((n, G), np.int64)
        blocks = (n + 31) // 32
        for b in range(n_ranges * G):
            g, rng_ = b % G, b // G
            b0, b1 = blocks * rng_ // n_ranges, blocks * (rng_ + 1) // n_ranges
            for w in range(nw):
                blk = b0 + w
                while blk < b1:
                    for j in range(unroll):
                        bj = blk + j * nw
                        if bj < b1:
                            lo, hi = bj * 32, min(bj * 32 + 32, n)
                            if lo < hi:
                                seen[lo:hi, g] += 1
                    blk += nw * unroll
        assert (seen == 1).all(), (n, G)


def test_cpulist_parsing_and_binding_is_a_no_op_without_topology():
    from mpitree_b200.utils import topology

    assert topology.parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert topology.parse_cpulist("") == []
    assert topology.parse_cpulist("5") == [5]
    # no CUDA device here: neither source reports anything and the process keeps its CPU set
    import os

    before = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
    assert topology.gpu_cpu_affinity_sysfs(0) is None
    assert topology.bind_to_gpu_numa(0) is None
    if before is not None:
        assert os.sched_getaffinity(0) == before
