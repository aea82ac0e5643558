"""CUDA engine vs the host engine / fp32-64 references (run on the B200 box: ``-m gpu``)."""
import numpy as np
import pytest

from mpitree_b200.ops.binning import BinMapper
from mpitree_b200.ops.cpu_engine import SplitParams
from mpitree_b200.tree import DecisionTreeClassifier

from .goldens import IRIS2_DEPTH3, IRIS2_DEPTH5_P1

pytestmark = pytest.mark.gpu


def _same_tree(a, b):
    assert a.export_text() == b.export_text()
    fa, fb = a.flat_tree_, b.flat_tree_
    assert np.array_equal(fa.counts, fb.counts)
    assert np.array_equal(fa.feature, fb.feature) and np.array_equal(fa.bin, fb.bin)
    assert np.array_equal(fa.left, fb.left) and np.array_equal(fa.right, fb.right)
    assert np.allclose(np.nan_to_num(fa.gain, nan=-1.0), np.nan_to_num(fb.gain, nan=-1.0), rtol=0, atol=1e-12)


def test_native_library_loaded():
    from mpitree_b200.ops import _lib

    lib = _lib.load()
    assert lib.mtb_version() == 100
    assert lib.mtb_sm_count(0) >= 100


def test_encode_float_matches_host():
    import torch

    from mpitree_b200.ops.cuda_engine import DeviceDataset

    rng = np.random.default_rng(0)
    for dtype, F in ((np.float32, 5), (np.float64, 37)):
        X = rng.normal(size=(3000, F)).astype(dtype)
        X[:, 0] = rng.integers(0, 7, size=3000)
        mapper = BinMapper.fit(X, max_bins=64)
        want = mapper.transform(X)
        data = DeviceDataset.allocate(3000, F, 2, torch.device("cuda"))
        data.codes[0].zero_()
        data.encode_float(torch.from_numpy(X).cuda(), mapper)
        got = data.codes_rowmajor().cpu().numpy()
        assert np.array_equal(got, want)


def test_encode_u8_identity_and_lut():
    import torch

    from mpitree_b200.ops.cuda_engine import DeviceDataset

    rng = np.random.default_rng(1)
    X = rng.integers(0, 256, size=(5000, 40), dtype=np.uint8)
    data = DeviceDataset.allocate(5000, 40, 2, torch.device("cuda"))
    data.encode_u8(torch.from_numpy(X).cuda())
    assert np.array_equal(data.codes_rowmajor().cpu().numpy(), X)
    lut = np.stack([rng.permutation(256).astype(np.uint8) for _ in range(40)])
    data.encode_u8(torch.from_numpy(X).cuda(), torch.from_numpy(lut).cuda())
    want = np.take_along_axis(lut.T, X.astype(np.int64), axis=0)
    assert np.array_equal(data.codes_rowmajor().cpu().numpy(), want)


@pytest.mark.parametrize("depth", [0, 1, 3, 5, None])
@pytest.mark.parametrize("cols", [slice(0, 2), slice(0, 4)])
def test_iris_gpu_equals_cpu(iris, depth, cols):
    X, y = iris.data[:, cols], iris.target
    g = DecisionTreeClassifier(max_depth=depth, device="cuda").fit(X, y)
    c = DecisionTreeClassifier(max_depth=depth, device="cpu").fit(X, y)
    assert g.engine_ == "cuda" and c.engine_ == "cpu"
    _same_tree(g, c)
    assert np.array_equal(g.predict(X), c.predict(X))
    assert np.array_equal(g.predict_proba(X), c.predict_proba(X))


def test_goldens_on_gpu(iris):
    X, y = iris.data[:, :2], iris.target
    clf = DecisionTreeClassifier(max_depth=3, device="cuda").fit(X, y)
    assert clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names) == IRIS2_DEPTH3
    clf = DecisionTreeClassifier(max_depth=5, device="cuda").fit(X, y)
    assert clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names, precision=1) == IRIS2_DEPTH5_P1


def test_gpu_matches_reference_oracle(ref, iris):
    X, y = iris.data, iris.target
    ours = DecisionTreeClassifier(device="cuda").fit(X, y)
    theirs = ref.DecisionTreeClassifier().fit(X, y)
    assert ours.export_text() == theirs.export_text()
    assert np.array_equal(ours.predict(X), theirs.predict(X))


CASES = [
    # n, F, values, classes, depth, mss
    (500, 3, 4, 2, 6, 2),
    (2000, 16, 8, 3, 5, 2),
    (5000, 33, 256, 10, 6, 2),
    (20000, 64, 256, 10, 5, 2),
    (3000, 7, 3, 5, None, 10),
    (4096, 100, 64, 4, 4, 50),
    (1500, 20, 17, 37, 5, 2),
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("fused", [True, False])
def test_random_data_gpu_equals_cpu(case, fused):
    from mpitree_b200.ops import cuda_engine
    from mpitree_b200.parallel.scheduler import fit_flat_tree_host

    n, F, vals, k, depth, mss = case
    rng = np.random.default_rng(hash(case) % (2**32))
    X = rng.integers(0, vals, size=(n, F)).astype(np.float32)
    w = rng.normal(size=F)
    y = np.digitize(X @ w + rng.normal(scale=vals, size=n), np.quantile(X @ w, np.linspace(0, 1, k + 1)[1:-1]))
    y[:k] = np.arange(k)
    params = SplitParams(max_depth=depth, min_samples_split=mss)
    flat_g, mapper = cuda_engine.fit_from_host(X, y, k, params, builder_kwargs=dict(fused=fused))
    codes = mapper.transform(X)
    flat_c = fit_flat_tree_host(codes, y, mapper.n_bins, k, params)
    assert flat_g.n_nodes == flat_c.n_nodes
    assert np.array_equal(flat_g.feature, flat_c.feature)
    assert np.array_equal(flat_g.bin, flat_c.bin)
    assert np.array_equal(flat_g.counts, flat_c.counts)
    assert np.array_equal(flat_g.left, flat_c.left)
    assert np.array_equal(flat_g.depth, flat_c.depth)


def test_continuous_quantile_bins_gpu_equals_cpu():
    rng = np.random.default_rng(5)
    X = rng.normal(size=(30000, 12)).astype(np.float32)
    y = ((X[:, 2] > 0.3) ^ (X[:, 7] < -0.2)).astype(np.int64)
    g = DecisionTreeClassifier(max_depth=6, device="cuda").fit(X, y)
    c = DecisionTreeClassifier(max_depth=6, device="cpu").fit(X, y)
    _same_tree(g, c)
    assert g.score(X, y) > 0.97


def test_gini_gpu_equals_cpu(iris):
    g = DecisionTreeClassifier(criterion="gini", device="cuda").fit(iris.data, iris.target)
    c = DecisionTreeClassifier(criterion="gini", device="cpu").fit(iris.data, iris.target)
    _same_tree(g, c)


@pytest.mark.parametrize("dtype", [np.float32, np.float64, np.uint8])
def test_predict_kernel_matches_host_walk(dtype):
    rng = np.random.default_rng(7)
    if dtype == np.uint8:
        X = rng.integers(0, 256, size=(50000, 24), dtype=np.uint8)
    else:
        X = rng.normal(size=(50000, 24)).astype(dtype)
    y = (X[:, 1].astype(np.float64) > np.median(X[:, 1])).astype(np.int64) + (X[:, 5] > np.median(X[:, 5]))
    clf = DecisionTreeClassifier(max_depth=8, device="cuda").fit(X, y)
    flat = clf.flat_tree_
    want = flat.apply_raw(X)
    assert np.array_equal(clf.apply(X), want)
    assert np.array_equal(clf.predict(X), flat.expanded_counts().argmax(1)[want])


def test_torch_tensor_api_cuda_and_pinned():
    import torch

    rng = np.random.default_rng(11)
    X = rng.integers(0, 256, size=(40000, 32), dtype=np.uint8)
    y = ((X[:, 0] > 127).astype(np.int64) * 2 + (X[:, 9] > 60)).astype(np.int64)
    want = DecisionTreeClassifier(max_depth=5, device="cpu").fit(X, y)
    for Xt, yt in (
        (torch.from_numpy(X).cuda(), torch.from_numpy(y).cuda()),
        (torch.from_numpy(X).pin_memory(), torch.from_numpy(y).pin_memory()),
    ):
        got = DecisionTreeClassifier(max_depth=5).fit(Xt, yt)
        assert got.export_text() == want.export_text()
        pred = got.predict(Xt)
        assert isinstance(pred, torch.Tensor) and pred.is_cuda
        assert np.array_equal(pred.cpu().numpy(), want.predict(X))


def test_large_single_node_counts():
    """A level with one node and many rows exercises multi-CTA tile merging."""
    rng = np.random.default_rng(2)
    n = 1_000_000
    X = rng.integers(0, 256, size=(n, 16), dtype=np.uint8)
    y = rng.integers(0, 10, size=n)
    clf = DecisionTreeClassifier(max_depth=2, device="cuda").fit(X, y)
    flat = clf.flat_tree_
    assert np.array_equal(flat.counts[0], np.bincount(y, minlength=10))
    f, b = int(flat.feature[0]), int(flat.bin[0])
    assert np.array_equal(flat.counts[1], np.bincount(y[X[:, f] <= b], minlength=10))
    assert np.array_equal(flat.counts[2], np.bincount(y[X[:, f] > b], minlength=10))
    for node in (1, 2):
        if flat.feature[node] >= 0:
            l, r = flat.left[node], flat.right[node]
            assert np.array_equal(flat.counts[l] + flat.counts[r], flat.counts[node])


@pytest.mark.parametrize("small_rows", [16, 300, 4096])
@pytest.mark.parametrize("shape", [(5000, 7, 3, 5, None, 2), (20000, 40, 64, 10, 9, 2), (3000, 130, 256, 4, None, 20)])
def test_subtree_phase_equals_level_wise(shape, small_rows):
    """Phase 2: nodes handed to the one-CTA-per-subtree kernel give the same tree as the level-wise build."""
    from mpitree_b200.ops import cuda_engine
    from mpitree_b200.parallel.scheduler import fit_flat_tree_host

    n, F, vals, k, depth, mss = shape
    rng = np.random.default_rng(F * 7 + small_rows)
    X = rng.integers(0, vals, size=(n, F)).astype(np.float32)
    w = rng.normal(size=F)
    s = X @ w + rng.normal(scale=vals, size=n)
    y = np.digitize(s, np.quantile(s, np.linspace(0, 1, k + 1)[1:-1]))
    y[:k] = np.arange(k)
    params = SplitParams(max_depth=depth, min_samples_split=mss)
    got, mapper = cuda_engine.fit_from_host(X, y, k, params, builder_kwargs=dict(small_rows=small_rows))
    if small_rows >= mss:  # smaller nodes are leaves by min_samples_split and are never handed off
        assert cuda_engine.fit_from_host.last_stats.handed_off > 0
    want = fit_flat_tree_host(mapper.transform(X), y, mapper.n_bins, k, params)
    assert got.n_nodes == want.n_nodes
    assert np.array_equal(got.feature, want.feature) and np.array_equal(got.bin, want.bin)
    assert np.array_equal(got.counts, want.counts) and np.array_equal(got.left, want.left)
    assert np.array_equal(got.depth, want.depth) and np.array_equal(got.parent, want.parent)


def test_unbounded_depth_many_small_nodes():
    """Fully grown tree on noisy data: tens of thousands of nodes, far beyond the slab capacity."""
    rng = np.random.default_rng(9)
    n = 60000
    X = rng.integers(0, 256, size=(n, 24), dtype=np.uint8)
    y = rng.integers(0, 3, size=n)
    clf = DecisionTreeClassifier(device="cuda").fit(X, y)
    flat = clf.flat_tree_
    assert flat.n_nodes > 20000
    assert clf.score(X, y) == 1.0  # distinct rows: a fully grown tree memorises the data
    leaves = flat.feature < 0
    assert int(flat.counts[leaves].sum()) == n


def test_klogk_device_function_matches_numpy_twin():
    """g(k) = k log2 k of the gain phase (csrc/impurity.cuh::KlogK) against its numpy twin and plain numpy."""
    import ctypes

    import torch

    from mpitree_b200.ops import _lib
    from mpitree_b200.ops.cpu_engine import klogk
    from mpitree_b200.ops.cuda_engine import klogk_table

    rng = np.random.default_rng(0)
    k = np.concatenate([np.arange(0, 70000), 2 ** np.arange(32, dtype=np.int64), 2 ** np.arange(1, 33, dtype=np.int64) - 1,
                        rng.integers(0, 2**32, size=200_000)]).astype(np.uint32)
    kd = torch.from_numpy(k.view(np.int32)).cuda()
    out = torch.empty(k.shape[0], dtype=torch.float64, device="cuda")
    tab = klogk_table(torch.device("cuda", torch.cuda.current_device()))
    rc = _lib.load().mtb_klogk_eval(ctypes.c_void_p(kd.data_ptr()), ctypes.c_void_p(out.data_ptr()),
                                    ctypes.c_int64(k.shape[0]), ctypes.c_void_p(tab.data_ptr()),
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    got = out.cpu().numpy()
    want = klogk(k)
    big = k > 1
    assert np.all(got[~big] == 0.0)
    assert np.max(np.abs(got[big] - want[big]) / want[big]) < 5e-16  # fused vs separately rounded Horner steps
    ref = k[big].astype(np.float64) * np.log2(k[big].astype(np.float64))
    assert np.max(np.abs(got[big] - ref) / ref) < 1e-15


@pytest.mark.parametrize("C", [15, 16, 64, 128])
def test_many_classes_with_256_bins(C):
    """B = 256 with many classes: the gain slab of one (node, feature) pair is B x C words and must still
    fit next to (or instead of) the histogram pipeline's shared memory (fewer gain groups, smaller tiles)."""
    rng = np.random.default_rng(C)
    n, F = 40_000, 6
    X = rng.integers(0, 256, size=(n, F), dtype=np.uint8)
    y = ((X[:, 0].astype(np.int64) * 3 + X[:, 1] + rng.integers(0, 40, size=n)) * C // 1100).clip(0, C - 1)
    y[:C] = np.arange(C)  # every class occurs
    g = DecisionTreeClassifier(max_depth=5, device="cuda").fit(X, y)
    c = DecisionTreeClassifier(max_depth=5, device="cpu").fit(X, y)
    assert g.engine_ == "cuda"
    _same_tree(g, c)


def test_string_and_mixed_columns_run_on_the_gpu_and_match_the_reference(ref):
    """Object / string ("categorical") columns: ordinal codes on the host, the tree on the device; the
    reference splits such columns with ``<=`` on the sorted distinct values (decision_tree.py:73-77)."""
    rng = np.random.default_rng(3)
    n = 400
    colour = rng.choice(np.array(["red", "green", "blue", "amber"], dtype=object), size=n)
    size = rng.choice(np.array(["S", "M", "L"], dtype=object), size=n)
    x = rng.integers(0, 12, size=n)
    y = ((colour <= "blue").astype(int) + (x > 6) + (size == "M")).astype(np.int64)
    X = np.empty((n, 3), dtype=object)
    X[:, 0], X[:, 1], X[:, 2] = colour, x.astype(object), size
    ours = DecisionTreeClassifier(max_depth=4, device="cuda").fit(X, y)
    assert ours.engine_ == "cuda"
    cpu = DecisionTreeClassifier(max_depth=4, device="cpu").fit(X, y)
    theirs = ref.DecisionTreeClassifier(max_depth=4).fit(X, y)
    assert np.array_equal(ours.predict(X), theirs.predict(X))
    assert np.array_equal(ours.predict_proba(X), theirs.predict_proba(X))
    a, b = ours.flat_tree_, cpu.flat_tree_
    assert np.array_equal(a.feature, b.feature) and np.array_equal(a.counts, b.counts) and a.thresholds == b.thresholds
    # all-string matrix: thresholds are strings (the reference cannot print them; the structure must agree)
    Xs = np.stack([colour, size], axis=1)
    g = DecisionTreeClassifier(max_depth=3, device="cuda").fit(Xs, y)
    r = ref.DecisionTreeClassifier(max_depth=3).fit(Xs, y)
    assert g.engine_ == "cuda" and np.array_equal(g.predict(Xs), r.predict(Xs))


@pytest.mark.parametrize("n,vals,seed", [(1500, 3, 0), (1500, 4, 1), (3000, 3, 2), (3000, 4, 3)])
def test_large_tie_heavy_nodes_match_the_reference(ref, n, vals, seed):
    """Root nodes above the 1,024-row limit of the exact-arithmetic path take the k log2 k cost form; with
    3-4 distinct values per feature and duplicated / mirrored columns many candidates tie exactly, and the
    reference's tie-breaks (first threshold, first feature) must survive.  CPU engine and CUDA engine."""
    rng = np.random.default_rng(seed)
    F = 6
    X = rng.integers(0, vals, size=(n, F)).astype(np.float64)
    X[:, 3] = X[:, 0]                 # duplicate column: exact ties between features
    X[:, 4] = (vals - 1) - X[:, 1]    # mirrored column: the same partitions with left and right swapped
    y = ((X[:, 0] + X[:, 1] + rng.integers(0, 2, size=n)) % 3).astype(np.int64)
    theirs = ref.DecisionTreeClassifier(max_depth=3).fit(X, y)
    for device in ("cpu", "cuda"):
        ours = DecisionTreeClassifier(max_depth=3, device=device).fit(X, y)
        assert ours.engine_ == device
        assert ours.export_text() == theirs.export_text(), device
        assert np.array_equal(ours.predict(X), theirs.predict(X))


def test_wide_bins_on_the_gpu_match_the_reference(ref):
    rng = np.random.default_rng(1)
    n = 900
    X = np.round(rng.normal(size=(n, 4)), 3)
    X[:, 3] = rng.integers(0, 7, size=n)
    y = (X[:, 0] - 0.7 * X[:, 2] > 0.05).astype(np.int64) + 2 * (X[:, 3] > 3)
    theirs = ref.DecisionTreeClassifier(max_depth=4).fit(X, y)
    ours = DecisionTreeClassifier(max_depth=4, device="cuda", max_bins=2048).fit(X, y)
    assert ours.engine_ == "cuda"
    assert ours.export_text() == theirs.export_text()
    assert np.array_equal(ours.predict(X), theirs.predict(X))
