"""``mtb_predict_noalloc``: the prediction entry point for callers that cannot allocate.

Some users of the C library cannot call ``cudaMalloc`` at prediction time: a serving process
that split the whole GPU between its components at start-up, or a plugin that runs inside
another application's allocator and only receives input/output pointers.  For them the
library keeps the node-record scratch in a fixed 1 MiB device array of its own.  This test
drives that entry point through the C ABI (ctypes) and checks that it returns exactly what
the regular ``mtb_pack_nodes`` + ``mtb_predict`` path returns, on two calls in a row that use
different trees and input types (the shared array still holds the first call's records when
the second call starts).  The file sorts last so a ``-x`` run has already run everything else.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _tree_arrays(flat, torch, dev, dtype_code):
    thr = np.where(np.isnan(flat.numeric_thresholds()), 0.0, flat.numeric_thresholds())
    if dtype_code == 0:
        t32 = thr.astype(np.float32)
        t32 = np.where(t32.astype(np.float64) > thr, np.nextafter(t32, np.float32(-np.inf)), t32)
        thr_t = torch.from_numpy(t32)
    elif dtype_code == 1:
        thr_t = torch.from_numpy(np.floor(thr).astype(np.int32))
    else:
        thr_t = torch.from_numpy(thr)
    labels = flat.expanded_counts().argmax(axis=1).astype(np.int32)
    return (
        torch.from_numpy(flat.feature.astype(np.int32)).to(dev),
        thr_t.to(dev),
        torch.from_numpy(flat.left.astype(np.int32)).to(dev),
        torch.from_numpy(labels).to(dev),
    )


def _both_paths(lib, torch, dev, X, flat, dtype_code):
    vp = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    feature, thr, left, label = _tree_arrays(flat, torch, dev, dtype_code)
    n, F, nn = X.shape[0], X.shape[1], flat.n_nodes
    # existing path: caller-allocated scratch
    packed = torch.empty(nn * 4, dtype=torch.int32, device=dev)
    leaf_a = torch.full((n,), -7, dtype=torch.int32, device=dev)
    lab_a = torch.full((n,), -7, dtype=torch.int32, device=dev)
    assert lib.mtb_pack_nodes(vp(feature), vp(left), vp(label), ctypes.c_int64(nn), vp(packed), stream) == 0
    assert lib.mtb_predict(vp(X), ctypes.c_int(dtype_code), ctypes.c_int64(n), ctypes.c_int64(F),
                           ctypes.c_int64(X.stride(0)), vp(packed), vp(thr), ctypes.c_int64(nn), vp(leaf_a), vp(lab_a),
                           stream) == 0
    # new path: no allocation inside the library
    leaf_b = torch.full((n,), -9, dtype=torch.int32, device=dev)
    lab_b = torch.full((n,), -9, dtype=torch.int32, device=dev)
    rc = lib.mtb_predict_noalloc(vp(X), ctypes.c_int(dtype_code), ctypes.c_int64(n), ctypes.c_int64(F),
                                 ctypes.c_int64(X.stride(0)), vp(feature), vp(thr), vp(left), vp(label),
                                 ctypes.c_int64(nn), vp(leaf_b), vp(lab_b), stream)
    assert rc == 0, lib.mtb_last_error()
    torch.cuda.synchronize()
    return leaf_a.cpu().numpy(), lab_a.cpu().numpy(), leaf_b.cpu().numpy(), lab_b.cpu().numpy()


def test_noalloc_predict_matches_regular_path_twice():
    import torch

    from mpitree_b200.ops import _lib
    from mpitree_b200.tree import DecisionTreeClassifier

    lib = _lib.load()
    lib.mtb_last_error.restype = ctypes.c_char_p
    dev = torch.device("cuda", torch.cuda.current_device())
    rng = np.random.default_rng(42)

    # call 1: a deep tree on float32 features
    X1 = rng.normal(size=(60000, 12)).astype(np.float32)
    y1 = ((X1[:, 0] > 0.1).astype(np.int64) + (X1[:, 3] * X1[:, 5] > 0) + (X1[:, 7] > 1.0))
    big = DecisionTreeClassifier(max_depth=9, device="cuda").fit(X1, y1).flat_tree_
    a_leaf, a_lab, b_leaf, b_lab = _both_paths(lib, torch, dev, torch.from_numpy(X1).to(dev), big, 0)
    assert np.array_equal(a_leaf, b_leaf) and np.array_equal(a_lab, b_lab)
    assert np.array_equal(a_leaf, big.apply_raw(X1))

    # call 2, right after: a much smaller tree on uint8 features; stale records from call 1
    # are still in the library's array beyond this tree's node count and must not matter
    X2 = rng.integers(0, 256, size=(30000, 20), dtype=np.uint8)
    y2 = (X2[:, 2] > 99).astype(np.int64) * 2 + (X2[:, 11] > 180)
    small = DecisionTreeClassifier(max_depth=3, device="cuda").fit(X2, y2).flat_tree_
    assert small.n_nodes < big.n_nodes
    a_leaf, a_lab, b_leaf, b_lab = _both_paths(lib, torch, dev, torch.from_numpy(X2).to(dev), small, 1)
    assert np.array_equal(a_leaf, b_leaf) and np.array_equal(a_lab, b_lab)
    assert np.array_equal(a_leaf, small.apply_raw(X2))


def test_noalloc_predict_rejects_oversized_tree():
    import torch

    from mpitree_b200.ops import _lib

    lib = _lib.load()
    lib.mtb_last_error.restype = ctypes.c_char_p
    dev = torch.device("cuda", torch.cuda.current_device())
    x = torch.zeros((4, 4), dtype=torch.float32, device=dev)
    z = torch.zeros(8, dtype=torch.int32, device=dev)
    vp = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    rc = lib.mtb_predict_noalloc(vp(x), ctypes.c_int(0), ctypes.c_int64(4), ctypes.c_int64(4), ctypes.c_int64(4),
                                 vp(z), vp(x), vp(z), vp(z), ctypes.c_int64(65537), vp(z), vp(z), ctypes.c_void_p(0))
    assert rc != 0
    assert b"65536" in lib.mtb_last_error()
