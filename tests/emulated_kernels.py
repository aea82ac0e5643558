"""Python stand-ins for the level-wise CUDA kernels (single rank), for GPU-less tests of the driver.

Each function takes the same ``MtbCtx`` the native entry point receives and performs the same
state transition on the buffers the context points at (CPU tensors in these tests): slot tables,
tree arrays, scalars, row planes, hand-off list.  What is *inside* a kernel is not imitated (no
tiles, no sibling subtraction, no atomics): a node's histogram is simply counted from its rows and
the split is found with the host engine's exact search, which the CUDA kernels are validated
against on hardware (``tests/test_gpu_engine.py``).  The contract between kernels and host driver —
which arrays mean what, in which order children are numbered, where rows land — follows
``csrc/build.cu`` (``finalize_phase``, ``k_partition``, ``commit_phase``).
"""
import ctypes

import numpy as np

from mpitree_b200.ops.cpu_engine import best_split

NONE, BUILD, DERIVE, HANDOFF = 0, 1, 2, 3
_CT = {np.int32: ctypes.c_int32, np.uint32: ctypes.c_uint32, np.uint8: ctypes.c_uint8, np.float64: ctypes.c_double}


def view(ptr, dtype, count):
    return np.ctypeslib.as_array(ctypes.cast(int(ptr), ctypes.POINTER(_CT[dtype])), shape=(int(count),))


class Ctx:
    """Typed views of everything an ``MtbCtx`` points at."""

    def __init__(self, c):
        self.c = c
        m, F, C, G, n_cap, node_cap = int(c.m_cap), int(c.F), int(c.C), int(c.G), int(c.n_cap), int(c.node_cap)
        self.F, self.C, self.G, self.B, self.n_cap = F, C, G, int(c.B), n_cap
        i32 = lambda p, k: view(p, np.int32, k)  # noqa: E731
        self.scal = i32(c.scalars, 8)
        self.slot = {k: i32(getattr(c, "slot_" + k), m) for k in ("node", "begin", "end", "mode", "parent", "sibling")}
        self.nslot = {k: i32(getattr(c, "nslot_" + k), m) for k in ("node", "begin", "end", "mode", "parent", "sibling", "src")}
        self.rec_feature, self.rec_bin, self.part_flags = i32(c.rec_feature, m), i32(c.rec_bin, m), i32(c.part_flags, m)
        self.cursors = view(c.cursors, np.uint32, 2 * m)
        self.part_child = i32(c.part_child, m) if c.part_child else None
        self.t = {k: i32(getattr(c, "t_" + k), node_cap) for k in ("feature", "bin", "left", "parent", "depth")}
        self.t_gain = view(c.t_gain, np.float64, node_cap)
        self.t_counts = view(c.t_counts, np.uint32, node_cap * C).reshape(node_cap, C)
        self.planes = [view(p, np.uint8, G * n_cap * 16).reshape(G, n_cap, 16) for p in (c.codes0, c.codes1)]
        self.labels = [view(p, np.uint8, n_cap) for p in (c.labels0, c.labels1)]
        if c.ho_cap:
            self.ho = {k: i32(getattr(c, "ho_" + k), int(c.ho_cap)) for k in ("node", "begin", "end", "side")}

    def rows(self, side, begin, end):
        codes = np.stack([self.planes[side][f // 16, begin:end, f % 16] for f in range(self.F)], axis=1)
        return codes, self.labels[side][begin:end].astype(np.int64)


def _terminal(cnt, depth, c):
    return (cnt > 0).sum() <= 1 or (c.max_depth >= 0 and depth >= c.max_depth) or int(cnt.sum()) < c.min_samples_split


def search_and_finalize(c, reduce=None):
    """plan + histogram + gain + finalize of the current level (``mtb_level_fused`` without the kernels' insides).

    ``reduce`` (several ranks): callable that sums a list of arrays across the ranks, standing in for the
    peer push + owner reduction + record broadcast; every rank then finalizes redundantly, as the kernels do.
    """
    criterion = "gini" if c.criterion else "entropy"
    x = Ctx(c)
    if int(c.depth) == 0:  # ``init_root``: the level-0 kernel writes the root slot and the scalars itself
        for k, v in zip(("node", "begin", "end", "mode", "parent", "sibling"), (0, 0, int(c.n_local), BUILD, -1, -1)):
            x.slot[k][0] = v
        x.scal[:] = (1, 0, 1, 0, 0, 0, 0, 0)
        x.t["left"][0], x.t["parent"][0], x.t["depth"][0] = -1, -1, 0
    n = int(x.scal[0])
    if n == 0:
        x.scal[1] = 0
        return
    depth, cur, F, C, B = int(c.depth), int(c.cur), x.F, x.C, x.B
    x.cursors[: 2 * n] = 0
    nodes0 = int(x.scal[2])
    n_split = n_pair = units = 0
    searched = [s for s in range(n) if int(x.slot["mode"][s]) in (BUILD, DERIVE)]
    hists = []
    for s in searched:
        codes, y = x.rows(cur, int(x.slot["begin"][s]), int(x.slot["end"][s]))
        hist = np.zeros((F, B, C), dtype=np.int64)
        for f in range(F):
            np.add.at(hist[f], (codes[:, f], y), 1)
        hists.append(hist)
    if reduce is not None:
        hists = reduce(hists)
    hist_of = dict(zip(searched, hists))
    for s in range(n):
        mode = int(x.slot["mode"][s])
        x.rec_feature[s], x.rec_bin[s], x.part_flags[s] = -1, -1, 0
        if x.part_child is not None:
            x.part_child[s] = -1
        if mode not in (BUILD, DERIVE):
            continue
        node, b, e = int(x.slot["node"][s]), int(x.slot["begin"][s]), int(x.slot["end"][s])
        hist = hist_of[s]
        cnt = hist[0].sum(axis=0)
        x.t_counts[node] = cnt
        split = None
        if not _terminal(cnt, depth, c):
            split = best_split(hist, np.full(F, B), criterion)
        x.t["feature"][node], x.t["bin"][node], x.t_gain[node] = -1, -1, np.nan
        if split is None:
            continue
        f, bin_, gain, left = split
        right = cnt - left
        tl, tr = _terminal(left, depth + 1, c), _terminal(right, depth + 1, c)
        hl = (not tl) and int(left.sum()) <= c.small_rows
        hr = (not tr) and int(right.sum()) <= c.small_rows
        pair = not (tl and tr)
        x.rec_feature[s], x.rec_bin[s] = f, bin_
        x.t["feature"][node], x.t["bin"][node], x.t_gain[node] = f, bin_, gain
        rows_needed = c.part_child == 0 or c.max_depth < 0 or depth + 2 < c.max_depth
        keep_l, keep_r = (not tl) and (rows_needed or hl), (not tr) and (rows_needed or hr)
        x.part_flags[s] = ((1 if keep_l else 0) | (2 if keep_r else 0)) if pair else 0
        lid = nodes0 + 2 * n_split
        rid = lid + 1
        n_split += 1
        x.t["left"][node] = lid
        for child, counts in ((lid, left), (rid, right)):
            x.t_counts[child] = counts
            x.t["parent"][child], x.t["depth"][child] = node, depth + 1
            x.t["feature"][child], x.t["bin"][child], x.t["left"][child], x.t_gain[child] = -1, -1, -1, np.nan
        if not pair:
            continue
        a, bb = 2 * n_pair, 2 * n_pair + 1
        n_pair += 1
        units += (e - b + 1023) // 1024
        if x.part_child is not None:
            x.part_child[s] = a
        x.nslot["node"][a], x.nslot["node"][bb] = lid, rid
        x.nslot["parent"][a] = x.nslot["parent"][bb] = s
        x.nslot["sibling"][a], x.nslot["sibling"][bb] = bb, a
        x.nslot["src"][a], x.nslot["src"][bb] = 2 * s, 2 * s + 1
        if hl or hr:
            ma = NONE if tl else (HANDOFF if hl else BUILD)
            mb = NONE if tr else (HANDOFF if hr else BUILD)
        elif not tl and not tr:
            small_left = int(left.sum()) <= int(right.sum())
            ma, mb = (BUILD, DERIVE) if small_left else (DERIVE, BUILD)
        else:
            ma, mb = (NONE if tl else BUILD), (NONE if tr else BUILD)
        x.nslot["mode"][a], x.nslot["mode"][bb] = ma, mb
    x.scal[1], x.scal[2], x.scal[4] = 2 * n_pair, nodes0 + 2 * n_split, units


def partition_and_commit(c):
    """``k_partition`` + ``commit_phase``: kept children's rows go to the other ping-pong side, left rows
    from the parent's first row upwards, right rows from its last row downwards; the next level's slots get
    their row ranges; handed-off children are listed; the next slot count becomes current."""
    x = Ctx(c)
    n, cur = int(x.scal[0]), int(c.cur)
    for s in range(n):
        flags = int(x.part_flags[s])
        f, bin_ = int(x.rec_feature[s]), int(x.rec_bin[s])
        if f < 0:
            continue
        b, e = int(x.slot["begin"][s]), int(x.slot["end"][s])
        go_left = x.planes[cur][f // 16, b:e, f % 16] <= bin_
        rows = np.arange(b, e)
        for keep, sel, cursor in ((flags & 1, go_left, 2 * s), (flags & 2, ~go_left, 2 * s + 1)):
            k = int(sel.sum())
            x.cursors[cursor] = k  # the kernel counts every row of the side, kept or not
            # The kernel skips rows nobody will read again (``keep`` clear): on the fused path a child's
            # histogram is accumulated while its parent's rows stream by.  The stand-in searches a child
            # from its own rows, so it always writes them; the two sides never overlap.
            if k == 0:
                continue
            dst = np.arange(b, b + k) if cursor % 2 == 0 else np.arange(e - 1, e - 1 - k, -1)
            x.planes[cur ^ 1][:, dst] = x.planes[cur][:, rows[sel]]
            x.labels[cur ^ 1][dst] = x.labels[cur][rows[sel]]
    for j in range(int(x.scal[1])):
        p, side = int(x.nslot["src"][j]) >> 1, int(x.nslot["src"][j]) & 1
        b = e = 0
        if x.nslot["mode"][j] != NONE:
            if side == 0:
                b = int(x.slot["begin"][p])
                e = b + int(x.cursors[2 * p])
            else:
                e = int(x.slot["end"][p])
                b = e - int(x.cursors[2 * p + 1])
        x.nslot["begin"][j], x.nslot["end"][j] = b, e
        if x.nslot["mode"][j] == HANDOFF:
            k = int(x.scal[5])
            x.scal[5] = k + 1
            if k < int(c.ho_cap):
                x.ho["node"][k], x.ho["begin"][k], x.ho["end"][k], x.ho["side"][k] = int(x.nslot["node"][j]), b, e, cur ^ 1
    x.scal[0] = x.scal[1]


def dispatch(reduce_of_thread=None):
    """Replacement for ``cuda_engine._call``: routes native entry-point names to the stand-ins.

    ``reduce_of_thread``: optional dict ``thread id -> reduce callable`` for thread-rank tests (one patched
    ``_call`` serves all ranks of the process)."""
    import threading

    def call(name, *args):
        if name == "mtb_pack_nodes":
            return pack_nodes(*args)
        if name == "mtb_predict":
            return predict(*args)
        ctx = args[0]._obj
        reduce = (reduce_of_thread or {}).get(threading.get_ident())
        if name == "mtb_level_fused" or name == "mtb_level_finalize":
            search_and_finalize(ctx, reduce)
        elif name == "mtb_level_fused2":  # (parents, children): partition the parents' rows, then search the children
            partition_and_commit(ctx)
            search_and_finalize(args[1]._obj, reduce)
        elif name == "mtb_level_commit":
            partition_and_commit(ctx)
        elif name in ("mtb_level_plan", "mtb_level_hist", "mtb_level_gain", "mtb_level_partition"):
            pass  # folded into the two stand-ins above
        else:
            raise NotImplementedError(name)

    return call


def pack_nodes(feature, left, label, n, out, stream):
    """``mtb_pack_nodes``: 16-byte node records {feature, left child, leaf label, 0}."""
    n = n.value
    rec = view(out.value, np.int32, 4 * n).reshape(n, 4)
    rec[:, 0], rec[:, 1] = view(feature.value, np.int32, n), view(left.value, np.int32, n)
    rec[:, 2], rec[:, 3] = view(label.value, np.int32, n), 0
    return 0


def predict(X, dtype, n, F, ldx, packed, thr, n_nodes, out_leaf, out_label, stream):
    """``mtb_predict``: walk every row down the packed tree (``x <= threshold`` goes to the left child)."""
    n, F, ldx, n_nodes = n.value, F.value, ldx.value, n_nodes.value
    np_dt, ct = {0: (np.float32, ctypes.c_float), 1: (np.uint8, ctypes.c_uint8), 2: (np.float64, ctypes.c_double)}[dtype.value]
    rows = np.ctypeslib.as_array(ctypes.cast(X.value, ctypes.POINTER(ct)), shape=(n * ldx,)).reshape(n, ldx)[:, :F]
    rec = view(packed.value, np.int32, 4 * n_nodes).reshape(n_nodes, 4)
    t_ct = {0: ctypes.c_float, 1: ctypes.c_int32, 2: ctypes.c_double}[dtype.value]
    t = np.ctypeslib.as_array(ctypes.cast(thr.value, ctypes.POINTER(t_ct)), shape=(n_nodes,))
    node = np.zeros(n, dtype=np.int64)
    active = np.flatnonzero(rec[node, 0] >= 0)
    while active.size:
        cur = node[active]
        go_left = rows[active, rec[cur, 0]] <= t[cur]
        node[active] = rec[cur, 1] + np.where(go_left, 0, 1)
        active = active[rec[node[active], 0] >= 0]
    if out_leaf.value:
        view(out_leaf.value, np.int32, n)[:] = node
    if out_label.value:
        view(out_label.value, np.int32, n)[:] = rec[node, 2]
    return 0
