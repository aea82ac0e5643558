"""CUDA engine: device data preparation, the level loop and batched predict.

Host-side driver for the sm_100a kernels in ``mpitree_b200/csrc`` (K1 encode,
K2 split search, K3 partition, K4 predict).  It replaces the reference's recursive
``_make_tree`` (``mpitree/tree/decision_tree.py:93-166`` / ``:364-479``) with a
breadth-first loop that issues a fixed, small number of launches per tree level and
reads back 8 integers per level.

All device buffers are torch tensors (caching allocator, stream semantics); kernels
are reached through the C ABI in ``include/mpitree_b200.h``.
"""

from __future__ import annotations

import ctypes
import os
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np

from ..models.flat_tree import FlatTree
from .binning import BinMapper
from .cpu_engine import SplitParams, probe_dot_mode

__all__ = [
    "available",
    "supports",
    "fit_from_host",
    "fit_from_codes",
    "fit_from_torch",
    "predict_leaf_ids",
    "DeviceDataset",
    "TreeBuilder",
    "KERNEL_LAUNCHES",
    "clear_caches",
]

MAX_CLASSES = 128
SMALL_ROWS_DEFAULT = 2048  # hand a node to the subtree kernel at or below this many rows
SMALL_ROWS_MAX = 8192      # row indices live in shared memory as uint16 next to the histogram tile
NVLS_MAX_SLOTS = 2048      # histogram slabs per rank bound to the multicast object in reduce="nvls"
SAMPLE_ROWS = 1 << 18
ENCODE_CHUNK_ROWS = 1 << 22

#: running count of native kernel launches issued by this process (bench.py reports it)
KERNEL_LAUNCHES = {"count": 0}


def available() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except ImportError:  # pragma: no cover
        return False


def supports(*, numeric: bool, n_classes: int) -> Tuple[bool, str]:
    # (object / string columns are fine: they are ordinal-coded on the host, where the strings live, and the
    # tree is grown from the codes on the device: ``fit_from_codes``)
    if n_classes > MAX_CLASSES:
        return False, f"more than {MAX_CLASSES} classes"
    return True, ""


def _call(name: str, *args) -> None:
    from . import _lib as L

    lib = L.load()
    rc = getattr(lib, name)(*args)
    KERNEL_LAUNCHES["count"] += 1
    if rc != 0:
        L.check(rc, name)


def _stream_ptr():
    import torch

    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t) -> int:
    return int(t.data_ptr())


PLOGP_N = 1024
_PLOGP_CACHE: dict = {}


def plogp_table(device):
    """Device table of ``(k/n) * log2(k/n)`` for ``1 <= k <= n <= PLOGP_N`` computed by numpy.

    Rounding-level ties between mathematically equal split costs occur in small
    nodes; evaluating those with the host's own ``np.log2`` values makes the device
    arithmetic bit-identical to the reference's numpy arithmetic there
    (``mpitree/tree/decision_tree.py:49-51``).
    """
    import torch

    key = str(device)
    if key not in _PLOGP_CACHE:
        rows = []
        for n in range(1, PLOGP_N + 1):
            p = np.arange(1, n + 1) / n
            rows.append(p * np.log2(p))
        _PLOGP_CACHE[key] = torch.from_numpy(np.concatenate(rows)).to(device)
    return _PLOGP_CACHE[key]


KLOGK_N = 128  # interval centres of the k*log2(k) evaluation (csrc/impurity.cuh)


def klogk_table(device):
    """Device constants of the arithmetic ``k * log2(k)`` evaluation used for large nodes.

    ``1/m0`` for the 128 interval centres ``m0 = 1 + (i + 1/2)/128``, ``e + log2 m0`` for every (exponent,
    interval) and the six polynomial coefficients of ``log2(1 + r)``; see ``cpu_engine.klogk`` for the
    numpy twin of the device function (``csrc/impurity.cuh::KlogK``).
    """
    import torch

    from .cpu_engine import klogk_constants

    key = ("klogk", str(device))
    if key not in _PLOGP_CACHE:
        _PLOGP_CACHE[key] = torch.from_numpy(klogk_constants()).to(device)
    return _PLOGP_CACHE[key]


# ---------------------------------------------------------------------------
# device data set: tiled codes + labels
# ---------------------------------------------------------------------------
_WORKSPACE: dict = {}


@dataclass
class DeviceDataset:
    """Bin codes ``[G][n_cap][16]`` (two ping-pong sides) and labels on one GPU."""

    n_local: int
    n_features: int
    n_classes: int
    n_cap: int
    G: int
    codes: list  # two uint8 tensors of G*n_cap*16 (+pad) bytes
    labels: list  # two uint8 tensors of n_cap (+pad) bytes
    device: object
    cur: int = 0

    @classmethod
    def allocate(cls, n_local: int, n_features: int, n_classes: int, device) -> "DeviceDataset":
        import torch

        G = (n_features + 15) // 16
        n_cap = ((max(n_local, 1) + 1023) // 1024) * 1024
        # the two ping-pong sides are the largest allocations of a fit: keep the last set around
        # (repeated fits of the same shape then skip cudaMalloc/cudaFree entirely)
        key = (str(device), G * n_cap * 16 + 256, n_cap + 256)
        cached = _WORKSPACE.pop("buffers", None)
        if cached is not None and cached[0] == key:
            codes, labels = cached[1], cached[2]
        else:
            del cached
            codes = [torch.empty(key[1], dtype=torch.uint8, device=device) for _ in range(2)]
            labels = [torch.empty(key[2], dtype=torch.uint8, device=device) for _ in range(2)]
        return cls(n_local, n_features, n_classes, n_cap, G, codes, labels, device)

    def release(self) -> None:
        """Hand the big buffers back to the one-entry workspace cache (call when the fit is done)."""
        if self.codes is not None:
            key = (str(self.device), self.G * self.n_cap * 16 + 256, self.n_cap + 256)
            _WORKSPACE["buffers"] = (key, self.codes, self.labels)
            self.codes = self.labels = None

    # -- encode ---------------------------------------------------------------
    def encode_float(self, X, mapper: BinMapper, row0: int = 0, mismatch=None) -> None:
        """``X``: CUDA tensor ``(n, F)`` float32/float64, row-major."""
        import torch

        assert X.is_cuda and X.dim() == 2
        if X.stride(1) != 1:
            X = X.contiguous()
        key = id(mapper)
        if getattr(self, "_edges_key", None) != key:
            edges = mapper.edges_matrix(np.float64)
            self._edges = torch.from_numpy(edges).to(self.device)
            self._nbins = torch.from_numpy(mapper.n_bins.astype(np.int32)).to(self.device)
            self._edges_key = key
        dtype = {torch.float32: 0, torch.float64: 1}[X.dtype]
        if (
            os.environ.get("MTB_ENCODE", "quads") != "scalar"
            and dtype == 0
            and mismatch is None
            and self.n_features % 4 == 0
            and X.stride(0) % 4 == 0
            and X.data_ptr() % 16 == 0
        ):
            _call(
                "mtb_encode_f32_quads" if os.environ.get("MTB_ENCODE") == "quads" else "mtb_encode_f32_slots",
                ctypes.c_void_p(_p(X)),
                ctypes.c_int64(X.shape[0]),
                ctypes.c_int64(self.n_features),
                ctypes.c_int64(X.stride(0)),
                ctypes.c_void_p(_p(self._edges)),
                ctypes.c_void_p(_p(self._nbins)),
                ctypes.c_void_p(_p(self.codes[self.cur])),
                ctypes.c_int64(self.n_cap),
                ctypes.c_int64(row0),
                _stream_ptr(),
            )
            return
        _call(
            "mtb_encode_float",
            ctypes.c_void_p(_p(X)),
            ctypes.c_int(dtype),
            ctypes.c_int64(X.shape[0]),
            ctypes.c_int64(self.n_features),
            ctypes.c_int64(X.stride(0)),
            ctypes.c_void_p(_p(self._edges)),
            ctypes.c_void_p(_p(self._nbins)),
            ctypes.c_void_p(_p(self.codes[self.cur])),
            ctypes.c_int64(self.n_cap),
            ctypes.c_int64(row0),
            ctypes.c_void_p(_p(mismatch) if mismatch is not None else 0),
            _stream_ptr(),
        )

    def encode_u8(self, X, lut=None, row0: int = 0) -> None:
        """``X``: CUDA uint8 tensor ``(n, F)``; ``lut`` optional uint8 ``(F, 256)`` value->code."""
        assert X.is_cuda and X.dim() == 2
        if X.stride(1) != 1:
            X = X.contiguous()
        _call(
            "mtb_encode_u8",
            ctypes.c_void_p(_p(X)),
            ctypes.c_int64(X.shape[0]),
            ctypes.c_int64(self.n_features),
            ctypes.c_int64(X.stride(0)),
            ctypes.c_void_p(_p(lut) if lut is not None else 0),
            ctypes.c_void_p(_p(self.codes[self.cur])),
            ctypes.c_int64(self.n_cap),
            ctypes.c_int64(row0),
            _stream_ptr(),
        )

    def set_labels(self, y_codes, row0: int = 0) -> None:
        import torch

        n = y_codes.shape[0]
        self.labels[self.cur][row0 : row0 + n].copy_(y_codes.to(torch.uint8), non_blocking=True)

    def codes_rowmajor(self):
        """Debug/test helper: current codes as an ``(n_local, F)`` uint8 tensor."""
        c = self.codes[self.cur][: self.G * self.n_cap * 16].view(self.G, self.n_cap, 16)
        return c[:, : self.n_local].permute(1, 0, 2).reshape(self.n_local, self.G * 16)[:, : self.n_features]


# ---------------------------------------------------------------------------
# tree builder
# ---------------------------------------------------------------------------
@dataclass
class BuildStats:
    levels: int = 0
    launches: int = 0
    level_ms: List[Dict[str, float]] = field(default_factory=list)
    hist_bytes: int = 0
    m_cap: int = 0
    FT: int = 0
    stages: int = 0
    smem_bytes: int = 0
    reduce: str = "none"
    world: int = 1
    handed_off: int = 0
    tile_layout: str = "interleaved"


_EPOCH = {"next": 0}  # barrier epochs are monotonic for the life of the process


class TreeBuilder:
    """Level-wise builder bound to one :class:`DeviceDataset` (one rank).

    Cross-rank histogram reduction when ``world > 1``:

    * ``reduce="p2p"`` — the product path: one cooperative launch per level pushes
      partial slabs into the owners' staging areas over NVLink, synchronises with
      in-kernel flags, evaluates gains feature-parallel on the owners and broadcasts
      16-byte-class best-split records to every rank's mailbox.  No NCCL call.
    * ``reduce="nccl"`` — in-house baseline: local-hist kernel -> ``ncclAllReduce`` ->
      gain kernel (every rank evaluates every feature).
    * ``peers=<LocalRegion view>`` — several virtual ranks in one process, driven in
      lock-step by :func:`build_lockstep` (same kernels, phases launched separately).
    * ``reduce="nvls"`` — staged: the ``p2p`` protocol with the slabs in a
      multicast-bound region, owners read cross-rank sums with ``multimem.ld_reduce``.

    Other switches: ``fused`` / ``fuse_partition`` select the launch decomposition (one cooperative
    launch per level by default), ``async_levels`` enqueues a bounded-depth fit without read-backs,
    ``small_rows`` is the hand-off size of phase 2 (``None`` = automatic, 0 = never),
    ``h_cap`` / ``hist_budget_bytes`` size the histogram slots, ``tile_layout`` selects the shared-memory tile
    layout ("interleaved" by default, "planar"), ``profile`` records per-phase times.
    """

    def __init__(
        self,
        data: DeviceDataset,
        n_bins: int,
        params: SplitParams,
        *,
        n_global: Optional[int] = None,
        fused: bool = True,
        reduce: str = "auto",
        comm=None,
        peers=None,
        profile: bool = False,
        hist_budget_bytes: Optional[int] = None,
        h_cap: Optional[int] = None,
        min_free_bytes: Optional[int] = None,
        async_levels: bool = True,
        small_rows: Optional[int] = None,
        fuse_partition: bool = True,
        tile_layout: Optional[str] = None,
    ):
        import torch

        from . import _lib as L

        self.torch = torch
        self.L = L
        self.lib = L.load()
        self.data = data
        self.params = params
        self.fused = fused
        self.profile = profile
        self.comm = comm
        self.world = 1 if comm is None else comm.size
        self.rank = 0 if comm is None else comm.rank
        self.region = None
        if peers is not None:  # virtual ranks in one process
            self.world, self.rank = peers.world, peers.rank
            reduce = "lockstep"
        elif self.world == 1:
            reduce = "none"
        elif reduce == "auto":
            reduce = "p2p"
        # "nvls" (staged): the p2p protocol with the histogram slabs in multicast-bound memory, so
        # the owners read cross-rank sums from the switch instead of receiving pushed partials
        want_nvls = reduce == "nvls"
        if want_nvls:
            reduce = "p2p"
        if reduce not in ("none", "p2p", "nccl", "lockstep"):
            raise ValueError("reduce must be 'auto', 'p2p', 'nccl' or 'nvls'")
        self.reduce = reduce
        self.nvls = False
        dev = data.device
        F, C, G = data.n_features, data.n_classes, data.G
        if C > MAX_CLASSES:
            raise ValueError(f"the CUDA engine supports at most {MAX_CLASSES} classes")
        B = _bin_stride(n_bins)
        if B > 256:
            raise ValueError("at most 256 bins per feature")
        self.F, self.C, self.G, self.B = F, C, G, B
        n_global = int(n_global if n_global is not None else data.n_local)
        self.n_global = n_global
        P = self.world
        peer_mode = reduce in ("p2p", "lockstep")
        self.fpo = (F + P - 1) // P if peer_mode else F

        ctx = L.MtbCtx()
        ctx.n_cap, ctx.n_local, ctx.F, ctx.G, ctx.B, ctx.C = data.n_cap, data.n_local, F, G, B, C
        ctx.max_depth = -1 if params.max_depth is None else int(params.max_depth)
        ctx.min_samples_split = int(params.min_samples_split)
        ctx.criterion = 0 if params.criterion == "entropy" else 1
        ctx.dot_mode = probe_dot_mode()
        self._plogp = plogp_table(dev)
        ctx.plogp, ctx.plogp_n = _p(self._plogp), PLOGP_N
        self._klogk = klogk_table(dev)
        ctx.klogk, ctx.klogk_n = _p(self._klogk), KLOGK_N
        ctx.world, ctx.rank = (P, self.rank) if peer_mode else (1, 0)
        ctx.fpo = self.fpo
        # in-kernel peer barriers give up after this long (a rank that died or raised between two
        # launches would otherwise leave the others spinning for ever); 0 = wait for ever
        ctx.barrier_timeout_ns = int(float(os.environ.get("MTB_BARRIER_TIMEOUT_S", "30")) * 1e9)

        # features per shared-memory tile: the largest that leaves >= 2 TMA stages
        self.stats = BuildStats(reduce=reduce, world=P)
        for ft in (16, 8, 4, 2, 1):
            ctx.FT = ft
            th, sm, st, rp = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
            if self.lib.mtb_hist_config(ctypes.byref(ctx), ctypes.byref(th), ctypes.byref(sm), ctypes.byref(st), ctypes.byref(rp)) == 0:
                self.stats.FT, self.stats.stages, self.stats.smem_bytes = ft, st.value, sm.value
                break
        else:
            raise ValueError("bins x classes too large for a shared-memory histogram tile")
        # shared-memory tile layout: "interleaved" = [B*C][16], two banks per feature (default when a tile
        # holds a whole 16-feature group; root level 2.41 vs 2.80 ms on 100M x 64); "planar" = [f][B][C]
        # (MTB_TILE=planar or the keyword)
        layout = tile_layout or os.environ.get("MTB_TILE", "interleaved")
        if layout not in ("planar", "interleaved"):
            raise ValueError(f"unknown tile layout {layout!r}")
        ctx.tile_layout = 1 if (layout == "interleaved" and self.stats.FT == 16) else 0
        self.stats.tile_layout = "interleaved" if ctx.tile_layout else "planar"

        slab_words = G * 16 * B * C
        self.slab_words = slab_words
        if h_cap is None:
            agreed = min_free_bytes is not None or hist_budget_bytes is not None
            if hist_budget_bytes is None:
                free = min_free_bytes if min_free_bytes is not None else torch.cuda.mem_get_info(dev)[0]
                hist_budget_bytes = int(free * 0.5)
            depth_cap = 1 << 30 if params.max_depth is None else 1 << max(int(params.max_depth) - 1, 0)
            rows_cap = max(2, n_global // max(int(params.min_samples_split), 2) + 1)
            # single rank: two ping-pong sides; peers: one side of local partials + staging + 2/P reduced
            per_slot = slab_words * 4 * (2.0 + 2.0 / P if peer_mode else 2.0)
            h_cap = min(depth_cap, rows_cap, max(2, int(hist_budget_bytes // per_slot)))
            if want_nvls:
                h_cap = min(h_cap, NVLS_MAX_SLOTS)  # the slabs are bound to a multicast object: keep it modest
            h_cap = max(2, h_cap + (h_cap & 1))
            if comm is not None and comm.size > 1 and not agreed:
                h_cap = min(comm.allgather_object(h_cap))  # every rank must use the same layout
        self.h_cap = int(h_cap)  # histogram slabs per ping-pong side
        m_cap = 2 * self.h_cap + 2  # slot-array capacity: a level may emit two children per slot
        self.m_cap = m_cap
        self.stats.m_cap = self.h_cap
        self.stats.hist_bytes = 2 * self.h_cap * slab_words * 4
        node_cap = 2 * n_global + 1 if params.max_depth is None else min((1 << (int(params.max_depth) + 1)) - 1, 2 * n_global + 1)
        self.node_cap = max(3, min(node_cap, 1 << 16))
        # With a bounded depth whose worst-case frontier and node count fit, the whole fit is
        # enqueued without reading anything back between levels (kernels of levels that turn
        # out to be empty return immediately).
        worst_frontier = None if params.max_depth is None else 1 << max(int(params.max_depth) - 1, 0)
        if os.environ.get("MTB_ASYNC") == "0":
            async_levels = False
        self.async_levels = bool(
            async_levels and not profile and params.max_depth is not None and int(params.max_depth) >= 1
            and worst_frontier <= self.h_cap and node_cap <= (1 << 22) and reduce != "lockstep"
        )
        # phase 2 (hand-off): nodes with <= small_rows global rows are finished by the subtree kernel.
        # Default: only when the level-wise frontier could outgrow the histogram slot capacity.
        if small_rows is None:
            small_rows = 0 if (worst_frontier is not None and worst_frontier <= self.h_cap) else SMALL_ROWS_DEFAULT
            # A single rank has no rows to gather, so finishing small nodes with the subtree kernel only has
            # to beat more level-wise launches over thousands of tiny slots (whose gain runs in the exact
            # small-node arithmetic): 50M x 128, depth 16, one B200: 104 ms with the hand-off, 174 ms without.
            if self.world == 1 and reduce == "none" and params.max_depth is not None and int(params.max_depth) >= 12:
                small_rows = SMALL_ROWS_DEFAULT
            if small_rows and os.environ.get("MTB_SMALL_ROWS"):  # tuning knob for the hand-off size
                small_rows = max(1, int(os.environ["MTB_SMALL_ROWS"]))
        if small_rows and reduce in ("lockstep", "nccl"):
            small_rows = 0
        if small_rows and self.lib.mtb_subtree_config(ctypes.c_int64(B), ctypes.c_int64(C)) == 0:
            small_rows = 0
        self.small_rows = int(min(small_rows, SMALL_ROWS_MAX))
        ctx.small_rows = self.small_rows
        if self.small_rows:
            self.async_levels = False
            self.ho_cap = int(min(max(n_global // 2 + 1, 16), 1 << 22))
            self.ho = {k: torch.zeros(self.ho_cap, dtype=torch.int32, device=dev) for k in ("node", "begin", "end", "side")}
            ctx.ho_cap = self.ho_cap
            ctx.ho_node, ctx.ho_begin, ctx.ho_end, ctx.ho_side = (_p(self.ho[k]) for k in ("node", "begin", "end", "side"))
        if self.async_levels:
            self.node_cap = max(3, node_cap)

        i32 = torch.int32
        z = lambda n, dt=i32: torch.zeros(max(int(n), 1), dtype=dt, device=dev)  # noqa: E731
        if reduce == "p2p":
            # peers never read the previous level's local partials (reduced slabs live in rstore)
            if want_nvls and fused and self.stats.FT == 16:
                reg = _nvls_region(self.h_cap * slab_words * 4, comm)
                one = reg.as_tensor()[: self.h_cap * slab_words * 4].view(i32)
                ctx.mc_hist[0] = ctx.mc_hist[1] = int(reg.mc_ptr)
                self.nvls = True
                self.stats.reduce = "nvls"
            else:
                one = torch.empty(self.h_cap * slab_words, dtype=i32, device=dev)
            self.hist = [one, one]
        else:
            # no initial fill: every level zeroes the slabs it is going to accumulate into (prepare_level /
            # _build_async) and overwrites the derived ones; hundreds of MB of memset per fit saved
            self.hist = [torch.empty(self.h_cap * slab_words, dtype=i32, device=dev) for _ in range(2)]
        # the per-slot int32 tables share one zero-filled allocation (one memset instead of twenty);
        # every table starts on a 256-byte boundary
        stride = _round_up(m_cap + 64, 64)  # (with at least 256 bytes of slack behind it, as separate allocations had)
        names = [f"slot{side}_{k}" for side in range(2) for k in ("node", "begin", "end", "mode", "parent", "sibling")]
        names += ["nslot_src", "rec_feature", "rec_bin", "part_flags", "part_child"]
        sizes = {n: stride for n in names}
        sizes.update(cursors=_round_up(2 * m_cap + 64, 64), unit_prefix=_round_up(2 * (m_cap + 1) + 64, 64), scalars=128)
        arena = z(sum(sizes.values()))
        tables, off = {}, 0
        for name, size in sizes.items():
            tables[name] = arena[off : off + size]
            off += size
        self._arena = arena
        self.slots = [{k: tables[f"slot{side}_{k}"][:m_cap] for k in ("node", "begin", "end", "mode", "parent", "sibling")}
                      for side in range(2)]
        self.nslot_src = tables["nslot_src"][:m_cap]
        self.rec_feature, self.rec_bin, self.part_flags = (tables[k][:m_cap] for k in ("rec_feature", "rec_bin", "part_flags"))
        self.cursors = tables["cursors"][: 2 * m_cap]
        self.scalars = tables["scalars"][:8]
        self.unit_prefix = tables["unit_prefix"][: 2 * (m_cap + 1)]
        # fused partition + next-level histogram (levels >= 1 of the fused paths)
        self.fuse_partition = bool(fuse_partition and fused and reduce in ("none", "p2p"))
        if self.fuse_partition:
            max_units = data.n_cap // 1024 + m_cap + 2
            self.part_child = tables["part_child"][:m_cap]
            self.mask_words = z(max_units * 68)  # 272-byte unit records (32 bit words, 32 prefixes, 2 bases, pad)
        # one cooperative launch per level (these kernels zero the slabs they accumulate into)
        self._coop = reduce == "p2p" or (reduce == "none" and fused)
        ctx.zero_next = 1 if self.fuse_partition else 0
        self._ctx_prev = None
        self._bound_tree = None  # tree arrays the context currently points at (see _bind)
        self._alloc_tree(self.node_cap)

        # record mailbox: fb_cost | fb_bin | fb_left | node_cnt (layout shared with the kernels)
        MF = m_cap * F
        self._rec_off = (0, MF * 8, MF * 12, MF * 12 + MF * C * 4)
        rec_bytes = _round_up(self._rec_off[3] + m_cap * C * 4, 256)
        # staging cells [source][slot][fpo*B*C] u32 (or u16-packed) + one format word per (source, slot)
        stage_bytes = _round_up(P * self.h_cap * (self.fpo * B * C + 1) * 4, 256) if peer_mode else 0
        self._layout = (0, stage_bytes, stage_bytes + rec_bytes)  # stage | mailbox | flags
        total = stage_bytes + rec_bytes + 256
        if reduce == "p2p":
            ptrs = _shared_region(total, comm).ptrs
        elif reduce == "lockstep":
            if peers.nbytes < total:
                raise ValueError(f"LocalRegion too small: need {total} bytes")
            ptrs = peers.ptrs
        else:
            self._mailbox = torch.zeros(total, dtype=torch.uint8, device=dev)
            ptrs = [_p(self._mailbox)]
        me = self.rank if peer_mode else 0
        self._mailbox_ptr = ptrs[me] + self._layout[1]
        if peer_mode:
            for q in range(P):
                ctx.peer_hist[q] = ptrs[q] + self._layout[0]
                ctx.peer_rec[q] = ptrs[q] + self._layout[1]
                ctx.peer_flags[q] = ptrs[q] + self._layout[2]
        ctx.h_cap = self.h_cap
        if peer_mode:
            rs_words = self.h_cap * self.fpo * B * C
            self.rstore = [z(rs_words), z(rs_words)]
            ctx.rstore[0], ctx.rstore[1] = _p(self.rstore[0]), _p(self.rstore[1])
        self.ctx = ctx
        self.side = 0  # which slot set is "current"
        self.hcur = 0
        self.n_slots, self.n_nodes, self.depth = 1, 1, 0
        self._timers = None
        if profile:
            self._timers = torch.zeros(16, dtype=torch.int64, device=dev)
            ctx.timers = _p(self._timers)

    @staticmethod
    def region_bytes(world: int, F: int, B: int, C: int, h_cap: int) -> int:
        """Bytes each rank's symmetric region needs (for :class:`LocalRegion` sizing)."""
        m_cap = 2 * h_cap + 2
        MF = m_cap * F
        rec = _round_up(MF * 12 + MF * C * 4 + m_cap * C * 4, 256)
        fpo = (F + world - 1) // world
        return _round_up(world * h_cap * (fpo * B * C + 1) * 4, 256) + rec + 256

    def close(self) -> None:
        if self.region is not None:
            self.region.close()
            self.region = None

    # ------------------------------------------------------------------ tree
    def _alloc_tree(self, cap: int, old=None) -> None:
        """Tree arrays (SoA, ``cap`` nodes) as views of one int32 buffer, so that a finished tree comes
        back in a single device-to-host copy.  No initial fill: the kernels write every field of every
        node they create (the root's by ``init_root``)."""
        torch = self.torch
        dev = self.data.device
        C = self.C
        cap = int(cap)
        st = _round_up(cap, 64)  # every array starts on a 256-byte boundary
        buf = torch.empty((7 + C) * st, dtype=torch.int32, device=dev)
        new = {k: buf[i * st : i * st + cap] for i, k in enumerate(("feature", "bin", "left", "parent", "depth"))}
        new["gain"] = buf[5 * st : 7 * st].view(torch.float64)[:cap]
        new["counts"] = buf[7 * st : 7 * st + cap * C]
        if old is not None:
            for k, t in old.items():
                new[k][: t.shape[0]].copy_(t)
        self.tree = new
        self._tree_buf = buf
        self.node_cap = cap

    def _bind(self) -> None:
        """Point the context at this level's buffers.

        Most pointers never change during a fit: they are written once (again after the tree arrays
        were reallocated); per level only the ping-pong selectors and the two slot tables swap.
        """
        c, d = self.ctx, self.data
        if self._bound_tree is not self.tree:
            c.codes0, c.codes1 = _p(d.codes[0]), _p(d.codes[1])
            c.labels0, c.labels1 = _p(d.labels[0]), _p(d.labels[1])
            c.hist0, c.hist1 = _p(self.hist[0]), _p(self.hist[1])
            c.nslot_src = _p(self.nslot_src)
            mb = self._mailbox_ptr
            c.fb_cost, c.fb_bin, c.fb_left, c.node_cnt = (mb + o for o in self._rec_off)
            c.rec_feature, c.rec_bin, c.part_flags, c.cursors = (
                _p(self.rec_feature), _p(self.rec_bin), _p(self.part_flags), _p(self.cursors),
            )
            t = self.tree
            c.t_feature, c.t_bin, c.t_left, c.t_parent, c.t_depth = (
                _p(t["feature"]), _p(t["bin"]), _p(t["left"]), _p(t["parent"]), _p(t["depth"]),
            )
            c.t_gain, c.t_counts = _p(t["gain"]), _p(t["counts"])
            c.scalars, c.unit_prefix = _p(self.scalars), _p(self.unit_prefix)
            c.m_cap, c.node_cap = self.m_cap, self.node_cap
            if self.fuse_partition:
                c.part_child, c.mask_words = _p(self.part_child), _p(self.mask_words)
            keys = ("node", "begin", "end", "mode", "parent", "sibling")
            self._slot_ptrs = [tuple(_p(tab[k]) for k in keys) for tab in self.slots]
            self._bound_tree = self.tree
        c.cur, c.hcur = d.cur, self.hcur
        (c.slot_node, c.slot_begin, c.slot_end, c.slot_mode, c.slot_parent, c.slot_sibling) = self._slot_ptrs[self.side]
        (c.nslot_node, c.nslot_begin, c.nslot_end, c.nslot_mode, c.nslot_parent, c.nslot_sibling) = self._slot_ptrs[self.side ^ 1]

    def _launch(self, name: str) -> None:
        _call(name, ctypes.byref(self.ctx), _stream_ptr())
        self.stats.launches += 1

    # ------------------------------------------------------------ level steps
    def start(self) -> None:
        """Begin a fit.  Nothing is written to the device: the level-0 kernel initialises the root slot,
        the scalars and the root node (``init_root`` in csrc/build.cu)."""
        self.n_slots, self.n_nodes, self.depth = 1, 1, 0
        self.side, self.hcur = 0, 0
        if getattr(self, "_host_scal", None) is None:
            self._host_scal = self.torch.empty(8, dtype=self.torch.int32).pin_memory()
        self._phase_ns = []
        self._ctx_prev = None
        self.stats.levels = self.stats.launches = self.stats.handed_off = 0
        self.stats.level_ms = []

    def prepare_level(self) -> None:
        n_slots, n_nodes = self.n_slots, self.n_nodes
        if n_slots > self.h_cap:
            raise RuntimeError(
                f"frontier of {n_slots} nodes exceeds the histogram slot capacity {self.h_cap}; "
                "raise hist_budget_bytes, lower max_depth or raise min_samples_split"
            )
        if n_nodes + 2 * n_slots > self.node_cap:
            self._alloc_tree(max(2 * self.node_cap, n_nodes + 2 * n_slots), self.tree)
        self.ctx.depth = self.depth
        self.ctx.n_slots_host = n_slots
        self._bind()
        # slabs of BUILD slots must start at zero (DERIVE slots are overwritten); the cooperative
        # level kernels do that themselves
        if not self._coop:
            self.hist[self.hcur][: n_slots * self.slab_words].zero_()

    def search_level(self) -> None:
        """plan + histogram + reduction + gain + finalize for the current level."""
        if self._coop:
            if self.reduce == "p2p":
                self.ctx.epoch = _EPOCH["next"]
                _EPOCH["next"] += 2
            if self.depth == 0:
                self._root_pass()
            if self.fuse_partition and self._ctx_prev is not None:
                # partition of the previous level's rows + this level's histograms + search, one launch
                _call("mtb_level_fused2", ctypes.byref(self._ctx_prev), ctypes.byref(self.ctx), _stream_ptr())
                self.stats.launches += 1
            else:
                self._launch("mtb_level_fused")
        else:
            self._launch("mtb_level_plan")
            self._launch("mtb_level_hist")
            if self.reduce == "nccl":
                import torch.distributed as dist

                dist.all_reduce(self.hist[self.hcur][: self.n_slots * self.slab_words])
            self._launch("mtb_level_gain")
            self._launch("mtb_level_finalize")

    def _root_pass(self) -> None:
        """Depth 0 of a fit whose rows were handed over as raw row-major bytes (``root_input``): planes,
        labels and the root histogram in one pass (``k_root_cluster``), or - when the shape does not
        qualify - the usual retile + label copy, with the histogram left to the level kernel."""
        self.ctx.root_hist_done = 0
        src = getattr(self, "root_input", None)
        if src is None:
            return
        self.root_input = None
        X, y = src
        entry = self.lib.mtb_root_retile_hist if _root_mode() == "cluster" else self.lib.mtb_root_direct_hist
        rc = entry(
            ctypes.byref(self.ctx), ctypes.c_void_p(_p(X)), ctypes.c_int64(X.stride(0)), ctypes.c_int64(X.shape[0]),
            ctypes.c_void_p(_p(y)), _stream_ptr()) if X.shape[0] else 1
        KERNEL_LAUNCHES["count"] += 1
        if rc == 0:
            self.ctx.root_hist_done = 1
            self.stats.launches += 1
            return
        if rc < 0:
            self.L.check(rc, "mtb_root_retile_hist")
        self.data.encode_u8(X, None, 0)
        self.data.set_labels(y)

    def read_level(self) -> int:
        self._host_scal.copy_(self.scalars, non_blocking=True)
        self.torch.cuda.current_stream().synchronize()
        if self._timers is not None:
            t = self._timers.cpu().numpy().astype(np.float64) / 1e6
            names = ("plan|part+hist", "hist|plan", "push+bar", "gain", "bar", "finalize")
            row = {n + "_ms": float(t[i + 1] - t[i]) for i, n in enumerate(names)}
            if t[8] > t[0]:
                row["prepass_ms"] = float(t[8] - t[0])
            self._phase_ns.append(row)
        self._check_peers()
        self.n_next, self.n_nodes = int(self._host_scal[1]), int(self._host_scal[2])
        return self.n_next

    def _check_peers(self) -> None:
        """Raise when an in-kernel peer barrier gave up (``scalars[7]``, see ``peer_barrier`` in csrc/build.cu)."""
        if self.reduce == "p2p" and int(self._host_scal[7]) != 0:
            clear_caches(collective=False)  # flags, epochs and abort words of the cached regions are no longer in step
            raise RuntimeError(
                f"rank {self.rank}: a peer did not reach an in-kernel barrier within "
                f"{self.ctx.barrier_timeout_ns / 1e9:.0f} s (a rank failed or is far behind); the fit was abandoned"
            )

    def _swap_sides(self) -> None:
        if self.fuse_partition:
            # the partition of this level's rows runs at the start of the next level's launch
            self._ctx_prev = type(self.ctx).from_buffer_copy(self.ctx)
        else:
            self._launch("mtb_level_partition")
            self._launch("mtb_level_commit")
        self.data.cur ^= 1
        self.side ^= 1
        self.hcur ^= 1

    def advance_level(self) -> None:
        if self.n_next > 0:
            self._swap_sides()
        self.n_slots = self.n_next
        self.depth += 1
        self.stats.levels += 1

    # ----------------------------------------------------------------- build
    def build(self) -> FlatTree:
        torch = self.torch
        if self.reduce == "lockstep":
            raise RuntimeError("virtual ranks are driven by build_lockstep()")
        self.start()
        if self.async_levels:
            return self._build_async()
        events = []
        while self.n_slots > 0:
            self.prepare_level()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if self.profile else None
            if ev:
                ev[0].record()
            self.search_level()
            if ev:
                ev[1].record()
            self.read_level()
            if ev:
                ev[2].record()
            n_slots = self.n_slots
            self.advance_level()
            if ev:
                ev[3].record()
                events.append((n_slots, ev))
        if events:
            torch.cuda.synchronize()
            for i, (ns, ev) in enumerate(events):
                d = {"slots": ns, "search_ms": ev[0].elapsed_time(ev[1]), "partition_ms": ev[2].elapsed_time(ev[3])}
                if i < len(self._phase_ns):
                    d.update(self._phase_ns[i])
                self.stats.level_ms.append(d)
        trunk = self._download(self.n_nodes)
        if self.small_rows:
            from .subtree_phase import finish_subtrees

            trunk = finish_subtrees(self, trunk)
        return trunk

    def _build_async(self) -> FlatTree:
        """All levels enqueued back to back; one read-back at the end."""
        for depth in range(int(self.params.max_depth)):
            self.depth = depth
            self.n_slots = min(1 << depth, self.h_cap)  # upper bound: sizes the zeroing and launches
            self.ctx.depth = depth
            self.ctx.n_slots_host = self.n_slots
            self._bind()
            if not self._coop:
                self.hist[self.hcur][: self.n_slots * self.slab_words].zero_()
            self.search_level()
            self._swap_sides()
            self.stats.levels += 1
        self._host_scal.copy_(self.scalars, non_blocking=True)
        self.torch.cuda.current_stream().synchronize()
        self._check_peers()
        self.n_nodes = int(self._host_scal[2])
        return self._download(self.n_nodes)

    def _download(self, n_nodes: int) -> FlatTree:
        cap, C = self.node_cap, self.C
        st = _round_up(cap, 64)
        if (7 + C) * st * 4 <= (8 << 20) or cap <= 2 * n_nodes:
            host = self._tree_buf.cpu().numpy()  # one copy of all seven arrays
            t = {k: host[i * st : i * st + n_nodes] for i, k in enumerate(("feature", "bin", "left", "parent", "depth"))}
            t["gain"] = host[5 * st : 7 * st].view(np.float64)[:n_nodes]
            t["counts"] = host[7 * st : 7 * st + n_nodes * C]
        else:
            t = {k: v[: n_nodes * (C if k == "counts" else 1)].cpu().numpy() for k, v in self.tree.items()}
        left = t["left"].astype(np.int32)
        right = np.where(left >= 0, left + 1, -1).astype(np.int32)
        counts = t["counts"].view(np.uint32).reshape(n_nodes, C).astype(np.int64)
        return FlatTree(
            feature=t["feature"].astype(np.int32),
            bin=t["bin"].astype(np.int32),
            left=left,
            right=right,
            parent=t["parent"].astype(np.int32),
            depth=t["depth"].astype(np.int32),
            counts=counts,
            gain=t["gain"].astype(np.float64),
            thresholds=[None] * n_nodes,
        )


_REGIONS: dict = {}


def _nvls_region(nbytes: int, comm):
    """Multicast-bound region for the histogram slabs, reused across fits like :func:`_shared_region`."""
    from ..parallel.nvls import NvlsRegion

    key = ("nvls", comm.size, comm.rank)
    reg = _REGIONS.get(key)
    if reg is None or reg.nbytes < int(nbytes):
        if reg is not None:
            reg.close()
        reg = NvlsRegion(int(nbytes), comm)
        _REGIONS[key] = reg
    return reg


def _shared_region(nbytes: int, comm):
    """Peer-mapped region reused across fits (mapping costs milliseconds; epochs keep flags valid)."""
    from ..parallel.symm import IpcRegion

    key = (comm.size, comm.rank)
    reg = _REGIONS.get(key)
    need = int(nbytes)  # a pure function of (h_cap, shapes): identical on every rank by construction
    if reg is None or reg.nbytes < need:
        if reg is not None:
            reg.close()
        reg = IpcRegion(need, comm)
        _REGIONS[key] = reg
    return reg


def fit_virtual_ranks(shards, n_classes: int, n_bins: int, params: SplitParams, *, h_cap: int = 64):
    """Fit one tree from several row shards held by *virtual ranks* on this GPU.

    ``shards``: list of ``(X_uint8_cuda, y_uint8_cuda)``.  Returns the per-rank trees
    (they must be identical).  See :func:`build_lockstep`.
    """
    from ..parallel.symm import LocalRegion

    P = len(shards)
    F = int(shards[0][0].shape[1])
    B = _bin_stride(n_bins)
    region = LocalRegion(TreeBuilder.region_bytes(P, F, B, n_classes, h_cap), P)
    n_global = sum(int(x.shape[0]) for x, _ in shards)
    builders = []
    for r, (X, y) in enumerate(shards):
        data = DeviceDataset.allocate(int(X.shape[0]), F, n_classes, X.device)
        data.encode_u8(X)
        data.set_labels(y)
        builders.append(TreeBuilder(data, n_bins, params, n_global=n_global, peers=region.view(r), h_cap=h_cap))
    return build_lockstep(builders)


def _bin_stride(n_bins: int) -> int:
    """Bins per class row of a histogram slab: a power of two, at least 32 (a lane of the gain phase
    owns ``B / 32`` consecutive bins and loads them with one aligned vector load)."""
    B = 32
    while B < int(n_bins):
        B *= 2
    return B


def _round_up(x: int, m: int) -> int:
    return ((int(x) + m - 1) // m) * m


def build_lockstep(builders: List[TreeBuilder]) -> List[FlatTree]:
    """Drive several virtual ranks (one process, one stream) through the multi-rank data path.

    The same kernels as the ``p2p`` path run, but as separate launches in phase order
    (all ranks' histograms, then all pushes, then all gains, ...), so no kernel ever
    waits on another: this is how the peer push / owner reduction / record broadcast
    logic is exercised with fewer GPUs than ranks.
    """
    for b in builders:
        b.start()
    while builders[0].n_slots > 0:
        for b in builders:
            b.prepare_level()
        for name in ("mtb_level_plan", "mtb_level_hist", "mtb_level_push", "mtb_level_gain", "mtb_level_finalize"):
            for b in builders:
                b._launch(name)
        nexts = {b.read_level() for b in builders}
        if len(nexts) != 1:
            raise RuntimeError(f"virtual ranks disagree on the next frontier: {sorted(nexts)}")
        for b in builders:
            b.advance_level()
    return [b._download(b.n_nodes) for b in builders]


# ---------------------------------------------------------------------------
# builder cache
# ---------------------------------------------------------------------------
_BUILDERS: dict = {}
_BUILDERS_MAX = 2


def _log_fit(builder, n_global: int, F: int, flat) -> None:
    """One INFO line per fit on rank 0 (``MTB_LOG=INFO``); per-level rows at DEBUG when the builder profiled."""
    from ..utils.log import get_logger

    log = get_logger()
    if not log.isEnabledFor(20):
        return
    st = builder.stats
    log.info("fit: %d rows x %d features, %d classes, %d bins | world %d, reduce %s, tile %s, FT %d, %d slots | "
             "%d levels, %d launches, %d nodes, %d handed off", n_global, F, builder.C, builder.B, st.world, st.reduce,
             st.tile_layout, st.FT, st.m_cap, st.levels, st.launches, flat.n_nodes, st.handed_off)
    for lv, row in enumerate(st.level_ms):
        log.debug("level %d: %s", lv, " ".join(f"{k}={v:.3f}" if isinstance(v, float) else f"{k}={v}" for k, v in row.items()))


def _params_key(params: SplitParams):
    return (params.max_depth, int(params.min_samples_split), params.criterion)


def _get_builder(key, data: DeviceDataset, make):
    """A :class:`TreeBuilder` for ``key``, reused across fits.

    Constructing a builder allocates a few dozen device buffers and (with several ranks) maps peer
    regions; for repeated fits of the same problem shape that is pure overhead (2.2 ms per fit at two
    ranks, more than the device time of a small fit).  ``key`` must be built from values every rank
    agrees on, so that all ranks hit or miss together (a miss runs collectives).  A reused builder
    only needs the new data set: every table it owns is rewritten by the kernels, starting with
    ``init_root``.
    """
    b = _BUILDERS.pop(key, None)
    if b is None:
        b = make()
        while len(_BUILDERS) >= _BUILDERS_MAX:
            _BUILDERS.pop(next(iter(_BUILDERS))).close()
    else:
        b.data = data
        b._bound_tree = None  # re-bind the data pointers
    _BUILDERS[key] = b  # most recently used last
    return b


def clear_caches(*, collective: bool = True) -> None:
    """Drop cached builders, peer regions and workspaces (frees their device memory).

    ``collective=False``: do not synchronise with the peers while closing the regions (used after a
    peer failed)."""
    for b in _BUILDERS.values():
        b.close()
    _BUILDERS.clear()
    _WORKSPACE.clear()
    for reg in _REGIONS.values():
        try:
            reg.close(collective=collective)
        except TypeError:  # regions without a collective close
            reg.close()
    _REGIONS.clear()


# ---------------------------------------------------------------------------
# front ends
# ---------------------------------------------------------------------------
class _Stages:
    """Optional wall-clock stage timer for ``fit`` (``MTB_TIMING=1``): syncs the device at each mark."""

    def __init__(self):
        self.on = os.environ.get("MTB_TIMING") == "1"
        self.t = time.perf_counter()
        self.rows = []

    def mark(self, name: str) -> None:
        if not self.on:
            return
        import torch

        torch.cuda.synchronize()
        now = time.perf_counter()
        self.rows.append((name, (now - self.t) * 1e3))
        self.t = now

    def report(self) -> None:
        if self.on:
            print("[mtb timing] " + "  ".join(f"{n}={ms:.2f}ms" for n, ms in self.rows), flush=True)


def _sample_rows(n: int, cap: Optional[int] = None) -> Optional[np.ndarray]:
    """Indices of an evenly strided sample of ``cap`` rows (``None`` = use every row)."""
    cap = SAMPLE_ROWS if cap is None else cap
    if n <= cap:
        return None
    return (np.arange(cap, dtype=np.int64) * n) // cap  # strictly increasing because n > cap


def _fit_mapper_numpy(X: np.ndarray, max_bins: int, comm=None) -> BinMapper:
    idx = _sample_rows(X.shape[0])
    sample = X if idx is None else X[idx]
    summaries = None
    if comm is not None and comm.size > 1:
        mine = BinMapper.local_summary(sample)
        summaries = comm.allgather_object(mine)  # every rank, own one included, in rank order
    mapper = BinMapper.fit(sample, max_bins=max_bins, summaries=summaries)
    # cover values the sample did not see: the last edge must be the column maximum.  With several
    # ranks the exchange is unconditional (a collective inside a rank-local test would hang when
    # only some shards are large enough to be sampled) and applies when any rank sampled.
    sampled = idx is not None
    col_max = X.max(axis=0) if X.shape[0] else None
    if comm is not None and comm.size > 1:
        got = comm.allgather_object((sampled, col_max))
        sampled = any(g[0] for g in got)
        maxima = [g[1] for g in got if g[1] is not None]
        col_max = np.max(np.stack(maxima), axis=0) if maxima else None
    if sampled and col_max is not None:
        for f, e in enumerate(mapper.edges):
            if len(e) and col_max[f] > e[-1]:
                if len(e) < max_bins:
                    mapper.edges[f] = np.append(e, col_max[f]).astype(e.dtype)
                else:
                    e[-1] = col_max[f]
                mapper.lossless[f] = False
    if sampled and X.dtype.kind in "fiu":
        import torch

        _repair_sampled_edges(torch.from_numpy(X), mapper, max_bins, comm)
    return mapper


def _fit_mapper_device(sample, max_bins: int, comm=None, col_max=None, exchange_col_max: bool = False) -> BinMapper:
    """Bin edges from a device-resident sample: one sort, then order statistics (K1, edge discovery).

    Per feature: all distinct values when there are at most ``max_bins`` (lossless), otherwise
    ``max_bins`` order statistics ("higher" quantiles, so every edge is a data value).  With
    several ranks each contributes its candidates and the merged set is reduced the same way on
    every rank (the cross-rank edge agreement of K1): the candidates of all ranks travel in ONE device
    all-gather (NCCL over NVLink) followed by one read-back; ``col_max`` (full-data maxima, extends the
    last edge when the sample missed the maximum) rides along when ``exchange_col_max`` is set (some rank
    sampled its rows: every rank then contributes its maxima, or -inf when it has no rows).
    """
    import torch

    K, F = int(sample.shape[0]), int(sample.shape[1])
    multi = comm is not None and comm.size > 1
    np_dtype = np.float64 if sample.dtype == torch.float64 else np.float32
    dev = sample.device
    if K == 0:  # a rank without rows still takes part in the agreement, with nothing to contribute
        distinct = torch.zeros(F, dtype=torch.float64, device=dev)
        uniq = torch.full((F, 256), float("inf"), dtype=torch.float64, device=dev)
        cand = torch.zeros((0, F), dtype=torch.float64, device=dev)
    else:
        srt = sample.t().contiguous().sort(dim=1).values  # [F, K], rows sorted
        # sorted rows end in +inf / nan when the input has them: one check on the extremes
        finite = torch.isfinite(srt[:, [0, -1]]).all()
        distinct_t = (srt[:, 1:] != srt[:, :-1]).sum(dim=1) + 1 if K > 1 else torch.ones(F, dtype=torch.int64, device=dev)
        # order statistics ("higher" quantiles): every edge is a data value
        q = np.linspace(0.0, 1.0, max_bins + 1)[1:]
        qidx = torch.from_numpy(np.ceil(q * (K - 1)).astype(np.int64).clip(0, K - 1)).to(dev)
        cand = srt[:, qidx].t().to(torch.float64)  # [max_bins, F]
        distinct = distinct_t.to(torch.float64)
        # features with few distinct values keep all of them (lossless bins): the first 256 distinct values
        # of every row, compacted on the device (rows with more are not used)
        first = torch.ones((F, K), dtype=torch.bool, device=dev)
        if K > 1:
            first[:, 1:] = srt[:, 1:] != srt[:, :-1]
        rank_in_row = first.cumsum(dim=1) - 1  # index among the row's distinct values
        uniq = torch.full((F, 257), float("inf"), dtype=torch.float64, device=dev)
        tgt = torch.where(first & (rank_in_row < 256), rank_in_row, torch.full_like(rank_in_row, 256))
        uniq.scatter_(1, tgt, srt.to(torch.float64))
        uniq = uniq[:, :256].contiguous()
        uniq[:, 0] = srt[:, 0].to(torch.float64)  # (scatter order among duplicates is unspecified only for column 256)
    cmax = torch.full((F,), float("-inf"), dtype=torch.float64, device=dev)
    if col_max is not None:
        cmax = torch.as_tensor(col_max, device=dev).to(torch.float64)
    n_cand = int(cand.shape[0])
    payload = torch.cat([
        torch.tensor([float(K), float(n_cand), 1.0 if K == 0 else 0.0], dtype=torch.float64, device=dev),
        (finite.to(torch.float64).reshape(1) if K else torch.ones(1, dtype=torch.float64, device=dev)),
        distinct, cmax, uniq.reshape(-1),
        torch.cat([cand, torch.zeros((max_bins - n_cand, F), dtype=torch.float64, device=dev)]).reshape(-1),
    ])
    gather = getattr(comm, "allgather_device", None) if multi else None
    gathered = gather(payload) if gather is not None else None
    if gathered is not None:
        table = gathered.cpu().numpy()
    elif multi:
        table = np.stack(comm.allgather_object(payload.cpu().numpy()))
    else:
        table = payload.cpu().numpy()[None]
    if not bool((table[:, 3] > 0).all()):
        raise ValueError("Input X contains NaN or infinity.")
    parts, maxima = [], []
    for row in table:
        k_rows, n_c = int(row[0]), int(row[1])
        o = 4
        d = row[o : o + F].astype(np.int64); o += F
        m = row[o : o + F]; o += F
        u = row[o : o + F * 256].reshape(F, 256).astype(np_dtype); o += F * 256
        c = row[o : o + max_bins * F].reshape(max_bins, F)[:n_c].astype(np_dtype)
        parts.append((d, u, c))
        if k_rows or np.isfinite(m).any():
            maxima.append(m)
    if exchange_col_max or col_max is not None:
        live = [m for m in maxima if np.isfinite(m).all()]
        col_max = np.max(np.stack(live), axis=0).astype(np_dtype) if live else None
    else:
        col_max = None
    return _merge_edge_candidates(parts, F, max_bins, np_dtype, col_max)


def _merge_edge_candidates(parts, F: int, max_bins: int, np_dtype, col_max) -> BinMapper:
    """Reduce every rank's ``(distinct counts, distinct values, order statistics)`` to one edge set.

    Pure function of ``parts`` (in rank order), so every rank derives identical bins."""
    edges, lossless, enumerated = [], np.zeros(F, dtype=bool), np.zeros(F, dtype=bool)
    for f in range(F):
        all_small = all(int(p[0][f]) <= 256 for p in parts)
        if all_small:
            vals = np.unique(np.concatenate([p[1][f][: int(p[0][f])] for p in parts]))
            if len(vals) <= max_bins:
                edges.append(vals.astype(np_dtype))
                lossless[f] = len(parts) == 1 and col_max is None
                enumerated[f] = True
                continue
        merged = np.sort(np.concatenate([p[2][:, f] for p in parts]))
        if len(merged) == 0:  # no rank has any row
            edges.append(np.zeros(0, dtype=np_dtype))
            continue
        if len(parts) > 1 or len(merged) > max_bins:
            pick = np.ceil(np.linspace(0.0, 1.0, max_bins + 1)[1:] * (len(merged) - 1)).astype(np.int64)
            merged = merged[pick]
        edges.append(np.unique(merged).astype(np_dtype))
    mapper = BinMapper(edges=edges, lossless=lossless, max_bins=max_bins, enumerated=enumerated)
    if col_max is not None:
        for f, e in enumerate(mapper.edges):
            if len(e) and col_max[f] > e[-1]:
                if len(e) < max_bins:
                    mapper.edges[f] = np.append(e, col_max[f]).astype(e.dtype)
                else:
                    e[-1] = col_max[f]
    return mapper


def _repair_sampled_edges(X, mapper: BinMapper, max_bins: int, comm=None) -> int:
    """After edges were fitted on a *sample* of the rows: make sure no value of a low-cardinality column was missed.

    A column whose sample showed at most ``max_bins`` distinct values gets one bin per value, and the
    search is then exact for it - but only if the sample saw every value.  A rare value the sample
    missed would silently share the next edge's bin, and candidate thresholds (and printed ones) would
    differ from the reference and from a fit on all rows.  So every such column is checked against all
    rows (one ``bucketize`` pass, chunked); columns with strays get their edges refitted from the
    distinct values of *all* rows of *all* ranks (or, if there turn out to be more than ``max_bins``,
    keep their edges and lose the ``lossless`` mark).  Every decision is taken from all-gathered
    values, so ranks stay in step.  ``X``: this rank's rows as a torch tensor (any device).  Returns the
    number of columns whose edges changed.
    """
    import torch

    cand = [f for f in range(mapper.n_features) if mapper.enumerated is not None and mapper.enumerated[f] and len(mapper.edges[f])]
    multi = comm is not None and comm.size > 1
    stray = np.zeros(len(cand), dtype=bool)
    n = int(X.shape[0])
    step = 1 << 24
    for i, f in enumerate(cand):
        e = torch.from_numpy(np.ascontiguousarray(mapper.edges[f])).to(X.device)
        for r0 in range(0, n, step):
            col = X[r0 : r0 + step, f].contiguous()
            if col.dtype != e.dtype:
                col = col.to(e.dtype)
            idx = torch.bucketize(col, e).clamp_(max=e.numel() - 1)
            if bool((e[idx] != col).any()):
                stray[i] = True
                break
    if multi:
        stray = np.any(np.stack(comm.allgather_object(stray)), axis=0)
    changed = 0
    for i, f in enumerate(cand):
        if not stray[i]:
            continue
        u = torch.unique(X[:, f].contiguous())
        vals = u[: 4 * 256 + 1].cpu().numpy() if u.numel() > 4 * 256 else u.cpu().numpy()  # more than that cannot be enumerated anyway
        if multi:
            vals = np.unique(np.concatenate(comm.allgather_object(vals)))
        vals = vals.astype(mapper.edges[f].dtype, copy=False)
        if len(vals) <= max_bins:
            mapper.edges[f] = vals
            mapper.lossless[f] = not multi
            changed += 1
        else:
            # too many distinct values for one bin each: the sampled edges stay, merged with the column's
            # extremes so that every value still falls into a bin whose upper edge is a data value
            hi = vals[-1]
            e = mapper.edges[f]
            if hi > e[-1]:
                mapper.edges[f] = np.append(e[: max_bins - 1], hi).astype(e.dtype)
                changed += 1
            mapper.lossless[f] = False
            mapper.enumerated[f] = False
    if changed:
        mapper.__dict__.pop("_edge_table", None)
    return changed


def _byte_lut(mapper: BinMapper) -> np.ndarray:
    """``(F, 256)`` uint8 table: byte value -> bin code (first edge >= value, clamped to the last bin)."""
    lut = np.zeros((mapper.n_features, 256), dtype=np.uint8)
    for f, e in enumerate(mapper.edges):
        if len(e):
            lut[f] = np.minimum(np.searchsorted(e, np.arange(256).astype(e.dtype), side="left"), len(e) - 1)
    return lut


_COPY_STREAMS: dict = {}


def _root_mode() -> str:
    """How the first level reads uint8 rows that already are bin codes: ``direct`` (default: planes, labels and
    root histogram in one pass straight from the row-major matrix, ``k_root_direct``), ``cluster`` (the same
    by thread-block clusters and a multicast TMA copy, ``k_root_cluster``: measured slower) or
    ``off`` (retile pass + label copy, histogram by the level kernel).  ``MTB_ROOT`` selects;
    ``MTB_ROOT_CLUSTER=1`` is the older spelling of ``cluster``."""
    if os.environ.get("MTB_ROOT_CLUSTER", "0") == "1":
        return "cluster"
    mode = os.environ.get("MTB_ROOT", "direct")
    return mode if mode in ("direct", "cluster", "off") else "direct"


def _copy_stream(dev):
    import torch

    key = str(dev)
    if key not in _COPY_STREAMS:
        _COPY_STREAMS[key] = torch.cuda.Stream(device=dev)
    return _COPY_STREAMS[key]


def _float_dtype_for(dtype):
    """Floating type that holds every value of ``dtype`` exactly enough to sit on the right side of a
    threshold: integers wider than 16 bits (and anything unknown) go to float64 - float32 would round
    values above 2**24 - narrow integers and half-precision floats to float32.  Fit and predict use
    the same rule (``fit_from_host`` does the same for numpy inputs)."""
    import torch

    narrow = (torch.int8, torch.int16, torch.float16, torch.bfloat16, torch.bool) + tuple(
        t for t in (getattr(torch, "uint16", None),) if t is not None)
    return torch.float32 if dtype in narrow else torch.float64


def _device():
    import torch

    return torch.device("cuda", torch.cuda.current_device())


def _dist_group(comm):
    """torch.distributed is the bootstrap: a >1-rank comm must be the NCCL world."""
    if comm is None or comm.size <= 1:
        return None
    import torch.distributed as dist

    if not dist.is_initialized():
        raise RuntimeError("multi-rank CUDA fits need an initialised torch.distributed (NCCL) process group")
    return None  # default group


def fit_from_host(X: np.ndarray, y_codes: np.ndarray, n_classes: int, params: SplitParams, *, max_bins: int = 256,
                  comm=None, builder_kwargs=None):
    """Fit on a numpy matrix (this rank's shard).  Returns ``(FlatTree, BinMapper)``."""
    import torch

    dev = _device()
    X = np.ascontiguousarray(X)
    if X.dtype not in (np.float32, np.float64, np.uint8):
        X = X.astype(np.float64 if X.dtype.itemsize > 4 or X.dtype.kind in "iu" and X.dtype.itemsize > 2 else np.float32)
    lut = None
    if X.dtype == np.uint8 and max_bins >= 256:
        # one bin per byte value: codes are the values themselves
        mapper = BinMapper(edges=[np.arange(256, dtype=np.uint8) for _ in range(X.shape[1])],
                           lossless=np.ones(X.shape[1], bool), max_bins=256)
    else:
        mapper = _fit_mapper_numpy(X, max_bins, comm)
        if X.dtype == np.uint8:  # fewer bins than byte values were asked for: encode through a value -> code table
            lut = torch.from_numpy(_byte_lut(mapper)).to(dev)
    Xf = X
    n, F = X.shape
    data = DeviceDataset.allocate(n, F, n_classes, dev)
    for r0 in range(0, n, ENCODE_CHUNK_ROWS):
        chunk = torch.from_numpy(Xf[r0 : r0 + ENCODE_CHUNK_ROWS]).to(dev, non_blocking=False)
        if X.dtype == np.uint8:
            data.encode_u8(chunk, lut, r0)
        else:
            data.encode_float(chunk, mapper, r0)
    data.set_labels(torch.from_numpy(np.asarray(y_codes)).to(dev))
    n_global, min_free = n, None
    if comm is not None and comm.size > 1:
        meta = comm.allgather_ints((n, int(torch.cuda.mem_get_info(dev)[0])))
        n_global, min_free = int(meta[:, 0].sum()), int(meta[:, 1].min())
    _dist_group(comm)
    builder = TreeBuilder(data, int(mapper.n_bins.max()), params, n_global=n_global, comm=comm,
                          min_free_bytes=min_free, **(builder_kwargs or {}))
    try:
        flat = builder.build()
    finally:
        builder.close()
        data.release()
    fit_from_host.last_stats = builder.stats
    _log_fit(builder, n_global, F, flat)
    return flat, mapper


def fit_from_codes(codes: np.ndarray, y_codes: np.ndarray, n_classes: int, n_bins: int, params: SplitParams, *,
                   comm=None, builder_kwargs=None) -> FlatTree:
    """Grow the tree from bin codes that were formed on the host (this rank's shard, ``uint8``).

    The path of object / string ("categorical") columns: the reference splits them ordinally with
    Python's ``<=`` on the sorted distinct values (``decision_tree.py:73-77``), so the host mapper turns
    every column into ordinal codes in sorted-unique order - strings cannot go to the device, their
    codes can - and the level-wise kernels take it from there exactly as for numeric data.  ``n_bins``:
    the largest number of bins of any column (agreed on across ranks).
    """
    import torch

    dev = _device()
    codes = np.ascontiguousarray(codes, dtype=np.uint8)
    n, F = codes.shape
    data = DeviceDataset.allocate(n, F, n_classes, dev)
    for r0 in range(0, n, ENCODE_CHUNK_ROWS):
        data.encode_u8(torch.from_numpy(codes[r0 : r0 + ENCODE_CHUNK_ROWS]).to(dev), None, r0)
    data.set_labels(torch.from_numpy(np.asarray(y_codes)).to(dev))
    n_global, min_free = n, None
    if comm is not None and comm.size > 1:
        meta = comm.allgather_ints((n, int(torch.cuda.mem_get_info(dev)[0])))
        n_global, min_free = int(meta[:, 0].sum()), int(meta[:, 1].min())
    _dist_group(comm)
    builder = TreeBuilder(data, int(n_bins), params, n_global=n_global, comm=comm, min_free_bytes=min_free,
                          **(builder_kwargs or {}))
    try:
        flat = builder.build()
    finally:
        builder.close()
        data.release()
    fit_from_host.last_stats = builder.stats
    _log_fit(builder, n_global, F, flat)
    return flat


def fit_from_torch(X, y, params: SplitParams, *, max_bins: int = 256, comm=None, replicated: bool = False,
                   builder_kwargs=None):
    """Fit on torch tensors (CUDA, or pinned/pageable host).  Returns ``(FlatTree, BinMapper, classes)``.

    Host tensors are streamed to the device in chunks on the current stream, so a
    pinned ``X`` is consumed at PCIe speed with no extra host copy.
    """
    import torch

    dev = _device()
    st = _Stages()
    if replicated and comm is not None and comm.size > 1:
        n = X.shape[0]
        lo, hi = (n * comm.rank) // comm.size, (n * (comm.rank + 1)) // comm.size
        X, y = X[lo:hi], y[lo:hi]
    n, F = int(X.shape[0]), int(X.shape[1])
    if X.dtype not in (torch.float32, torch.float64, torch.uint8):
        X = X.to(_float_dtype_for(X.dtype))

    # labels -> class codes (classes_ are agreed on globally)
    y_dev = y.to(dev, non_blocking=True)
    if y_dev.dtype.is_floating_point or y_dev.dtype == torch.bool:
        raise TypeError("class labels must be non-negative integers")
    multi = comm is not None and comm.size > 1
    if y_dev.numel() == 0 and not multi:
        raise ValueError("Found array with 0 sample(s)")
    # One read-back per fit: label range, and with several ranks every shard's size and free memory (the
    # one host-visible exchange of a fit; a rank of a sharded fit may hold no rows and still joins it).
    # Free memory is only needed when a builder has to be constructed; -1 = "not asked".
    free_b = -1 if _BUILDERS else int(torch.cuda.mem_get_info(dev)[0])
    if y_dev.numel():
        lo_hi = torch.stack(torch.aminmax(y_dev)).to(torch.int64)
    else:
        lo_hi = torch.tensor([0, -1], dtype=torch.int64, device=dev)
    gather = getattr(comm, "allgather_device", None) if multi else None
    gathered = None
    if gather is not None:
        gathered = gather(torch.cat([lo_hi, torch.tensor([n, free_b], dtype=torch.int64, device=dev)]))
    if gathered is not None:
        meta = gathered.cpu().numpy()  # (size, 4): y_min, y_max, n, free
    else:
        y_lo, y_hi = (int(v) for v in lo_hi.tolist())
        meta = comm.allgather_ints((y_lo, y_hi, n, free_b)) if multi else np.asarray([[y_lo, y_hi, n, free_b]], dtype=np.int64)
    have_rows = meta[:, 2] > 0
    y_min = int(meta[have_rows, 0].min()) if have_rows.any() else 0
    y_max = int(meta[have_rows, 1].max()) if have_rows.any() else -1
    if y_min < 0:  # decided from the gathered values: every rank raises, nobody is left waiting
        raise ValueError("class labels must be non-negative integers")
    st.mark("labels")
    lazy_classes = False
    n_global = int(meta[:, 2].sum())
    min_free = None if not multi else int(meta[:, 3].min())
    meta = meta[:, 1:]  # (y_max, n, free) as before
    if y_max < MAX_CLASSES:
        # label values are used as class codes; the labels that actually occur are read off the
        # root histogram afterwards (saves a counting pass over y)
        lazy_classes = True
        classes = np.arange(y_max + 1, dtype=np.int64)
        y_codes = y_dev
    else:
        if y_max < (1 << 16):  # small label range: one counting pass instead of a sort
            present = torch.bincount(y_dev.view(-1).to(torch.int64), minlength=y_max + 1)
            classes_t = torch.nonzero(present, as_tuple=False).view(-1)
        else:
            classes_t = torch.unique(y_dev)
        classes = classes_t.cpu().numpy().astype(np.int64)
        if comm is not None and comm.size > 1:
            classes = np.unique(np.concatenate(comm.allgather_object(classes)))
            classes_t = torch.from_numpy(classes).to(dev)
        if len(classes) > MAX_CLASSES:
            raise ValueError(f"the CUDA engine supports at most {MAX_CLASSES} classes")
        y_codes = torch.bucketize(y_dev.to(torch.int64), classes_t.to(torch.int64))

    lut = None
    if X.dtype == torch.uint8 and max_bins >= 256:
        mapper = BinMapper(edges=[np.arange(256, dtype=np.uint8) for _ in range(F)], lossless=np.ones(F, bool), max_bins=256)
    else:
        idx = _sample_rows(n)
        sample = X if idx is None else X[torch.from_numpy(idx).to(X.device)]
        if X.dtype == torch.uint8:
            sample = sample.to(torch.float32)  # edge discovery works on floats; the rows stay bytes
        # Some rank sampled its rows (the shard sizes come from the gather above, so the decision is the same
        # everywhere).  No pass over all rows for the column maxima: a value above the last sampled edge is
        # clamped into the last bin by the encoders, the last edge is never a threshold (its right side is
        # empty), and columns that are enumerated value by value are checked against all rows below.
        any_sampled = idx is not None or (multi and int(meta[:, 1].max()) > SAMPLE_ROWS)
        mapper = _fit_mapper_device(sample.to(dev, non_blocking=True), max_bins, comm)
        if any_sampled:
            _repair_sampled_edges(X, mapper, max_bins, comm)  # (uint8 rows compare against float edges)
        if X.dtype == torch.uint8:
            lut = torch.from_numpy(_byte_lut(mapper)).to(dev)

    st.mark("meta+mapper")
    data = DeviceDataset.allocate(n, F, len(classes), dev)
    st.mark("alloc")
    # uint8 rows that already are bin codes, resident on this GPU, whole 16-feature groups: the builder's
    # first launch can read them in the caller's row-major layout (planes + labels + root histogram in one
    # pass by thread-block clusters, k_root_cluster) instead of a retile pass and a label copy here
    root_input = None
    mode = _root_mode()
    if (X.is_cuda and X.device == dev and X.dtype == torch.uint8 and lut is None and X.stride(1) == 1
            and X.stride(0) == F and y_codes.dtype == torch.uint8 and y_codes.is_contiguous() and y_codes.device == dev
            and (F in (32, 64, 128) if mode == "cluster" else (mode == "direct" and F % 16 == 0 and F > 0))):
        root_input = (X, y_codes)
    if root_input is not None:
        pass
    elif X.is_cuda:
        (data.encode_u8(X, lut, 0) if X.dtype == torch.uint8 else data.encode_float(X, mapper, 0))
    elif dev.type != "cuda":  # driver tests without a GPU: the "device" is the host
        for r0 in range(0, n, ENCODE_CHUNK_ROWS):
            chunk = X[r0 : r0 + ENCODE_CHUNK_ROWS].to(dev, non_blocking=True)
            (data.encode_u8(chunk, lut, r0) if X.dtype == torch.uint8 else data.encode_float(chunk, mapper, r0))
    else:
        # host rows: the copies run on their own stream, so chunk k+1 is on the wire while chunk k is encoded
        main = torch.cuda.current_stream()
        side = _copy_stream(dev)
        side.wait_stream(main)
        for r0 in range(0, n, ENCODE_CHUNK_ROWS):
            with torch.cuda.stream(side):
                chunk = X[r0 : r0 + ENCODE_CHUNK_ROWS].to(dev, non_blocking=True)
                arrived = side.record_event()
            main.wait_event(arrived)
            chunk.record_stream(main)
            (data.encode_u8(chunk, lut, r0) if X.dtype == torch.uint8 else data.encode_float(chunk, mapper, r0))
    if root_input is None:
        data.set_labels(y_codes)
    st.mark("encode")
    _dist_group(comm)
    n_bins = int(mapper.n_bins.max())
    kw = dict(builder_kwargs or {})

    def make():
        free = min_free
        if multi and free is not None and free < 0:
            # some rank did not look at its free memory (it expected a cached builder): ask again, all together
            free = int(comm.allgather_ints((int(torch.cuda.mem_get_info(dev)[0]),)).min())
        return TreeBuilder(data, n_bins, params, n_global=n_global, comm=comm, min_free_bytes=free, **kw)

    # the key holds only values every rank sees identically (shard sizes come from the gather above)
    key = ("torch", str(dev), tuple(int(v) for v in meta[:, 1]), F, len(classes), n_bins, _params_key(params),
           0 if comm is None else (comm.size, comm.rank), tuple(sorted(kw.items())), os.environ.get("MTB_TILE"),
           os.environ.get("MTB_ASYNC"), os.environ.get("MTB_SMALL_ROWS"))
    cacheable = not kw.get("profile") and "peers" not in kw
    builder = _get_builder(key, data, make) if cacheable else make()
    builder.root_input = root_input if builder._coop else None
    if root_input is not None and builder.root_input is None:  # a launch decomposition without the root pass
        data.encode_u8(X, lut, 0)
        data.set_labels(y_codes)
    st.mark("builder_init")
    try:
        flat = builder.build()
    finally:
        if not cacheable:
            builder.close()
        data.release()
    st.mark("build")
    st.report()
    fit_from_torch.last_stats = builder.stats
    flat.max_shard_rows = int(meta[:, 1].max())  # agreed on by every rank (for decisions that involve collectives)
    _log_fit(builder, n_global, F, flat)
    if lazy_classes:
        # classes_ = labels that occur: read them off the root's class counts
        seen = np.flatnonzero(flat.counts[0] > 0)
        classes = seen.astype(np.int64)
        flat.counts = np.ascontiguousarray(flat.counts[:, seen])
    return flat, mapper, classes


# ---------------------------------------------------------------------------
# predict
# ---------------------------------------------------------------------------
class _PackedTree:
    """Device copy of a FlatTree for K4 (cached on the FlatTree object)."""

    def __init__(self, flat: FlatTree, dev):
        import torch

        n = flat.n_nodes
        self.n_nodes = n
        feature = torch.from_numpy(flat.feature.astype(np.int32)).to(dev)
        left = torch.from_numpy(flat.left.astype(np.int32)).to(dev)
        bad = (flat.left >= 0) & (flat.right != flat.left + 1)
        if bad.any():
            raise ValueError("predict kernel needs consecutive child ids (right == left + 1)")
        labels = flat.expanded_counts().argmax(axis=1).astype(np.int32)
        label = torch.from_numpy(labels).to(dev)
        self.packed = torch.empty(n * 4, dtype=torch.int32, device=dev)
        _call("mtb_pack_nodes", ctypes.c_void_p(_p(feature)), ctypes.c_void_p(_p(left)), ctypes.c_void_p(_p(label)),
              ctypes.c_int64(n), ctypes.c_void_p(_p(self.packed)), _stream_ptr())
        thr = flat.numeric_thresholds(np.float64)
        thr = np.where(np.isnan(thr), 0.0, thr)
        self.thr = {
            0: torch.from_numpy(thr.astype(np.float32)).to(dev),
            1: torch.from_numpy(np.floor(thr).astype(np.int32)).to(dev),
            2: torch.from_numpy(thr).to(dev),
        }
        # a float32 threshold must not move across any float32 input value: round down
        t32 = thr.astype(np.float32)
        t32 = np.where(t32.astype(np.float64) > thr, np.nextafter(t32, np.float32(-np.inf)), t32)
        self.thr[0] = torch.from_numpy(t32).to(dev)


def predict_leaf_ids(flat: FlatTree, X, *, want: str = "leaf"):
    """Leaf id per row.  ``X``: numpy array or torch tensor (host or CUDA)."""
    import torch

    dev = _device()
    cache = getattr(flat, "_packed", None)
    if cache is None or cache.packed.device != dev:
        cache = _PackedTree(flat, dev)
        flat._packed = cache
    is_np = isinstance(X, np.ndarray)
    Xt = torch.from_numpy(np.ascontiguousarray(X)) if is_np else X
    if Xt.dtype not in (torch.float32, torch.float64, torch.uint8):
        Xt = Xt.to(_float_dtype_for(Xt.dtype))
    dtype = {torch.float32: 0, torch.uint8: 1, torch.float64: 2}[Xt.dtype]
    n = int(Xt.shape[0])
    out = torch.empty(n, dtype=torch.int32, device=dev)
    step = n if Xt.is_cuda else ENCODE_CHUNK_ROWS
    for r0 in range(0, n, max(step, 1)):
        chunk = Xt[r0 : r0 + step]
        if not chunk.is_cuda:
            chunk = chunk.to(dev, non_blocking=True)
        if chunk.stride(1) != 1:
            chunk = chunk.contiguous()
        view = out[r0 : r0 + chunk.shape[0]]
        _call(
            "mtb_predict",
            ctypes.c_void_p(_p(chunk)), ctypes.c_int(dtype), ctypes.c_int64(chunk.shape[0]),
            ctypes.c_int64(chunk.shape[1]), ctypes.c_int64(chunk.stride(0)),
            ctypes.c_void_p(_p(cache.packed)), ctypes.c_void_p(_p(cache.thr[dtype])), ctypes.c_int64(cache.n_nodes),
            ctypes.c_void_p(_p(view) if want == "leaf" else 0), ctypes.c_void_p(_p(view) if want == "label" else 0),
            _stream_ptr(),
        )
    if is_np:
        return out.cpu().numpy().astype(np.int64)
    return out.to(torch.int64)
