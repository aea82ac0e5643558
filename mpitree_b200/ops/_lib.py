"""Build and load ``libmpitree_b200.so`` (the sm_100a native library) through ctypes.

The library is built **in-tree** (``mpitree_b200/lib/``) with plain ``nvcc`` so the
``.so`` travels with the source snapshot; there is no JIT cache.  Loading fails loudly:
on a machine with a GPU a missing library is an error, never a silent fallback.
"""

from __future__ import annotations

import ctypes
import hashlib
import os
import pathlib
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import Optional

PKG = pathlib.Path(__file__).resolve().parents[1]
CSRC = PKG / "csrc"
LIBDIR = PKG / "lib"
LIBNAME = "libmpitree_b200.so"
INCLUDE = PKG.parent / "include"

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-O3",
    "-std=c++17",
    "-Xcompiler",
    "-fPIC",
]

_lib: Optional[ctypes.CDLL] = None


class MtbCtx(ctypes.Structure):
    """Mirror of ``struct MtbCtx`` (include/mpitree_b200.h): all fields are int64."""

    _names = (
        "n_cap n_local F G B C FT max_depth min_samples_split criterion dot_mode world rank m_cap node_cap "
        "codes0 codes1 labels0 labels1 cur hist0 hist1 hcur "
        "slot_node slot_begin slot_end slot_mode slot_parent slot_sibling "
        "nslot_node nslot_begin nslot_end nslot_mode nslot_parent nslot_sibling nslot_src n_slots_host "
        "fb_cost fb_bin fb_left node_cnt rec_feature rec_bin part_flags cursors "
        "t_feature t_bin t_left t_parent t_depth t_gain t_counts scalars unit_prefix depth"
    ).split()
    _fields_ = (
        [(n, ctypes.c_int64) for n in _names]
        + [("peer_hist", ctypes.c_int64 * 16), ("peer_rec", ctypes.c_int64 * 16), ("peer_flags", ctypes.c_int64 * 16)]
        + [(n, ctypes.c_int64) for n in "epoch plogp plogp_n h_cap fpo timers klogk klogk_n".split()]
        + [("rstore", ctypes.c_int64 * 2)]
        + [(n, ctypes.c_int64) for n in
           "small_rows ho_cap ho_node ho_begin ho_end ho_side part_child mask_words zero_next tile_layout".split()]
        + [("mc_hist", ctypes.c_int64 * 2), ("barrier_timeout_ns", ctypes.c_int64), ("root_hist_done", ctypes.c_int64)]
    )


class MtbSubtreeArgs(ctypes.Structure):
    """Mirror of ``struct MtbSubtreeArgs`` (mpitree_b200/csrc/subtree.cu): all fields are int64."""

    _names = (
        "codes0 codes1 labels0 labels1 n_cap F G B C FT T max_depth min_samples_split criterion dot_mode "
        "plogp plogp_n klogk klogk_n n_items it_side it_begin it_end it_depth it_off it_counts "
        "o_feature o_bin o_left o_depth o_gain o_counts o_nn stack counter"
    ).split()
    _fields_ = [(n, ctypes.c_int64) for n in _names]


def sources():
    return sorted(CSRC.glob("*.cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def build(force: bool = False, verbose: bool = False) -> pathlib.Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link the shared library (idempotent)."""
    LIBDIR.mkdir(exist_ok=True)
    out = LIBDIR / LIBNAME
    stamp = LIBDIR / "build.sha256"
    digest = _digest()
    if not force and out.exists() and stamp.exists() and stamp.read_text().strip() == digest:
        return out
    nvcc = nvcc_path()
    objdir = LIBDIR / "obj"
    objdir.mkdir(exist_ok=True)

    def compile_one(src: pathlib.Path) -> pathlib.Path:
        obj = objdir / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-c", str(src), "-o", str(obj)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
        if verbose and res.stderr:
            print(res.stderr, file=sys.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as pool:
        objs = list(pool.map(compile_one, sources()))
    cmd = [nvcc, "-shared", "-o", str(out), *map(str, objs)]  # static cudart (nvcc default)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    stamp.write_text(digest)
    return out


def lib_path() -> pathlib.Path:
    override = os.environ.get("MTB_LIB")  # development: an experimental build of the same sources
    return pathlib.Path(override) if override else LIBDIR / LIBNAME


def load() -> ctypes.CDLL:
    """Load the native library (building it first if the sources changed and nvcc exists)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        try:
            build()
        except Exception as exc:  # noqa: BLE001
            raise RuntimeError(
                f"{LIBNAME} is missing and could not be built ({exc}); run `python -c 'import __graft_entry__ as g; g.build()'`"
            ) from exc
    # torch ships the CUDA runtime; make sure it is loaded first so symbols resolve to one copy
    import torch  # noqa: F401

    lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
    lib.mtb_last_error.restype = ctypes.c_char_p
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().mtb_last_error().decode(errors="replace")
        raise RuntimeError(f"native call {what} failed ({rc}): {msg}")
