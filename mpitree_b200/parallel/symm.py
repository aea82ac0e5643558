"""Symmetric peer memory: one device region per rank, mapped into every peer.

This is the data plane that replaces the reference's pickled ``comm.allgather``
(``mpitree/tree/decision_tree.py:456-466``): each rank owns a region holding its
histogram staging area, its best-split record mailbox and its barrier flags; every
peer maps that region (CUDA IPC over NVLink/NVSwitch) and kernels write into it
directly.  ``torch.distributed`` is used once, to exchange the 64-byte IPC handles.

Two implementations with the same attributes (``world``, ``rank``, ``ptrs``):

* :class:`IpcRegion`   — one process per GPU (the production layout);
* :class:`LocalRegion` — several *virtual* ranks inside one process / one GPU, used
  to run the multi-rank data path in lock-step where fewer GPUs than ranks exist
  (tests, development).
"""

from __future__ import annotations

import ctypes
from typing import List

__all__ = ["IpcRegion", "LocalRegion", "RawBuffer"]


class RawBuffer:
    """A ``cudaMalloc`` allocation owned by this process (``__cuda_array_interface__`` view)."""

    def __init__(self, nbytes: int):
        from ..ops import _lib as L

        self._L = L
        self.nbytes = int(nbytes)
        out = ctypes.c_void_p()
        L.check(L.load().mtb_device_malloc(ctypes.c_int64(self.nbytes), ctypes.byref(out)), "mtb_device_malloc")
        self.ptr = int(out.value)

    @property
    def __cuda_array_interface__(self):
        return {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 3, "strides": None}

    def as_tensor(self):
        import torch

        return torch.as_tensor(self, device="cuda")

    def zero_(self, stream_ptr) -> None:
        L = self._L
        L.check(L.load().mtb_memset_async(ctypes.c_void_p(self.ptr), 0, ctypes.c_int64(self.nbytes), stream_ptr),
                "mtb_memset_async")

    def free(self) -> None:
        if self.ptr:
            self._L.load().mtb_device_free(ctypes.c_void_p(self.ptr))
            self.ptr = 0

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.free()
        except Exception:
            pass


class IpcRegion:
    """``nbytes`` on every rank of ``comm``; ``ptrs[q]`` is rank q's region as seen from here."""

    def __init__(self, nbytes: int, comm):
        import torch

        from ..ops import _lib as L

        lib = L.load()
        self.world, self.rank = comm.size, comm.rank
        self.nbytes = int(nbytes)
        self.buf = RawBuffer(nbytes)
        self.buf.zero_(ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        hsize = lib.mtb_ipc_handle_size()
        handle = (ctypes.c_ubyte * hsize)()
        L.check(lib.mtb_ipc_get_handle(ctypes.c_void_p(self.buf.ptr), handle), "mtb_ipc_get_handle")
        info = comm.allgather_object((bytes(handle), torch.cuda.current_device()))
        self.ptrs: List[int] = []
        self._opened: List[int] = []
        for q, (h, dev) in enumerate(info):
            if q == self.rank:
                self.ptrs.append(self.buf.ptr)
                continue
            raw = (ctypes.c_ubyte * hsize).from_buffer_copy(h)
            out = ctypes.c_void_p()
            L.check(lib.mtb_ipc_open_handle(raw, ctypes.byref(out)), f"mtb_ipc_open_handle(rank {q})")
            self.ptrs.append(int(out.value))
            self._opened.append(int(out.value))
        comm.barrier()  # everyone has mapped everyone before any kernel writes to a peer
        self._comm = comm

    def close(self, *, collective: bool = True) -> None:
        """Unmap the peers and free the own region.  ``collective=False`` skips the barrier: the way out
        after a peer failed (nobody will answer, and nothing that still runs writes here)."""
        from ..ops import _lib as L

        lib = L.load()
        try:
            import torch

            torch.cuda.synchronize()
            if collective:
                self._comm.barrier()  # no peer may still be writing into a region that goes away
        except Exception:  # pragma: no cover
            pass
        for p in self._opened:
            lib.mtb_ipc_close(ctypes.c_void_p(p))
        self._opened = []
        self.buf.free()


class LocalRegion:
    """Regions of ``world`` virtual ranks living in this process (plain device pointers)."""

    def __init__(self, nbytes: int, world: int):
        import torch

        self.world = world
        self.nbytes = int(nbytes)
        self.tensors = [torch.zeros(self.nbytes, dtype=torch.uint8, device="cuda") for _ in range(world)]
        self.ptrs = [int(t.data_ptr()) for t in self.tensors]

    def view(self, rank: int) -> "_LocalView":
        return _LocalView(self, rank)


class _LocalView:
    def __init__(self, region: LocalRegion, rank: int):
        self.world, self.rank, self.ptrs, self.nbytes = region.world, rank, region.ptrs, region.nbytes
        self._region = region

    def close(self) -> None:
        return None


def multicast_supported(device: int = 0) -> bool:
    """Does the driver expose NVLS multicast objects (``cuMulticastCreate``) on ``device``?

    Round 1 ships the always-available peer (P2P load/store) data plane; the probe is what a
    ``multimem.red`` / ``multimem.ld_reduce`` fast path would key on.  Returns False whenever
    the attribute cannot be queried (no driver, no cuda-python, CPU box).
    """
    try:
        try:
            from cuda.bindings import driver as drv  # cuda-python >= 12.8
        except ImportError:
            from cuda import cuda as drv

        (err,) = drv.cuInit(0)
        if int(err) != 0:
            return False
        attr = getattr(drv.CUdevice_attribute, "CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED", None)
        if attr is None:
            return False
        err, dev = drv.cuDeviceGet(device)
        if int(err) != 0:
            return False
        err, val = drv.cuDeviceGetAttribute(attr, dev)
        return int(err) == 0 and int(val) == 1
    except Exception:  # noqa: BLE001
        return False
