"""Symmetric memory with an NVSwitch multicast alias (NVLS) — STAGED, not yet run on hardware.

:class:`NvlsRegion` has the attributes of :class:`mpitree_b200.parallel.symm.IpcRegion`
(``world``, ``rank``, ``nbytes``, ``ptrs[q]`` = rank q's region mapped here) plus ``mc_ptr``: one
virtual address that names *the same offset in every rank's region at once*.  Kernels use it with

* ``multimem.ld_reduce.add`` — one load returns the sum over all ranks, computed in the switch
  (SASS ``LDGMC.ADD``): the reduce-scatter half of a histogram all-reduce without a staging copy;
* ``multimem.st``            — one store lands in every rank's region: the all-gather half.

``csrc/comm.cu`` (``k_allreduce_nvls``) is the stand-alone collective built on both.  Integer
``ld_reduce`` has no vector form (ptxas: "Vector qualifier is not allowed"), so the kernel adds
``u64`` words: two packed u32 counts per lane per instruction, exact as long as no global count
reaches 2**32 (already an invariant of the u32 histograms).

Construction (every rank, same ``nbytes``):

1. ``cuMemCreate`` a physical allocation that can be exported as a POSIX file descriptor;
2. rank 0 ``cuMulticastCreate``-s the multicast object; descriptors of all allocations and of the
   multicast object are all-gathered with :func:`mpitree_b200.parallel.fdpass.allgather_fds`;
3. every rank adds its device to the multicast object, *then* (barrier) binds its allocation;
4. every allocation and the multicast object are mapped into this process's address space.

Memory allocated with ``cudaMalloc`` (what :class:`IpcRegion` shares) cannot be bound to a
multicast object, which is why this is a separate region type.  The reference has no counterpart
(``mpitree/tree/decision_tree.py:456-466`` gathers pickles through MPI).
"""

from __future__ import annotations

import os
from typing import List

__all__ = ["NvlsRegion", "nvls_available"]


def _driver():
    try:
        from cuda.bindings import driver as drv  # cuda-python >= 12.8
    except ImportError:  # pragma: no cover - older layout
        from cuda import cuda as drv
    return drv


def _ck(res, what: str):
    """cuda-python returns ``(CUresult, *values)``; raise with the call's name on failure."""
    err = res[0]
    if int(err) != 0:
        drv = _driver()
        try:
            _, name = drv.cuGetErrorName(err)
            name = name.decode() if isinstance(name, bytes) else str(name)
        except Exception:  # noqa: BLE001
            name = str(err)
        raise RuntimeError(f"{what} failed: {name}")
    rest = res[1:]
    return rest[0] if len(rest) == 1 else rest


def nvls_available(device: int = 0) -> bool:
    from .symm import multicast_supported

    return multicast_supported(device)


def _round_up(x: int, m: int) -> int:
    return ((int(x) + m - 1) // m) * m


class _DeviceSpan:
    """``__cuda_array_interface__`` view of raw device bytes (keeps nothing alive)."""

    def __init__(self, ptr: int, nbytes: int):
        self.ptr, self.nbytes = int(ptr), int(nbytes)

    @property
    def __cuda_array_interface__(self):
        return {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 3, "strides": None}


class NvlsRegion:
    """``nbytes`` of zeroed device memory per rank: peer-mapped (``ptrs``) and multicast-mapped (``mc_ptr``)."""

    def __init__(self, nbytes: int, comm):
        import torch

        from .fdpass import allgather_fds

        drv = self._drv = _driver()
        self.world, self.rank = comm.size, comm.rank
        self._comm = comm
        torch.cuda.current_stream().synchronize()  # also makes sure torch's primary context is current
        dev_index = torch.cuda.current_device()
        _ck(drv.cuInit(0), "cuInit")
        self._dev = _ck(drv.cuDeviceGet(dev_index), "cuDeviceGet")
        fd_type = drv.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR

        prop = drv.CUmemAllocationProp()
        prop.type = drv.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
        prop.location.type = drv.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        prop.location.id = dev_index
        prop.requestedHandleTypes = fd_type
        gran = _ck(
            drv.cuMemGetAllocationGranularity(prop, drv.CUmemAllocationGranularity_flags.CU_MEM_ALLOC_GRANULARITY_RECOMMENDED),
            "cuMemGetAllocationGranularity",
        )
        mcprop = drv.CUmulticastObjectProp()
        mcprop.numDevices = self.world
        mcprop.handleTypes = int(fd_type)
        mcprop.flags = 0
        mcprop.size = _round_up(max(nbytes, 1), int(gran))
        mc_gran = _ck(
            drv.cuMulticastGetGranularity(mcprop, drv.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED),
            "cuMulticastGetGranularity",
        )
        # one size for everything: a multiple of both granularities, identical on every rank
        step = max(int(gran), int(mc_gran))
        sizes = comm.allgather_object(_round_up(max(nbytes, 1), step))
        self.size = int(max(sizes))
        mcprop.size = self.size
        self.nbytes = int(nbytes)

        self._handles: List[object] = [None] * self.world
        self._vas: List[int] = [0] * self.world
        self._mc_handle = None
        self.mc_ptr = 0
        self._bound = False
        self.ptrs: List[int] = [0] * self.world

        # 1-2: allocate, create the multicast object on rank 0, exchange descriptors
        own = _ck(drv.cuMemCreate(self.size, prop, 0), "cuMemCreate")
        self._handles[self.rank] = own
        send = [int(_ck(drv.cuMemExportToShareableHandle(own, fd_type, 0), "cuMemExportToShareableHandle(mem)"))]
        if self.rank == 0:
            self._mc_handle = _ck(drv.cuMulticastCreate(mcprop), "cuMulticastCreate")
            send.append(int(_ck(drv.cuMemExportToShareableHandle(self._mc_handle, fd_type, 0),
                                "cuMemExportToShareableHandle(multicast)")))
        got = allgather_fds(comm, send, "nvls")
        for fd in send:
            os.close(fd)
        try:
            for q in range(self.world):
                if q != self.rank:
                    self._handles[q] = _ck(drv.cuMemImportFromShareableHandle(got[q][0], fd_type),
                                           f"cuMemImportFromShareableHandle(rank {q})")
            if self.rank != 0:
                self._mc_handle = _ck(drv.cuMemImportFromShareableHandle(got[0][1], fd_type),
                                      "cuMemImportFromShareableHandle(multicast)")
        finally:
            for fds in got:
                for fd in fds:
                    os.close(fd)

        # 3: all devices join before any memory is bound
        _ck(drv.cuMulticastAddDevice(self._mc_handle, self._dev), "cuMulticastAddDevice")
        comm.barrier()
        _ck(drv.cuMulticastBindMem(self._mc_handle, 0, own, 0, self.size, 0), "cuMulticastBindMem")
        self._bound = True
        comm.barrier()

        # 4: map everything read-write for this device
        access = drv.CUmemAccessDesc()
        access.location.type = drv.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        access.location.id = dev_index
        access.flags = drv.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE

        def map_handle(handle, what: str) -> int:
            va = _ck(drv.cuMemAddressReserve(self.size, step, 0, 0), f"cuMemAddressReserve({what})")
            _ck(drv.cuMemMap(va, self.size, 0, handle, 0), f"cuMemMap({what})")
            _ck(drv.cuMemSetAccess(va, self.size, [access], 1), f"cuMemSetAccess({what})")
            return int(va)

        for q in range(self.world):
            self._vas[q] = map_handle(self._handles[q], f"rank {q}")
        self.ptrs = list(self._vas)
        self.mc_ptr = map_handle(self._mc_handle, "multicast")
        _ck(drv.cuMemsetD8(self.ptrs[self.rank], 0, self.size), "cuMemsetD8")
        torch.cuda.synchronize()
        comm.barrier()  # every region is zeroed and mapped before any kernel touches a peer

    # -- views -----------------------------------------------------------------------
    def as_tensor(self):
        """This rank's region as a uint8 CUDA tensor (no copy; valid until :meth:`close`)."""
        import torch

        return torch.as_tensor(_DeviceSpan(self.ptrs[self.rank], self.nbytes), device="cuda")

    def close(self) -> None:
        drv = self._drv
        try:
            import torch

            torch.cuda.synchronize()
            self._comm.barrier()  # no peer may still address a region that goes away
        except Exception:  # pragma: no cover  # noqa: BLE001
            pass
        for va in [self.mc_ptr] + self._vas:
            if va:
                drv.cuMemUnmap(va, self.size)
                drv.cuMemAddressFree(va, self.size)
        self.mc_ptr, self._vas, self.ptrs = 0, [0] * self.world, [0] * self.world
        if self._bound:
            drv.cuMulticastUnbind(self._mc_handle, self._dev, 0, self.size)
            self._bound = False
        for h in self._handles + [self._mc_handle]:
            if h is not None:
                drv.cuMemRelease(h)
        self._handles, self._mc_handle = [None] * self.world, None
