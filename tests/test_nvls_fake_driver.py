"""``NvlsRegion`` (staged) against a fake driver: call order, handle bookkeeping, descriptor exchange.

The real thing needs NVSwitch hardware.  What can be pinned without it is the control flow: three
thread-ranks build a region through a stand-in for ``cuda.bindings.driver`` that enforces the
driver's documented ordering rules (every device added to the multicast object before any memory
is bound, nothing mapped twice, everything unmapped and released at the end) while real file
descriptors travel through :mod:`mpitree_b200.parallel.fdpass`.
"""
import enum
import itertools
import os
import tempfile
import threading
import types

import pytest

from mpitree_b200.parallel.comm import run_threaded


class _E(enum.IntEnum):
    X = 1


class FakeDriver:
    """Process-wide fake: handles are ints, a shareable handle is a temp file holding the handle number."""

    GRAN = 1 << 21

    def __init__(self):
        self.lock = threading.Lock()
        self.ids = itertools.count(100)
        self.mem, self.mc, self.maps, self.vas, self.log = {}, {}, {}, set(), []
        ns = types.SimpleNamespace
        self.CUmemAllocationHandleType = ns(CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR=_E.X)
        self.CUmemAllocationType = ns(CU_MEM_ALLOCATION_TYPE_PINNED=1)
        self.CUmemLocationType = ns(CU_MEM_LOCATION_TYPE_DEVICE=1)
        self.CUmemAllocationGranularity_flags = ns(CU_MEM_ALLOC_GRANULARITY_RECOMMENDED=1)
        self.CUmulticastGranularity_flags = ns(CU_MULTICAST_GRANULARITY_RECOMMENDED=1)
        self.CUmemAccess_flags = ns(CU_MEM_ACCESS_FLAGS_PROT_READWRITE=3)

    def _struct(self):
        return types.SimpleNamespace(location=types.SimpleNamespace(), allocFlags=types.SimpleNamespace())

    CUmemAllocationProp = CUmulticastObjectProp = CUmemAccessDesc = property(lambda self: self._struct)

    def cuInit(self, flags):
        return (0,)

    def cuDeviceGet(self, index):
        return 0, ("dev", threading.get_ident())  # one "device" per thread-rank

    def cuMemGetAllocationGranularity(self, prop, option):
        return 0, self.GRAN

    def cuMulticastGetGranularity(self, prop, option):
        assert prop.numDevices >= 1 and prop.size % self.GRAN == 0
        return 0, self.GRAN

    def cuMemCreate(self, size, prop, flags):
        assert size % self.GRAN == 0 and prop.requestedHandleTypes == _E.X
        with self.lock:
            h = next(self.ids)
            self.mem[h] = {"size": size, "refs": 1}
        return 0, h

    def cuMulticastCreate(self, prop):
        assert prop.size % self.GRAN == 0 and prop.handleTypes == int(_E.X)
        with self.lock:
            h = next(self.ids)
            self.mc[h] = {"size": prop.size, "want": prop.numDevices, "devices": set(), "bound": {}, "refs": 1}
        return 0, h

    def cuMemExportToShareableHandle(self, handle, kind, flags):
        assert kind == _E.X and (handle in self.mem or handle in self.mc)
        fh = tempfile.TemporaryFile()
        fh.write(str(handle).encode())
        fh.flush()
        return 0, os.dup(fh.fileno())

    def cuMemImportFromShareableHandle(self, fd, kind):
        handle = int(os.pread(fd, 32, 0).decode())
        with self.lock:
            (self.mem.get(handle) or self.mc[handle])["refs"] += 1
        return 0, handle

    def cuMulticastAddDevice(self, mc, dev):
        with self.lock:
            assert not self.mc[mc]["bound"], "a device was added after memory had been bound"
            self.mc[mc]["devices"].add(dev)
        return (0,)

    def cuMulticastBindMem(self, mc, mc_off, mem, mem_off, size, flags):
        with self.lock:
            m = self.mc[mc]
            assert len(m["devices"]) == m["want"], "memory bound before every device had joined"
            assert size == m["size"] == self.mem[mem]["size"] and mc_off == 0 and mem_off == 0
            m["bound"][threading.get_ident()] = mem
        return (0,)

    def cuMemAddressReserve(self, size, align, addr, flags):
        with self.lock:
            va = next(self.ids) * (1 << 32)
            self.vas.add(va)
        return 0, va

    def cuMemMap(self, va, size, offset, handle, flags):
        with self.lock:
            assert va in self.vas and va not in self.maps and (handle in self.mem or handle in self.mc)
            if handle in self.mc:
                assert len(self.mc[handle]["devices"]) == self.mc[handle]["want"]
            self.maps[va] = handle
        return (0,)

    def cuMemSetAccess(self, va, size, desc, count):
        assert va in self.maps and count == len(desc) == 1 and desc[0].flags == 3
        return (0,)

    def cuMemsetD8(self, va, value, size):
        assert self.maps[va] in self.mem
        return (0,)

    def cuMemUnmap(self, va, size):
        with self.lock:
            del self.maps[va]
        return (0,)

    def cuMemAddressFree(self, va, size):
        with self.lock:
            assert va not in self.maps
            self.vas.remove(va)
        return (0,)

    def cuMulticastUnbind(self, mc, dev, off, size):
        with self.lock:
            del self.mc[mc]["bound"][threading.get_ident()]
        return (0,)

    def cuMemRelease(self, handle):
        with self.lock:
            (self.mem.get(handle) or self.mc[handle])["refs"] -= 1
        return (0,)

    def cuGetErrorName(self, err):
        return 0, b"FAKE_ERROR"


def test_region_life_cycle_on_thread_ranks(monkeypatch):
    import torch

    from mpitree_b200.parallel import nvls

    fake = FakeDriver()
    monkeypatch.setattr(nvls, "_driver", lambda: fake)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: types.SimpleNamespace(synchronize=lambda: None))
    before = len(os.listdir("/proc/self/fd"))

    def work(comm):
        reg = nvls.NvlsRegion(300_000 + comm.rank, comm)  # sizes differ per rank: the common size is the rounded maximum
        out = (reg.size, len(set(reg.ptrs)), reg.mc_ptr not in reg.ptrs and reg.mc_ptr != 0, fake.maps[reg.mc_ptr] in fake.mc)
        mine = fake.maps[reg.ptrs[comm.rank]]
        comm.barrier()
        reg.close()
        comm.barrier()
        return out, mine

    out = run_threaded(3, work)
    assert all(o[0] == (FakeDriver.GRAN, 3, True, True) for o in out)
    assert len({o[1] for o in out}) == 3  # every rank mapped its own allocation at ptrs[rank]
    assert not fake.maps and not fake.vas
    assert all(m["refs"] == 0 for m in fake.mem.values()) and all(m["refs"] == 0 and not m["bound"] for m in fake.mc.values())
    assert len(fake.mc) == 1 and len(fake.mem) == 3
    assert len(os.listdir("/proc/self/fd")) <= before + 1  # descriptors of the exchange were all closed


def test_driver_errors_name_the_failing_call(monkeypatch):
    from mpitree_b200.parallel import nvls

    fake = FakeDriver()
    monkeypatch.setattr(nvls, "_driver", lambda: fake)
    with pytest.raises(RuntimeError, match="cuMemCreate failed: FAKE_ERROR"):
        nvls._ck((1, None), "cuMemCreate")
    assert nvls._ck((0, 7), "x") == 7 and nvls._ck((0,), "x") == ()


def test_peer_allreduce_nvls_front_end(monkeypatch):
    """``PeerAllReduce(nvls=True)``: buffer layout, the multicast pointer handed to the native call, epochs."""
    import ctypes

    import torch

    from mpitree_b200.ops import _lib as L
    from mpitree_b200.parallel import allreduce, nvls

    fake = FakeDriver()
    calls = []

    class Lib:
        def mtb_allreduce_u32_nvls(self, rank, world, bufs, flags, mc, count, cap, epoch, out, stream):
            calls.append(("nvls", rank.value, world.value, [bufs[i] for i in range(2)], [flags[i] for i in range(2)], mc.value,
                          count.value, cap.value, epoch.value))
            return 0

        def mtb_allreduce_u32(self, algo, rank, world, bufs, flags, count, cap, epoch, out, stream):
            calls.append(("peer", algo.value, epoch.value))
            return 0

    monkeypatch.setattr(nvls, "_driver", lambda: fake)
    monkeypatch.setattr(nvls.NvlsRegion, "as_tensor", lambda self: torch.zeros(self.nbytes, dtype=torch.uint8))
    monkeypatch.setattr(L, "load", lambda: Lib())
    monkeypatch.setattr(L, "check", lambda rc, what: None)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(torch.cuda, "current_stream",
                        lambda *a, **k: types.SimpleNamespace(synchronize=lambda: None, cuda_stream=0))

    def work(comm):
        ar = allreduce.PeerAllReduce(comm, 1001, nvls=True)
        assert ar.cap == 1004 and ar.region.nbytes == 2 * 1004 * 4 + ar.FLAG_BYTES
        ar.launch(1001, allreduce.ALGO_NVLS)
        ar.launch(10, allreduce.ALGO_TWOSHOT)
        ar.launch(10, allreduce.ALGO_NVLS)
        region = ar.region
        plain = types.SimpleNamespace(lib=ar.lib, L=L, rank=0, world=2, _bufs=ar._bufs, _flags=ar._flags, _mc=0, cap=ar.cap,
                                      epoch=0, torch=torch, result_view=lambda n: torch.zeros(n, dtype=torch.int32))
        with pytest.raises(ValueError, match="nvls=True"):
            allreduce.PeerAllReduce.launch(plain, 4, allreduce.ALGO_NVLS)
        ptrs, mc = list(region.ptrs), region.mc_ptr  # close() clears them
        comm.barrier()
        ar.close()
        return ptrs, mc, ar.epoch

    calls_before = len(calls)
    out = run_threaded(2, work)
    assert calls_before == 0 and out[0][2] == out[1][2] == 6
    first = [c for c in calls if c[0] == "nvls" and c[6] == 1001]
    assert len(first) == 2
    for c in first:
        ptrs, mc = (out[c[1]][0], out[c[1]][1])
        assert c[2] == 2 and c[3] == ptrs and c[4] == [p + 2 * 1004 * 4 for p in ptrs] and c[5] == mc and c[7] == 1004 and c[8] == 0
    assert sorted(c[2] for c in calls if c[0] == "peer") == [2, 2]  # the peer algorithms run on the same region, later epoch
    assert sorted(c[8] for c in calls if c[0] == "nvls" and c[6] == 10) == [4, 4]
