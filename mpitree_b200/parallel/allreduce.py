"""K5 front end: peer-memory u32 sum all-reduce (one-shot / two-shot) vs ``ncclAllReduce``.

There is no counterpart in the reference (its only collective is the pickled subtree
``allgather``, ``mpitree/tree/decision_tree.py:456-466``).  Histograms are integer counts, so
the peer reduction is bit-deterministic regardless of arrival order.
"""

from __future__ import annotations

import ctypes

__all__ = ["PeerAllReduce", "ALGO_ONESHOT", "ALGO_TWOSHOT", "ALGO_NVLS", "ALGO_AUTO"]


ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS = 0, 1, 2
ALGO_AUTO = -1
#: size policy of ``ALGO_AUTO`` (bytes), from the 8-GPU sweep (profiles/r2/allreduce_sweep_8gpu.txt): with
#: multicast-bound buffers and four ranks or more the in-switch kernel is the fastest of ours at every size
#: (16.7 us at 4 KB ... 770 GB/s bus bandwidth at 256 MB; NCCL: 25.5 us ... 644 GB/s); otherwise one-shot up
#: to 16 KB and two-shot above
AUTO_ONESHOT_MAX_BYTES = 16 << 10
AUTO_NVLS_MIN_BYTES = 0


class PeerAllReduce:
    """Symmetric buffers + flags for repeated all-reduces of up to ``max_count`` u32 values.

    Producers can write straight into :meth:`input_view` (zero-copy); :meth:`all_reduce`
    copies from an ordinary tensor first.

    ``nvls=True`` places the buffers in an
    :class:`mpitree_b200.parallel.nvls.NvlsRegion` so that ``algo=2`` can reduce inside the
    NVSwitch; algorithms 0 and 1 work on either kind of region.
    """

    FLAG_BYTES = 512 * 16 * 4  # one row of world flags per CTA index

    def __init__(self, comm, max_count: int, nvls: bool = False):
        import torch

        from ..ops import _lib as L

        self.L, self.lib = L, L.load()
        self.comm = comm
        self.world, self.rank = comm.size, comm.rank
        self.cap = ((int(max_count) + 3) // 4) * 4
        nbytes = 2 * self.cap * 4 + self.FLAG_BYTES
        if nvls:
            from .nvls import NvlsRegion

            self.region = NvlsRegion(nbytes, comm)
            self._mem = self.region.as_tensor()
            self._mc = int(self.region.mc_ptr)
        else:
            from .symm import IpcRegion

            self.region = IpcRegion(nbytes, comm)
            self._mem = self.region.buf.as_tensor()
            self._mc = 0
        self._bufs = (ctypes.c_void_p * 16)(*[p for p in self.region.ptrs])
        self._flags = (ctypes.c_void_p * 16)(*[p + 2 * self.cap * 4 for p in self.region.ptrs])
        self.epoch = 0
        self.torch = torch

    def input_view(self, count: int):
        return self._mem[: count * 4].view(self.torch.int32)

    def result_view(self, count: int):
        return self._mem[self.cap * 4 : (self.cap + count) * 4].view(self.torch.int32)

    def launch(self, count: int, algo: int = 1, out=None) -> None:
        """Reduce what is in :meth:`input_view`; result in ``out`` or :meth:`result_view`."""
        torch = self.torch
        if algo == ALGO_AUTO:
            nbytes = 4 * count
            algo = (ALGO_NVLS if (self._mc and self.world >= 4 and nbytes >= AUTO_NVLS_MIN_BYTES) else
                    ALGO_ONESHOT if nbytes <= AUTO_ONESHOT_MAX_BYTES else ALGO_TWOSHOT)
        if out is None and algo == 0:  # one-shot writes its sums straight to the destination
            out = self.result_view(count)
        dst = out.data_ptr() if out is not None else 0
        if algo == ALGO_NVLS:
            if not self._mc:
                raise ValueError("algo=2 needs PeerAllReduce(..., nvls=True)")
            rc = self.lib.mtb_allreduce_u32_nvls(
                ctypes.c_int(self.rank), ctypes.c_int(self.world), self._bufs, self._flags, ctypes.c_void_p(self._mc),
                ctypes.c_int64(count), ctypes.c_int64(self.cap), ctypes.c_uint32(self.epoch), ctypes.c_void_p(dst),
                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream),
            )
            self.epoch = (self.epoch + 2) & 0x7FFFFFFF
            self.L.check(rc, "mtb_allreduce_u32_nvls")
            return
        rc = self.lib.mtb_allreduce_u32(
            ctypes.c_int(algo), ctypes.c_int(self.rank), ctypes.c_int(self.world), self._bufs, self._flags,
            ctypes.c_int64(count), ctypes.c_int64(self.cap), ctypes.c_uint32(self.epoch), ctypes.c_void_p(dst),
            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream),
        )
        self.epoch = (self.epoch + 2) & 0x7FFFFFFF
        self.L.check(rc, "mtb_allreduce_u32")

    def all_reduce(self, tensor, algo: int = 1):
        """Sum ``tensor`` (int32/uint32, 1-D, CUDA) across ranks; returns a new tensor."""
        count = tensor.numel()
        if count > self.cap:
            raise ValueError(f"{count} elements exceed the symmetric buffer capacity {self.cap}")
        self.input_view(count).copy_(tensor.view(self.torch.int32).reshape(-1))
        out = self.torch.empty(count, dtype=self.torch.int32, device=tensor.device)
        self.launch(count, algo, out)
        return out.view(tensor.dtype)

    def close(self) -> None:
        self.region.close()
