"""Host-side communicators (bootstrap / control plane).

The reference talks to ``mpi4py``'s ``MPI.COMM_WORLD`` directly
(``mpitree/tree/decision_tree.py:313-317``) and uses exactly four operations:
``Get_rank``/``Get_size``, ``Split(color, key)`` (``:336-338``), a pickled
``allgather`` (``:456-466``) and ``Free`` (``:477``).  Here the process group is
``torch.distributed`` (one rank per GPU, ``nccl`` for device tensors and a
``gloo`` side group for host metadata) and it is used only for bootstrap,
metadata and the CPU engine; the GPU data plane is peer memory written from
inside kernels (see :mod:`mpitree_b200.parallel.symm`).

Three interchangeable implementations share one small interface
(``rank``, ``size``, ``allreduce_sum``, ``allgather_object``, ``broadcast_object``,
``barrier``, ``split``):

* :class:`SelfComm`    — one rank, no dependencies;
* :class:`TorchComm`   — a ``torch.distributed`` process group;
* :class:`ThreadComm`  — N ranks as threads of one process (tests, and the
  "emulate ranks on one device" mode), no sockets involved.
"""

from __future__ import annotations

import os
import threading
from typing import Any, List, Optional, Sequence

import numpy as np

__all__ = ["SelfComm", "TorchComm", "ThreadComm", "world", "init_process_group_from_env"]


class SelfComm:
    """Trivial single-rank communicator."""

    rank = 0
    size = 1

    def allreduce_sum(self, arr: np.ndarray) -> np.ndarray:
        return arr

    def allgather_object(self, obj: Any) -> List[Any]:
        return [obj]

    def allgather_ints(self, values) -> "np.ndarray":
        return np.asarray([[int(v) for v in values]], dtype=np.int64)

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        return obj

    def barrier(self) -> None:
        return None

    def split(self, color: int, key: int = 0) -> "SelfComm":
        return self

    def free(self) -> None:
        return None

    # mpi4py spellings, so code written against the reference's comm keeps working
    Get_rank = lambda self: self.rank  # noqa: E731
    Get_size = lambda self: self.size  # noqa: E731


class TorchComm:
    """Communicator over a ``torch.distributed`` process group.

    Host arrays travel over a gloo group (created lazily when the default group
    is NCCL); device tensors should use :mod:`mpitree_b200.parallel.symm`.
    """

    def __init__(self, group=None, ranks: Optional[Sequence[int]] = None):
        import torch.distributed as dist

        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised; call init_process_group_from_env()")
        self._dist = dist
        self._ranks = list(ranks) if ranks is not None else list(range(dist.get_world_size()))
        self._group = group if group is not None else _host_group()
        self.rank = self._ranks.index(dist.get_rank())
        self.size = len(self._ranks)

    @property
    def global_ranks(self) -> List[int]:
        return list(self._ranks)

    def allreduce_sum(self, arr: np.ndarray) -> np.ndarray:
        import torch

        t = torch.from_numpy(np.ascontiguousarray(arr))
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self._group)
        return t.numpy()

    def allgather_object(self, obj: Any) -> List[Any]:
        out = [None] * self.size
        self._dist.all_gather_object(out, obj, group=self._group)
        return out

    def allgather_ints(self, values) -> "np.ndarray":
        """Gather a few integers from every rank -> ``(size, k)`` int64 array.

        On a NCCL world this is one tiny device all-gather (tens of microseconds) instead of
        the pickled gloo exchange, which costs milliseconds at 8 ranks.
        """
        import torch

        vals = [int(v) for v in values]
        world_is_default = self._ranks == list(range(self._dist.get_world_size()))
        if world_is_default and self._dist.get_backend() == "nccl" and torch.cuda.is_available():
            t = torch.tensor(vals, dtype=torch.int64, device="cuda")
            out = torch.empty(self.size * len(vals), dtype=torch.int64, device="cuda")
            self._dist.all_gather_into_tensor(out, t)
            return out.cpu().numpy().reshape(self.size, len(vals))
        return np.asarray(self.allgather_object(vals), dtype=np.int64)

    def allgather_device(self, t):
        """All-gather a small device tensor without touching the host: ``(k,)`` -> ``(size, k)``.

        Lets a caller fold device-side reductions (label range, ...) and host integers into one
        exchange with a single read-back.  Only when this communicator is the NCCL world (the device
        collective runs on the default group); otherwise ``None`` and callers use :meth:`allgather_ints`."""
        import torch

        world_is_default = self._ranks == list(range(self._dist.get_world_size()))
        if not (world_is_default and self._dist.get_backend() == "nccl" and t.is_cuda):
            return None
        out = torch.empty(self.size * t.numel(), dtype=t.dtype, device=t.device)
        self._dist.all_gather_into_tensor(out, t.contiguous())
        return out.view(self.size, t.numel())

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        box = [obj]
        self._dist.broadcast_object_list(box, src=self._ranks[src], group=self._group)
        return box[0]

    def barrier(self) -> None:
        self._dist.barrier(group=self._group)

    def split(self, color: int, key: int = 0) -> "TorchComm":
        """``MPI_Comm_split``: ranks with equal ``color`` form a group ordered by ``key``."""
        info = self.allgather_object((int(color), int(key), self._ranks[self.rank]))
        mine = None
        # every rank must create every subgroup, in the same order
        for c in sorted({c for c, _, _ in info}):
            members = [g for _, g in sorted((k, g) for cc, k, g in info if cc == c)]
            grp = self._dist.new_group(ranks=members, backend="gloo")
            if c == color:
                mine = TorchComm(group=grp, ranks=members)
        return mine

    def free(self) -> None:
        if self._group is not None and self._group is not _HOST_GROUP.get("g"):
            try:
                self._dist.destroy_process_group(self._group)
            except Exception:  # pragma: no cover - best effort
                pass

    Get_rank = lambda self: self.rank  # noqa: E731
    Get_size = lambda self: self.size  # noqa: E731


_HOST_GROUP: dict = {}


def _host_group():
    """A gloo group spanning the world (host metadata path)."""
    import torch.distributed as dist

    if "g" not in _HOST_GROUP:
        if dist.get_backend() == "gloo":
            _HOST_GROUP["g"] = dist.group.WORLD
        else:
            _HOST_GROUP["g"] = dist.new_group(backend="gloo")
    return _HOST_GROUP["g"]


def init_process_group_from_env(backend: Optional[str] = None) -> bool:
    """Initialise ``torch.distributed`` from torchrun's environment, if present.

    Returns True when a multi-process group exists afterwards.
    """
    import torch
    import torch.distributed as dist

    if dist.is_initialized():
        return dist.get_world_size() > 1
    if "RANK" not in os.environ or int(os.environ.get("WORLD_SIZE", "1")) <= 1:
        return False
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group(
            backend="nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        )
    else:
        dist.init_process_group(backend=backend)
    return True


def world():
    """The communicator for "all ranks of this job" (``COMM_WORLD`` equivalent)."""
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            return TorchComm()
    except ImportError:  # pragma: no cover
        pass
    return SelfComm()


# ---------------------------------------------------------------------------
# in-process multi-rank communicator (threads)
# ---------------------------------------------------------------------------
class _ThreadWorld:
    def __init__(self, size: int):
        self.size = size
        self.barrier = threading.Barrier(size)
        self.slots: List[Any] = [None] * size
        self.lock = threading.Lock()
        self.children: dict = {}


class ThreadComm:
    """``size`` ranks living in threads of one process.

    Create with :meth:`ThreadComm.create` and hand one communicator to each
    thread, or use :func:`run_threaded`.
    """

    def __init__(self, shared: _ThreadWorld, rank: int):
        self._w = shared
        self.rank = rank
        self.size = shared.size
        self._epoch = 0

    @classmethod
    def create(cls, size: int) -> List["ThreadComm"]:
        w = _ThreadWorld(size)
        return [cls(w, r) for r in range(size)]

    def _exchange(self, obj: Any) -> List[Any]:
        w = self._w
        w.slots[self.rank] = obj
        w.barrier.wait()
        out = list(w.slots)
        w.barrier.wait()
        return out

    def allreduce_sum(self, arr: np.ndarray) -> np.ndarray:
        parts = self._exchange(np.array(arr, copy=True))
        total = parts[0].copy()
        for p in parts[1:]:
            total += p
        return total

    def allgather_object(self, obj: Any) -> List[Any]:
        return self._exchange(obj)

    def allgather_ints(self, values) -> "np.ndarray":
        return np.asarray(self._exchange([int(v) for v in values]), dtype=np.int64)

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        return self._exchange(obj)[src]

    def barrier(self) -> None:
        self._w.barrier.wait()

    def split(self, color: int, key: int = 0) -> "ThreadComm":
        info = self._exchange((int(color), int(key), self.rank))
        members = [r for _, r in sorted((k, r) for c, k, r in info if c == color)]
        self._epoch += 1
        tag = (self._epoch, color)
        with self._w.lock:
            if tag not in self._w.children:
                self._w.children[tag] = _ThreadWorld(len(members))
            child = self._w.children[tag]
        self._w.barrier.wait()
        return ThreadComm(child, members.index(self.rank))

    def free(self) -> None:
        return None

    Get_rank = lambda self: self.rank  # noqa: E731
    Get_size = lambda self: self.size  # noqa: E731


def run_threaded(size: int, fn, *args, **kwargs) -> List[Any]:
    """Run ``fn(comm, *args, **kwargs)`` on ``size`` thread-ranks; return all results.

    An exception on any rank aborts the others (the reference deadlocks in
    ``allgather`` instead, ``decision_tree.py:456``) and is re-raised here.
    """
    comms = ThreadComm.create(size)
    results: List[Any] = [None] * size
    errors: List[Optional[BaseException]] = [None] * size

    def work(r: int) -> None:
        try:
            results[r] = fn(comms[r], *args, **kwargs)
        except BaseException as exc:  # noqa: BLE001
            errors[r] = exc
            comms[r]._w.barrier.abort()
            for child in list(comms[r]._w.children.values()):
                child.barrier.abort()

    threads = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(size)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    real = [e for e in errors if e is not None and not isinstance(e, threading.BrokenBarrierError)]
    if real:
        raise real[0]
    broken = [e for e in errors if e is not None]
    if broken:
        raise broken[0]
    return results
