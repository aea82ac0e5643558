"""Pass open file descriptors between the ranks of one node (``SCM_RIGHTS`` over unix sockets).

CUDA's virtual-memory-management handles (``cuMemCreate`` allocations and ``cuMulticastCreate``
multicast objects) are shared between processes as POSIX file descriptors, and a descriptor
number means nothing in another process: it has to travel as ancillary data of a unix-domain
socket message.  ``torch.distributed`` has no such channel, so this module provides the one
collective that :mod:`mpitree_b200.parallel.nvls` needs: an all-gather of descriptor lists.

The control-plane communicator (``comm.allgather_object`` / ``comm.barrier``) is only used to
agree on a socket name and to fence set-up and tear-down.  The reference has no counterpart
(its ranks only exchange pickles, ``mpitree/tree/decision_tree.py:456-466``).
"""

from __future__ import annotations

import array
import os
import select
import socket
import struct
from typing import List, Sequence

__all__ = ["allgather_fds"]

_MAX_FDS = 8
_COUNTER = [0]


def _address(token: str, tag: str, rank: int) -> str:
    # leading NUL = Linux abstract namespace: nothing to unlink, private to the network namespace
    return f"\0mtb-fd-{token}-{tag}-{rank}"


def allgather_fds(comm, fds: Sequence[int], tag: str = "x", timeout: float = 60.0) -> List[List[int]]:
    """Every rank contributes ``fds`` (0 .. 8 open descriptors); returns one list per rank.

    Entry ``rank`` of the result holds duplicates of this rank's own descriptors, so the caller
    owns (and must close) every descriptor in the result, and still owns the ones it passed in.
    All ranks of ``comm`` must run on the same host.
    """
    world, rank = comm.size, comm.rank
    fds = [int(f) for f in fds]
    if len(fds) > _MAX_FDS:
        raise ValueError(f"at most {_MAX_FDS} descriptors per rank")
    if world == 1:
        return [[os.dup(f) for f in fds]]
    _COUNTER[0] += 1
    token = comm.allgather_object(f"{os.getpid()}-{_COUNTER[0]}-{os.urandom(4).hex()}" if rank == 0 else None)[0]
    sock = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
    try:
        sock.bind(_address(token, tag, rank))
        sock.setblocking(False)
        comm.barrier()  # every rank is bound before anybody sends
        out: List[List[int]] = [[] for _ in range(world)]
        out[rank] = [os.dup(f) for f in fds]
        todo = [(rank + s) % world for s in range(1, world)]
        header = struct.pack("ii", rank, len(fds))
        got = 0
        waited = 0.0
        while todo or got < world - 1:
            progress = False
            if todo:
                try:
                    # (socket.send_fds ignores its address argument, so build the message here)
                    anc = [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", fds))] if fds else []
                    sock.sendmsg([header], anc, 0, _address(token, tag, todo[0]))
                    todo.pop(0)
                    progress = True
                except (BlockingIOError, InterruptedError):
                    pass
            if got < world - 1:
                try:
                    msg, rfds, _flags, _addr = socket.recv_fds(sock, 64, _MAX_FDS)
                    src, cnt = struct.unpack("ii", msg[:8])
                    if cnt != len(rfds):
                        raise RuntimeError(f"rank {src} announced {cnt} descriptors, {len(rfds)} arrived")
                    out[src] = list(rfds)
                    got += 1
                    progress = True
                except (BlockingIOError, InterruptedError):
                    pass
            if not progress:
                select.select([sock], [sock] if todo else [], [], 0.05)
                waited += 0.05
                if waited > timeout:
                    raise TimeoutError(f"descriptor exchange '{tag}' stalled on rank {rank} ({got}/{world - 1} received)")
        comm.barrier()  # nobody closes its socket while a peer may still be sending to it
        return out
    finally:
        sock.close()
