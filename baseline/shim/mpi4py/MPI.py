"""``mpi4py.MPI`` subset: COMM_WORLD / Intracomm with Get_rank, Get_size, Split, allgather, Free."""

from __future__ import annotations

import os


class Intracomm:
    def __init__(self, ranks=None, group=None):
        self._ranks = ranks  # global ranks in this communicator, ordered by key
        self._group = group

    # -- lazy world ---------------------------------------------------------
    def _ensure(self):
        if self._ranks is not None:
            return
        world = int(os.environ.get("WORLD_SIZE", "1"))
        if world > 1:
            import torch.distributed as dist

            if not dist.is_initialized():
                os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
                os.environ.setdefault("MASTER_PORT", "29511")
                dist.init_process_group(backend="gloo")
            self._ranks = list(range(dist.get_world_size()))
            self._group = dist.group.WORLD
        else:
            self._ranks = [0]

    def _global_rank(self):
        if len(self._ranks) == 1 and int(os.environ.get("WORLD_SIZE", "1")) <= 1:
            return 0
        import torch.distributed as dist

        return dist.get_rank()

    def Get_rank(self):
        self._ensure()
        return self._ranks.index(self._global_rank())

    def Get_size(self):
        self._ensure()
        return len(self._ranks)

    def allgather(self, obj):
        self._ensure()
        if len(self._ranks) == 1:
            return [obj]
        import torch.distributed as dist

        out = [None] * len(self._ranks)
        dist.all_gather_object(out, obj, group=self._group)
        return out

    def Split(self, color=0, key=0):
        self._ensure()
        if len(self._ranks) == 1:
            return Intracomm([self._ranks[0]], None)
        import torch.distributed as dist

        info = self.allgather((int(color), int(key), self._global_rank()))
        mine = None
        for c in sorted({c for c, _, _ in info}):
            members = [g for _, g in sorted((k, g) for cc, k, g in info if cc == c)]
            # new_group is collective over the WORLD: only safe when the parent is the
            # world or when all world ranks reach this call; the reference's recursion
            # splits sibling groups in lock-step only for balanced trees, so subgroup
            # creation is restricted to the members (use_local_synchronization).
            grp = dist.new_group(ranks=members, backend="gloo", use_local_synchronization=True) if c == color else None
            if c == color:
                mine = Intracomm(members, grp)
        return mine

    def Free(self):
        return None


COMM_WORLD = Intracomm()
