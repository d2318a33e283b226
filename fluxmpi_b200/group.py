"""Rank groups — the ``comm`` argument of the communication primitives.

The reference's ``allreduce!/bcast!/reduce!`` and ``Iallreduce!/Ibcast!`` take any MPI communicator
(/root/reference/src/mpi_extensions.jl:97,119,141 and :52-60,84-88); everything above them uses
``MPI.COMM_WORLD``. Here a :class:`Group` is the communicator of a subset of the world's ranks:

    g = fm.Group([0, 1])            # collective over the WORLD: every rank makes the same calls
    if g.member:
        fm.allreduce_(x, "sum", comm=g)      # only ranks 0 and 1 take part
        fm.bcast_(y, root=1, comm=g)         # roots are ranks *within the group* (MPI semantics)

What a group owns
  * host side: its own shared-memory world (``csrc/host_backend.cpp``, segment ``<job>_g<id>``), so
    two disjoint groups run host collectives concurrently and independently of the world;
  * device side: its own set of signal pads in every member's base segment and ``fmpi_devcomm``
    tables that hold only the members' windows (``parallel/device_world.py``). The multicast object
    spans the whole world, so groups smaller than the world use the peer-to-peer kernels.

Every group has staging of its own in the members' windows (a non-member can be inside a world
collective while the members run a group collective), so collectives of different communicators
are independent of each other; calls on ONE communicator follow the usual rule: the same sequence
on every member.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

from . import _native, runtime

_counter = 0


class Group:
    def __init__(self, ranks: Sequence[int]):
        global _counter
        if not runtime.Initialized():
            raise runtime.FluxMPINotInitializedError()
        S = runtime.state()
        ranks = [int(x) for x in ranks]
        if not ranks or sorted(set(ranks)) != sorted(ranks) or min(ranks) < 0 or max(ranks) >= S.world:
            raise ValueError(f"Group needs distinct ranks in [0, {S.world}), got {ranks}")
        if S.nnodes > 1:
            raise NotImplementedError("rank groups are node-local (one shared-memory / NVSwitch domain)")
        _counter += 1
        self.id = _counter                      # the same on every rank: creation is collective
        self.ranks = sorted(ranks)
        self.size = len(self.ranks)
        self.member = S.rank in self.ranks
        self.rank = self.ranks.index(S.rank) if self.member else -1
        self.host_comm: Optional[int] = None
        self._dev = None                        # (comms, refs) per channel, built on first device use
        self._freed = False
        dw = S.device_world
        if dw is not None and getattr(dw, "native", False) and self.id >= dw.MAX_COMMS:
            raise RuntimeError(f"at most {dw.MAX_COMMS - 1} groups per job (signal-pad sets)")
        if self.member and self.size > 1:
            h = C.c_void_p()
            name = ("/fmpi_" + "".join(ch if ch.isalnum() else "_" for ch in S.job) + f"_g{self.id}")[:200]

            def make():
                _native.host_check(_native.host().fmpi_host_init(
                    name.encode(), self.rank, self.size, S.slot_bytes, S.timeout_s, C.byref(h)), "group host world")

            runtime.host_run(make)
            self.host_comm = h.value
        runtime.host_barrier()                  # everyone (members or not) leaves together
        _groups.append(self)

    # -- introspection -------------------------------------------------------------------
    def __repr__(self):
        return f"Group(id={self.id}, ranks={self.ranks}, rank={self.rank})"

    def world_rank(self, group_rank: int) -> int:
        return self.ranks[group_rank]

    def _check(self):
        if not self.member:
            raise ValueError(f"rank {runtime.local_rank()} is not a member of {self!r}")
        if self._freed:
            raise ValueError("this group has been freed")

    # -- device side ---------------------------------------------------------------------
    def device_refs(self):
        """ctypes references to this group's device communicators, one per signal channel."""
        if self._dev is None:
            dw = runtime.state().device_world
            comms = [dw.comm_for(dw.base, ch, data_offset=dw.sig_bytes, data_bytes=dw.base.size - dw.sig_bytes,
                                 comm_id=self.id, members=self.ranks) for ch in range(_native.NUM_CHANNELS)]
            self._dev = (comms, [C.byref(c) for c in comms])
        return self._dev[1]

    def free(self):
        """Release the group's host world (collective over the members)."""
        if self._freed:
            return
        self._freed = True
        if self.host_comm is not None:
            _native.host().fmpi_host_finalize(self.host_comm)
            self.host_comm = None
        if self in _groups:
            _groups.remove(self)


_groups: list = []


def free_all():
    for g in list(_groups):
        try:
            g.free()
        except Exception:                       # noqa: BLE001
            pass


def resolve(comm):
    """None for the world (``COMM_WORLD`` / ``None``), else the Group (checked for membership)."""
    if comm is None or isinstance(comm, str):
        return None
    if isinstance(comm, Group):
        comm._check()
        return comm
    raise TypeError(f"comm must be COMM_WORLD or a fluxmpi_b200.Group, got {type(comm).__name__}")
