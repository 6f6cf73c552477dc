"""Block-rotation schedule of the multi-GPU trainer as plain host logic.

The same arithmetic is used by csrc/dist.cu (ge_dist_train_epoch); it lives here too so that the
N>1 host logic can be exercised without GPUs (tests/test_schedule.py: conflict-freedom proof by
enumeration, and a world_size-2 gloo run that actually rotates blocks between processes)."""
from __future__ import annotations


def block_size(V: int, nb: int) -> int:
    return (V + nb - 1) // nb


def row_range(V: int, nb: int, rank: int, world: int):
    """ids owned by `rank`: row-blocks rank*m .. rank*m+m-1 (ge_dist_row_range)."""
    assert nb % world == 0
    Bs, m = block_size(V, nb), nb // world
    lo, hi = min(V, rank * m * Bs), min(V, (rank + 1) * m * Bs)
    return lo, hi - lo


def window(rank: int, world: int, nb: int, s: int):
    """C blocks held by `rank` during sub-epoch s, in window order."""
    m = nb // world
    return [(rank * m + s + k) % nb for k in range(m)]


def work(rank: int, world: int, nb: int, s: int):
    """(row_block, col_block) pairs `rank` updates in sub-epoch s: virtual rank v = rank*m+k owns
    row-block v and takes column block (v + s) mod nb."""
    m = nb // world
    return [(rank * m + k, (rank * m + k + s) % nb) for k in range(m)]


def rotation(rank: int, world: int):
    """(send_to, recv_from) after every sub-epoch: the window's first block goes to rank-1."""
    return (rank - 1) % world, (rank + 1) % world


class Ring:
    """Slot bookkeeping of the C-block ring buffer (m + 1 slots when world > 1)."""

    def __init__(self, rank: int, world: int, nb: int):
        self.rank, self.world, self.nb, self.m = rank, world, nb, nb // world
        self.slots = self.m + 1 if world > 1 else self.m
        self.head = 0
        self.block_in_slot = {k: rank * self.m + k for k in range(self.m)}

    def slot(self, k: int) -> int:
        return (self.head + k) % self.slots

    def send_slot(self) -> int:
        return self.head

    def recv_slot(self) -> int:
        return (self.head + self.m) % self.slots

    def rotate(self, received_block: int):
        if self.world > 1:
            del self.block_in_slot[self.head]
            self.block_in_slot[self.recv_slot()] = received_block
        self.head = (self.head + 1) % self.slots
