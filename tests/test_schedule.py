"""N>1 host logic without GPUs: the block-rotation schedule is conflict-free and complete, and a
world_size-2 gloo run really rotates the C blocks between two processes."""
import os
import sys

import numpy as np
import pytest

from graph_embeddings_b200 import schedule as S


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_schedule_is_conflict_free_and_complete(world):
    nb = 8
    seen = set()
    rings = [S.Ring(r, world, nb) for r in range(world)]
    for s in range(nb):
        rows, cols = [], []
        for r in range(world):
            w = S.work(r, world, nb, s)
            win = S.window(r, world, nb, s)
            assert [cb for _, cb in w] == win                       # a rank only updates blocks it holds
            assert [rings[r].block_in_slot[rings[r].slot(k)] for k in range(len(win))] == win
            rows += [rb for rb, _ in w]
            cols += [cb for _, cb in w]
            seen.update(w)
        assert sorted(rows) == list(range(nb)) and sorted(cols) == list(range(nb))   # no shared W / C block
        firsts = [S.window(r, world, nb, s)[0] for r in range(world)]
        for r in range(world):
            to, frm = S.rotation(r, world)
            rings[r].rotate(firsts[frm])
    assert len(seen) == nb * nb                                     # every block exactly once per epoch
    for r in range(world):                                          # every C block is home again
        assert sorted(rings[r].block_in_slot.values()) == S.window(r, world, nb, 0)


def test_row_ranges_partition_the_ids():
    for V in (7, 8, 1000, 10_000_001):
        for world in (1, 2, 4, 8):
            got = [S.row_range(V, 8, r, world) for r in range(world)]
            assert got[0][0] == 0 and sum(c for _, c in got) == V
            for (a, ca), (b, _) in zip(got, got[1:]):
                assert a + ca == b or ca == 0


def _gloo_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nb, Bs = 8, 5
    ring = S.Ring(rank, world, nb)
    slots = torch.zeros(ring.slots, Bs)
    for k in range(ring.m):                                         # block b carries the value b
        slots[k] = float(rank * ring.m + k)
    visits = []
    for s in range(nb):
        for k, (rb, cb) in enumerate(S.work(rank, world, nb, s)):
            assert slots[ring.slot(k)][0].item() == cb              # the slot really holds block cb
            slots[ring.slot(k)] += 0.0
            visits.append((rb, cb))
        to, frm = S.rotation(rank, world)
        first_block = S.window(frm, world, nb, s)[0]
        reqs = [dist.isend(slots[ring.send_slot()].clone(), to), dist.irecv(slots[ring.recv_slot()], frm)]
        for q in reqs:
            q.wait()
        ring.rotate(first_block)
    home = sorted(slots[ring.slot(k)][0].item() for k in range(ring.m))
    gathered = [None] * world
    dist.all_gather_object(gathered, (visits, home))
    if rank == 0:
        out.put(gathered)
    dist.destroy_process_group()


def test_gloo_world2_rotation():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    allv = [v for visits, _ in res for v in visits]
    assert len(allv) == 64 and len(set(allv)) == 64
    assert res[0][1] == [0.0, 1.0, 2.0, 3.0] and res[1][1] == [4.0, 5.0, 6.0, 7.0]
