"""Host-side logic of the slab decomposition (cfd_b200/dist.py) on CPU: partition arithmetic, and the halo
exchange / reduction plumbing over torch.distributed with the gloo backend, world sizes 2 and 3."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cfd_b200.dist import HALO_E, HALO_W, TorchDistComm, halo_ops, partition, slab_extent


def test_partition_covers_everything_once():
    for ni, world in ((12, 1), (12, 2), (4098, 8), (130, 3), (67, 5)):
        parts = partition(ni, world)
        assert parts[0][0] == 1 and parts[-1][1] == ni
        for (a, b), (c, d) in zip(parts, parts[1:]):
            assert c == b + 1 and b >= a
        assert max(b - a for a, b in parts) - min(b - a for a, b in parts) <= 1


def test_slab_extent_halos():
    ni, world = 130, 4
    for r in range(world):
        goff, nloc, hw, he = slab_extent(ni, world, r)
        lo, hi = partition(ni, world)[r]
        assert hw == (0 if r == 0 else HALO_W) and he == (0 if r == world - 1 else HALO_E)
        assert goff == lo - 1 - hw and nloc == hi - lo + 1 + hw + he
    with pytest.raises(ValueError):
        slab_extent(6, 4, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ni, ld):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = TorchDistComm()
        rng = np.random.default_rng(7)
        glob = torch.from_numpy(rng.random((ni, ld)))
        goff, nloc, hw, he = slab_extent(ni, world, rank)
        lo, hi = partition(ni, world)[rank]
        nown = hi - lo + 1
        for (w, e) in ((1, 1), (2, 1), (1, 0), (0, 1)):
            loc = torch.full((nloc, ld), -1.0, dtype=torch.float64)
            loc[hw:hw + nown] = glob[lo - 1:hi]                    # owned lines only; halos unknown
            comm.exchange(halo_ops(loc, rank, world, hw, he, nown, w, e))
            want = glob[goff:goff + nloc].clone()
            mask = torch.zeros(nloc, dtype=torch.bool)
            mask[hw:hw + nown] = True
            if rank > 0:
                mask[hw - min(w, hw):hw] = True
            if rank < world - 1:
                mask[hw + nown:hw + nown + min(e, he)] = True
            assert torch.equal(loc[mask], want[mask]), (rank, w, e)
            assert bool((loc[~mask] == -1.0).all())               # nothing else was touched
        r = torch.tensor([float(rank + 1)] * 8, dtype=torch.float64)
        comm.allreduce_sum(r)
        assert r[0].item() == world * (world + 1) / 2
        b = torch.tensor([3.25 if rank == world - 1 else 0.0], dtype=torch.float64)
        comm.broadcast(b, world - 1)
        assert b.item() == 3.25
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_halo_exchange_gloo(world):
    mp.spawn(_worker, args=(world, _free_port(), 23, 16), nprocs=world, join=True)


def test_heat_slab_extent():
    from cfd_b200.dist import heat_slab_extent
    ny, world = 97, 4
    tot = 0
    for r in range(world):
        goff, nloc, hs, hn = heat_slab_extent(ny, world, r)
        assert hs == (r > 0) and hn == (r < world - 1)
        tot += nloc - hs - hn
    assert tot == ny


def test_colour_filter_sender_receiver_agree():
    from cfd_b200.dist import colour_filter
    for ni, world in ((40, 3), (4098, 8), (67, 5)):
        parts = partition(ni, world)
        for istart in (2, 3):
            for colour in (0, 1):
                for (alo, ahi), (blo, bhi) in zip(parts, parts[1:]):
                    fa, fb = colour_filter(alo, ahi, istart, colour), colour_filter(blo, bhi, istart, colour)
                    assert fa[3] == fb[0]      # a's last line changed  <=>  b's west halo is stale
                    assert fa[1] == fb[2]      # a's east halo is stale <=>  b's first line changed
