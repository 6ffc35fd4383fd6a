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


def test_jslab_extent_and_line_counts():
    from cfd_b200.dist import JHALO_N, JHALO_S, _nlines, jslab_extent
    nj, world = 130, 4
    for r in range(world):
        joff, nloc, hs, hn = jslab_extent(nj, world, r)
        lo, hi = partition(nj, world)[r]
        assert hs == (0 if r == 0 else JHALO_S) and hn == (0 if r == world - 1 else JHALO_N)
        assert joff == lo - 1 - hs and nloc == hi - lo + 1 + hs + hn
    with pytest.raises(ValueError):
        jslab_extent(9, 4, 1)
    # lines istart..ni-1 of one colour: the count every rank must agree on for the all-gathered buffers
    for ni in (12, 13, 4098):
        for istart in (2, 3):
            assert _nlines(ni, istart, 0) + _nlines(ni, istart, 1) == ni - istart
            assert _nlines(ni, istart, 0) == len([i for i in range(istart, ni) if (i - istart) % 2 == 0])


@pytest.mark.parametrize("world", [2, 3, 8])
def test_partition_method_algebra(world):
    """The reduced system the j-slab driver all-gathers (dist.py / jslab_kernels.cu), in numpy on one line:
    piece solutions y with zero cut ends, spike vectors w, z, 2*world interface unknowns, x = y + w*x_s + z*x_n."""
    rng = np.random.default_rng(world)
    n = 8 * world + 5
    an, as_ = rng.random(n) * 0.4, rng.random(n) * 0.4          # x_j = an x_{j+1} + as x_{j-1} + c   (ap = 1)
    c = rng.random(n)
    bs, bn = 0.7, -0.3                                          # Dirichlet values below / above the line

    def solve(lo, hi, rhs, xs, xn):                             # unknowns lo..hi-1 with end values xs, xn
        m = hi - lo
        A = np.eye(m)
        b = rhs[lo:hi].copy()
        for k in range(m):
            j = lo + k
            if k > 0:
                A[k, k - 1] = -as_[j]
            else:
                b[k] += as_[j] * xs
            if k < m - 1:
                A[k, k + 1] = -an[j]
            else:
                b[k] += an[j] * xn
        return np.linalg.solve(A, b)

    want = solve(0, n, c, bs, bn)
    cuts = [p[0] - 1 for p in partition(n, world)] + [n]
    zero = np.zeros(n)
    Y, W, Z = [], [], []
    for r in range(world):
        lo, hi = cuts[r], cuts[r + 1]
        Y.append(solve(lo, hi, c, bs if r == 0 else 0.0, bn if r == world - 1 else 0.0))
        W.append(solve(lo, hi, zero, 1.0, 0.0))
        Z.append(solve(lo, hi, zero, 0.0, 1.0))
    N = 2 * world
    A, b = np.eye(N), np.zeros(N)
    for r in range(world):
        lo_, hi_ = 2 * r, 2 * r + 1
        if r > 0:
            A[lo_, hi_ - 2], A[hi_, hi_ - 2] = -W[r][0], -W[r][-1]
        if r < world - 1:
            A[lo_, lo_ + 2], A[hi_, lo_ + 2] = -Z[r][0], -Z[r][-1]
        b[lo_], b[hi_] = Y[r][0], Y[r][-1]
    x = np.linalg.solve(A, b)
    got = np.concatenate([Y[r] + (W[r] * x[2 * (r - 1) + 1] if r > 0 else 0.0) + (Z[r] * x[2 * (r + 1)] if r < world - 1 else 0.0)
                          for r in range(world)])
    assert np.abs(got - want).max() < 1e-13


# ---- j-slab driver plumbing over gloo: halo ROWS (packed / unpacked through the context) and the all-gathers ------------
class _StubCtx:
    """Stands in for the CUDA context: fields are host arrays (ni, nj_local), rows_pack / rows_unpack move rows
    j_first .. j_first+n-1 (1-based) to / from a [n, ni] buffer given by address, like cfd_rows_pack / cfd_rows_unpack."""

    def __init__(self, fields):
        self.f = fields

    def rows_pack(self, name, j_first, nrows, ptr):
        import ctypes
        rows = np.ascontiguousarray(self.f[name][:, j_first - 1:j_first - 1 + nrows].T)
        ctypes.memmove(ptr, rows.ctypes.data, rows.nbytes)

    def rows_unpack(self, name, j_first, nrows, ptr):
        import ctypes
        ni = self.f[name].shape[0]
        buf = np.empty((nrows, ni))
        ctypes.memmove(buf.ctypes.data, ptr, buf.nbytes)
        self.f[name][:, j_first - 1:j_first - 1 + nrows] = buf.T


def _jslab_worker(rank, world, port, ni, nj):
    from types import SimpleNamespace
    from cfd_b200.dist import JSlabCavity, jslab_extent
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(11)
        glob = {n: rng.random((ni, nj)) for n in ("u", "p", "dv")}
        joff, njl, hs, hn = jslab_extent(nj, world, rank)
        lo, hi = partition(nj, world)[rank]
        nown = hi - lo + 1
        loc = {n: np.full((ni, njl), -1.0) for n in glob}
        for n in glob:
            loc[n][:, hs:hs + nown] = glob[n][:, lo - 1:hi]            # owned rows only; halo rows unknown
        drv = object.__new__(JSlabCavity)                              # the driver's communication methods, no CUDA
        drv.torch, drv.dist, drv.comm, drv._device = torch, dist, TorchDistComm(), "cpu"
        drv.slab = SimpleNamespace(rank=rank, world=world, ni=ni, hs=hs, hn=hn, nown=nown, ctx=_StubCtx(loc))
        drv.rows((("u", 1, 1), ("p", 1, 0)))
        drv.rows((("dv", 0, 1),))
        for n, (ns, nn) in (("u", (1, 1)), ("p", (1, 0)), ("dv", (0, 1))):
            want = np.full((ni, njl), -1.0)
            want[:, hs:hs + nown] = glob[n][:, lo - 1:hi]
            if rank > 0 and ns:
                want[:, hs - 1] = glob[n][:, lo - 2]
            if rank < world - 1 and nn:
                want[:, hs + nown] = glob[n][:, hi]
            assert np.array_equal(loc[n], want), (rank, n)
        # the two all-gathers of the partition method: [ni, 4] spike ends, [nlines, 2] piece ends, in rank order
        wz_local = torch.full((ni, 4), float(rank + 1), dtype=torch.float64)
        wz_all = torch.zeros((world, ni, 4), dtype=torch.float64)
        y_local = torch.arange(10, dtype=torch.float64).reshape(5, 2) + 100 * rank
        y_all = torch.zeros((world, 5, 2), dtype=torch.float64)
        drv.gather([(wz_local, wz_all), (y_local, y_all)])
        for r in range(world):
            assert bool((wz_all[r] == r + 1).all()) and y_all[r, 0, 0].item() == 100 * r
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_jslab_rows_and_gathers_gloo(world):
    mp.spawn(_jslab_worker, args=(world, _free_port(), 9, 31), nprocs=world, join=True)


def test_iteration_stage_list_orders_exchanges_and_sweeps():
    """The stage list every slab driver walks (host logic only): halos before the u assembly, du exchanged before the
    u sweeps, 2*nswp colour passes per equation with pp = 0 folded into the first pressure passes, the reference
    pressure before the correction, and -- general family -- the temperature equation after the correction."""
    from cfd_b200.dist import iteration_stages
    nswp = (3, 2, 5, 4)
    st = iteration_stages(nswp, solve_t=True)
    kinds = [s[0] for s in st]
    assert kinds[0] == "halos" and {f[0] for f in st[0][1]} == {"u", "v", "p"}
    assert kinds[1] == "compute" and st[2][:2] == ("halo", "du")
    sweeps = [s for s in st if s[0] == "sweep"]
    for name, n in zip(("u", "v", "pp", "t"), nswp):
        mine = [s for s in sweeps if s[1] == name]
        assert len(mine) == 2 * n
        assert [s[4] for s in mine] == [0, 1] * n                                   # red, black, red, black ...
        assert [s[5] for s in mine] == ([True, True] + [False] * (2 * n - 2) if name == "pp" else [False] * (2 * n))
    assert [s[2:4] for s in sweeps if s[1] == "u"][0] == (3, 2) and [s[2:4] for s in sweeps if s[1] == "v"][0] == (2, 3)
    i_ppref = kinds.index("ppref")
    assert kinds[i_ppref + 1] == "compute"                                          # the correction
    assert max(i for i, s in enumerate(st) if s[0] == "sweep" and s[1] == "pp") < i_ppref
    assert st[i_ppref + 2][0] == "halos" and {f[0] for f in st[i_ppref + 2][1]} == {"u", "t"}
    assert min(i for i, s in enumerate(st) if s[0] == "sweep" and s[1] == "t") > i_ppref + 3
    cav = iteration_stages(nswp[:3])
    assert [s for s in cav if s[0] == "sweep" and s[1] == "t"] == [] and cav[-1][0] == "compute"


def test_turbulent_stage_list():
    """TEACH with k-epsilon + energy (control.f90:210-217): after the correction the corrected velocities beside the
    slab are exchanged once, then calcte, calced, calct each assemble and sweep (the switched-off ones vanish), then
    props and the exchange of the two property fields it rewrites; the peer-memory form keeps the same order."""
    from types import SimpleNamespace
    from cfd_b200.dist import iteration_stages, peer_stages
    turb = SimpleNamespace(incalk=1, incald=1, incalt=0, inpro=1, nswpk=3, nswpd=2, nswpt=4)
    st = iteration_stages((3, 3, 5), turb=turb)
    kinds = [s[0] for s in st]
    i_ppref = kinds.index("ppref")
    tail = st[i_ppref + 2:]
    assert tail[0][0] == "halos" and {f for f in tail[0][1]} == {("u", 0, 1), ("v", 1, 1)}
    names = [s[1] for s in tail if s[0] == "sweep"]
    assert names == ["te"] * 6 + ["ed"] * 4                      # incalt = 0: no temperature passes
    assert [s[2:4] for s in tail if s[0] == "sweep"][0] == (2, 2)
    assert tail[-2][0] == "compute" and tail[-1] == ("halos", (("vis", 1, 1), ("gamh", 1, 1)))
    assert sum(1 for s in tail if s[0] == "compute") == 3         # assemble_te, assemble_ed, props
    ps = peer_stages(st)
    assert [s[1] for s in ps if s[0] == "sweeps"] == ["u", "v", "pp", "te", "ed"]
    assert [s[4] for s in ps if s[0] == "sweeps"] == [3, 3, 5, 3, 2]
    assert ps[-1] == ("push", (("vis", 1, 1), ("gamh", 1, 1)))
    turb.incalt, turb.inpro = 1, 0
    st = iteration_stages((3, 3, 5), turb=turb)
    assert [s[1] for s in st if s[0] == "sweep"][-8:] == ["t"] * 8 and st[-1][0] == "sweep"


def _heat_halo_worker(rank, world, port, ny, nx, halo):
    """HeatSlab.halo_ops with a deep halo (implicit steps) over gloo: host tensors stand in for the device rows."""
    from cfd_b200.dist import HeatSlab, heat_slab_extent
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        comm = TorchDistComm()
        glob = torch.from_numpy(np.random.default_rng(11).random((ny, nx)))
        s = HeatSlab.__new__(HeatSlab)                              # bookkeeping only: no GPU context
        s.rank, s.world = rank, world
        s.j_goff, s.ny, s.hs, s.hn = heat_slab_extent(ny, world, rank, halo)
        s.nown = s.ny - s.hs - s.hn
        loc = torch.full((s.ny, nx), -1.0, dtype=torch.float64)
        loc[s.hs:s.hs + s.nown] = glob[s.j_goff + s.hs:s.j_goff + s.hs + s.nown]
        s.rows = lambda: loc
        comm.exchange(s.halo_ops())
        assert torch.equal(loc, glob[s.j_goff:s.j_goff + s.ny]), rank     # every halo row filled from its owner
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,halo", [(2, 1), (2, 32), (3, 17)])
def test_heat_deep_halo_exchange_gloo(world, halo):
    mp.spawn(_heat_halo_worker, args=(world, _free_port(), 160, 12, halo), nprocs=world, join=True)


def test_heat_slab_extent_deep_halo():
    from cfd_b200.dist import heat_slab_extent
    assert heat_slab_extent(400, 2, 0, 128) == (0, 328, 0, 128)
    assert heat_slab_extent(400, 2, 1, 128) == (72, 328, 128, 0)
    with pytest.raises(ValueError):
        heat_slab_extent(100, 4, 1, 64)                             # a slab must own at least the halo depth
