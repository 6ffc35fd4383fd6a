"""Slab-decomposed SIMPLE iterations (cfd_b200/dist.py) against the single-context run.  Red/black line order
makes the iterates independent of the partition, so fields must agree bit for bit; only the residual sums are
reduced in a different order."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("shape", [(64, 40), (30, 130)])
def test_local_slabs_match_single_context(cfd, world, shape):
    from cfd_b200.dist import LocalSlabCavity
    g = cfd.CavityFVM(*shape, order="redblack")
    g.init_lid_driven(reynolds=100.0)
    s = LocalSlabCavity(*shape, world, reynolds=100.0)
    for _ in range(3):
        r1 = g.iterate(2)
        r2 = s.iterate(2)
        for a, b in zip(r1, r2):
            assert abs(a - b) <= 1e-12 * abs(a)
    for name in ("u", "v", "p", "pp"):
        assert np.array_equal(s.gather_field(name), g.download(name)), name
    s.close()
    g.close()


@pytest.mark.parametrize("world,shape", [(2, (40, 34)), (3, (66, 50)), (4, (130, 96))])
def test_teach_local_slabs_match_single_context(cfd, world, shape):
    """Backward-step family cut into i-slabs: same bits as the single red/black context (fields and residuals up to
    the order of the residual sums), including the outlet mass balance on the last slab and the step-wall hooks."""
    from cfd_b200.dist import LocalSlabCavity, TeachSlab
    ni, nj = shape
    g = cfd.TeachFVM(ni, nj, order="redblack")
    s = LocalSlabCavity(ni - 2, nj - 2, world, slab_cls=TeachSlab)
    for _ in range(3):
        r1 = g.iterate(3)
        r2 = s.iterate(3)
        for a, b in zip(r1, r2):
            assert abs(a - b) <= 1e-11 * abs(a)
    for name in ("u", "v", "p", "pp"):
        assert np.array_equal(s.gather_field(name), g.download(name)), name
    s.close()
    g.close()


@pytest.mark.parametrize("world,shape", [(2, (40, 34)), (3, (47, 66)), (4, (64, 130)), (5, (98, 2050))])
def test_turbulent_teach_local_slabs_match_single_context(cfd, world, shape):
    """The backward step as shipped (k-epsilon + energy + props, control.f90:210-217) cut into i-slabs: every field,
    the wall arrays along the top wall and the six residuals agree with the single red/black context -- fields and
    wall arrays bit for bit (per-cell arithmetic does not depend on the partition; yplusn beside a slab is restated
    by the neighbour from the owner's operands), residuals up to the order of the sums."""
    from cfd_b200.dist import LocalSlabCavity, TeachSlab
    ni, nj = shape
    g = cfd.TeachFVM(ni, nj, order="redblack", turbulent=True)
    s = LocalSlabCavity(ni - 2, nj - 2, world, slab_cls=TeachSlab, turbulent=True)
    for _ in range(3):
        r1 = g.iterate_turb(2)
        s.iterate(2)
        r2 = s.residuals()[:6]
        for a, b in zip(r1, r2):
            assert abs(a - b) <= 1e-11 * abs(a), (r1, r2)
    for name in ("u", "v", "p", "pp", "te", "ed", "t", "vis", "gamh"):
        assert np.array_equal(s.gather_field(name), g.download(name)), name
    for name in ("taun", "yplusn"):
        assert np.array_equal(s.gather_wall(name), g.wall_download(name)), name
    owner = [x for x in s.slabs if x.lo <= g.istep <= x.hi][0]
    for name in ("tauw", "xplusw"):
        assert np.array_equal(owner.ctx.wall_download(name), g.wall_download(name)), name
    s.close()
    g.close()


@pytest.mark.parametrize("world,shape", [(2, (30, 34)), (3, (40, 66)), (4, (64, 130)), (8, (34, 96))])
def test_local_jslabs_partitioned_thomas(cfd, world, shape):
    """Lines cut ACROSS contexts (j-slabs): piece solves + spike vectors + reduced system reproduce the single
    context's line solves up to rounding (the partition method reorders the elimination)."""
    from cfd_b200.dist import LocalJSlabCavity
    g = cfd.CavityFVM(*shape, order="redblack")
    g.init_lid_driven(reynolds=100.0)
    s = LocalJSlabCavity(*shape, world, reynolds=100.0)
    for _ in range(4):
        r1 = g.iterate(3)
        r2 = s.iterate(3)
        for a, b in zip(r1, r2):
            assert abs(a - b) <= 1e-9 * abs(a), (r1, r2)
    for name in ("u", "v", "p", "pp"):
        a, b = s.gather_field(name), g.download(name)
        assert np.abs(a - b).max() <= 1e-10 * max(1.0, np.abs(b).max()), (name, np.abs(a - b).max())
    s.close()
    g.close()


@pytest.mark.parametrize("world,n", [(2, 34), (3, 50), (4, 66), (2, 2100)])
def test_buoyancy_local_slabs_match_single_context(cfd, world, n):
    """General family (Boussinesq cavity with its temperature equation) cut into i-slabs: same bits as the single
    red/black context.  n = 2100 has lines long enough for the TMA-staged assembly kernels (general_tma_kernels.cu)."""
    from cfd_b200.dist import BuoyancySlab, LocalSlabCavity
    g = cfd.BuoyantCavity(n, n, order="redblack")
    s = LocalSlabCavity(n - 2, n - 2, world, slab_cls=BuoyancySlab)
    for _ in range(3):
        r1 = g.iterate(3)
        r2 = s.iterate(3)
        for a, b in zip(r1[:3], r2):
            assert abs(a - b) <= 1e-11 * abs(a) + 1e-300
    for name in ("u", "v", "p", "pp", "t"):
        assert np.array_equal(s.gather_field(name), g.download(name)), name
    s.close()
    g.close()


def _nccl_worker(rank, world, port, shape, q, kind="i", peer=True):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), CFD_B200_PEER="1" if peer else "0")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from cfd_b200.dist import BuoyancySlab, JSlabCavity, SlabCavity, TeachSlab
        if kind == "buoyancy":
            s = SlabCavity(shape[0] - 2, shape[1] - 2, slab_cls=BuoyancySlab)
        elif kind == "j":
            s = JSlabCavity(*shape, reynolds=100.0)
        elif kind == "teach":
            s = SlabCavity(shape[0] - 2, shape[1] - 2, slab_cls=TeachSlab)
        elif kind == "turb":
            s = SlabCavity(shape[0] - 2, shape[1] - 2, slab_cls=TeachSlab, turbulent=True)
        else:
            s = SlabCavity(*shape, reynolds=100.0)
        res = s.iterate(6)
        if kind == "turb":
            res = s.residuals()[:6]
        out = {n: s.gather_field(n) for n in (("u", "v", "p", "te", "ed", "t", "vis") if kind == "turb" else ("u", "v", "p"))}
        if kind == "turb":
            out["taun"] = s.gather_wall("taun")
        used_peer = bool(getattr(s, "peer", False))
        timed_out = s.ctx.peer_status()
        if rank == 0:
            q.put((res, out, used_peer, timed_out))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,peer,shape", [("i", False, (96, 130)), ("j", False, (96, 130)), ("teach", False, (96, 130)),
                                             ("buoyancy", False, (98, 98)), ("i", True, (96, 2048)),
                                             ("teach", True, (98, 2050)), ("buoyancy", True, (2050, 2050)),
                                             ("turb", False, (96, 130)), ("turb", True, (98, 2050)),
                                             ("icluster", True, (96, 8192))])
def test_nccl_slabs_match_single_context(cfd, kind, peer, shape, monkeypatch):
    """One process per GPU.  peer=True: halo lines over peer memory (stores from inside the flag-ordered sweep kernels
    and cfd_halo_push) -- lines of 2050 nodes (160 layout threads x 13 nodes), long enough for the TMA-staged kernels that path
    lives in; peer=False:
    the grouped NCCL send/recv stages."""
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs (run under gpurun --gpus 2)")
    if kind == "icluster":      # lines of 8194 nodes solved by clusters of two thread blocks, in the slabs AND in the single context
        monkeypatch.setenv("CFD_B200_CLUSTER", "1")       # (the spawned ranks inherit it)
        kind = "i"
    if kind == "buoyancy":
        g = cfd.BuoyantCavity(*shape, order="redblack")
    elif kind == "teach":
        g = cfd.TeachFVM(*shape, order="redblack")
    elif kind == "turb":
        g = cfd.TeachFVM(*shape, order="redblack", turbulent=True)
    else:
        g = cfd.CavityFVM(*shape, order="redblack")
        g.init_lid_driven(reynolds=100.0)
    if kind == "turb":
        want_res = g.iterate_turb(6)
        want = {n: g.download(n) for n in ("u", "v", "p", "te", "ed", "t", "vis")}
        want["taun"] = g.wall_download("taun")
    else:
        want_res = g.iterate(6)
        want = {n: g.download(n) for n in ("u", "v", "p")}
    g.close()
    sk = socket.socket()
    sk.bind(("127.0.0.1", 0))
    port = sk.getsockname()[1]
    sk.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, shape, q, kind, peer)) for r in range(world)]
    for p in procs:
        p.start()
    res, out, used_peer, timed_out = q.get(timeout=300)
    assert not timed_out
    if peer and kind != "j":
        assert used_peer, "the peer-memory path was asked for but the ranks fell back to NCCL"
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for n in want:
        if kind == "j":      # the partition method reorders the elimination: equal up to rounding
            assert np.abs(out[n] - want[n]).max() <= 1e-10 * max(1.0, np.abs(want[n]).max()), n
        else:
            assert np.array_equal(out[n], want[n]), n
    for a, b in zip(res, want_res):
        assert abs(a - b) <= (1e-9 if kind == "j" else 1e-11) * abs(b)


@pytest.mark.parametrize("world", [2, 3])
def test_heat_row_slabs_match_single_grid(cfd, world):
    from cfd_b200.dist import LocalSlabHeat
    nx, ny = 130, 97
    t0 = np.asfortranarray(np.random.default_rng(3).random((nx, ny)))
    h = cfd.Heat2D(nx, ny, tleft=1.0, ttop=2.0)
    h.upload(t0)
    h.step(7)
    s = LocalSlabHeat(nx, ny, world, t0, tleft=1.0, ttop=2.0)
    s.step(7)
    assert np.array_equal(s.gather(), h.download())
    s.close()
    h.close()


@pytest.mark.parametrize("world,fo", [(2, 5.0), (3, 1.0), (4, 0.4)])
def test_implicit_heat_row_slabs_match_single_grid(cfd, world, fo):
    """Implicit (LOD) steps on row slabs: the y-lines cross the slabs, each slab carries cfd_heat_implicit_halo rows of
    halo (the distance over which the Thomas recurrences forget their start: m^W < 2^-70) refreshed before every step.
    Equal to the whole-grid production step to rounding (north_star: 1e-12 relative per step), and to the exact
    (one thread per line, reference operation order) mode likewise."""
    from cfd_b200.dist import LocalSlabHeat, implicit_halo_rows
    nx, ny = 1024, 1024
    kw = dict(fo=fo, tleft=1.0, ttop=2.0, tbottom=0.5)
    t0 = np.asfortranarray(np.random.default_rng(5).random((nx, ny)))
    halo = implicit_halo_rows(nx, ny, **kw)
    assert 32 <= halo <= ny // world
    h = cfd.Heat2D(nx, ny, **kw)
    h.upload(t0)
    hx = cfd.Heat2D(nx, ny, **kw)
    hx.upload(t0)
    s = LocalSlabHeat(nx, ny, world, t0, halo=halo, **kw)
    for _ in range(3):
        h.implicit_step(2)
        hx.implicit_step(2, exact=True)
        s.implicit_step(2)
        want, got = h.download(), s.gather()
        scale = np.abs(want).max()
        assert np.abs(got - want).max() <= 1e-12 * scale
        assert np.abs(got - hx.download()).max() <= 1e-12 * scale
    s.close()
    h.close()
    hx.close()


def test_implicit_heat_slab_refuses_a_shallow_halo(cfd):
    from cfd_b200.dist import HeatSlab
    s = HeatSlab(512, 512, 1, 2, halo=1, fo=5.0)
    with pytest.raises(Exception):
        s.h.implicit_step(1)
    s.h.close()


def _heat_lod_worker(rank, world, port, q):
    import os
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from cfd_b200.dist import SlabHeat, implicit_halo_rows
        nx = ny = 1024
        kw = dict(fo=2.0, tleft=1.0, ttop=2.0)
        t0 = np.asfortranarray(np.random.default_rng(9).random((nx, ny)))
        s = SlabHeat(nx, ny, t0=t0, halo=implicit_halo_rows(nx, ny, **kw), **kw)
        s.implicit_step(4)
        torch.cuda.synchronize()
        part = np.ascontiguousarray(s.slab.owned(s.slab.h.download()))
        out = [None] * world if rank == 0 else None
        dist.gather_object(part, out, dst=0)
        if rank == 0:
            q.put(np.asfortranarray(np.concatenate(out, axis=1)))
    finally:
        dist.destroy_process_group()


def test_implicit_heat_row_slabs_over_nccl(cfd):
    """One process per GPU: implicit (LOD) steps on row slabs with the deep halo exchanged over NCCL."""
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least 2 GPUs (run under gpurun --gpus 2)")
    nx = ny = 1024
    h = cfd.Heat2D(nx, ny, fo=2.0, tleft=1.0, ttop=2.0)
    h.upload(np.asfortranarray(np.random.default_rng(9).random((nx, ny))))
    h.implicit_step(4)
    want = h.download()
    h.close()
    sk = socket.socket()
    sk.bind(("127.0.0.1", 0))
    port = sk.getsockname()[1]
    sk.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_heat_lod_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
