"""Multi-GPU driver for the cavity family: 1-D slab decomposition along i.

The S-N TDMA lines of lisolv (cavity_fvm.f90:62-80) run along j, so cutting the grid into slabs of whole lines
keeps every tridiagonal system on one GPU; only the line Gauss-Seidel coupling crosses ranks, and with
red/black line ordering that is one halo line of phi per colour pass.  In the device line layout a halo is a
whole contiguous line (LD doubles) -- no packing.  Per outer iteration (cavity_fvm.f90:554-557):

    exchange u,v,p -> assemble_u -> 3 x (colour 0 | exchange u | colour 1 | exchange u) -> exchange du
    assemble_v -> 3 x (... v ...) -> assemble_p -> 3 x (... pp ...) -> broadcast pp(ipref,jpref) -> correct
    residual sums: one all-reduce of 8 packed doubles, only when the host asks for them.

One process per GPU (torch.distributed, NCCL over NVLink; gloo for the CPU tests of this plumbing), or -- for
single-GPU tests of the slab kernels -- several slabs in one process with `LocalComm`.  The sequential-order
debug mode does not shard (replicas only).
"""
from __future__ import annotations

import os

import numpy as np

from .cavity import CavityFVM, setgrid

HALO_W, HALO_E = 2, 1        # calcu reads den(i-2) (cavity_fvm.f90:113); everything else needs one line


def partition(ni_global: int, world: int):
    """Owned global line ranges [lo, hi] (1-based, inclusive) per rank: near-equal contiguous slabs."""
    base, rem = divmod(ni_global, world)
    out, lo = [], 1
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((lo, lo + n - 1))
        lo += n
    return out


def slab_extent(ni_global: int, world: int, rank: int):
    """(goff, ni_local, halo_w, halo_e) of rank's local arrays: owned lines plus halos, clipped at the walls."""
    lo, hi = partition(ni_global, world)[rank]
    hw = min(HALO_W, lo - 1) if rank > 0 else 0
    he = min(HALO_E, ni_global - hi) if rank < world - 1 else 0
    if world > 1 and hi - lo + 1 < HALO_W:
        raise ValueError(f"slab of rank {rank} owns {hi - lo + 1} lines; every slab needs at least {HALO_W}")
    return lo - 1 - hw, (hi - lo + 1) + hw + he, hw, he


class _DevArray:
    """Expose a raw device range to torch through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f8", "data": (ptr, False), "version": 3}


def device_tensor(ptr: int, shape):
    import torch
    return torch.as_tensor(_DevArray(ptr, shape), device="cuda")


# ----------------------------------------------------------------------------------------------------------
# communicators
# ----------------------------------------------------------------------------------------------------------
class TorchDistComm:
    """Halo exchange / reductions over torch.distributed (NCCL on GPUs, gloo on CPU tensors)."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

    def exchange(self, ops):
        """ops: list of (kind, tensor, peer) with kind 'send' | 'recv'; one grouped batch."""
        if not ops:
            return
        dist = self.dist
        reqs = dist.batch_isend_irecv([dist.P2POp(dist.isend if k == "send" else dist.irecv, t, p) for k, t, p in ops])
        for r in reqs:
            r.wait()

    def allreduce_sum(self, t):
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)

    def broadcast(self, t, src):
        self.dist.broadcast(t, src=src)

    def allgather(self, dst, src):
        """dst[r] = rank r's src (dst contiguous, [world, ...]).  One fused call over NCCL; gloo has no tensor form."""
        if self.dist.get_backend() == "nccl":
            self.dist.all_gather_into_tensor(dst, src)
        else:
            self.dist.all_gather(list(dst.unbind(0)), src)


def colour_filter(lo: int, hi: int, istart: int, colour: int):
    """After the colour pass `colour` of a sweep starting at line istart only lines with (i-istart)%2 == colour
    changed.  Returns (west_halo_stale, east_halo_stale, my_first_changed, my_last_changed) for a slab owning
    global lines lo..hi: which of the four single-line messages are needed at all."""
    def changed(i):
        return i >= istart and (i - istart) % 2 == colour
    return changed(lo - 1), changed(hi + 1), changed(lo), changed(hi)


def halo_ops(lines, rank: int, world: int, hw: int, he: int, nown: int, w: int = 1, e: int = 1, only=None):
    """P2P ops that refresh `w` west and `e` east halo lines of `lines` (tensor [ni_local, LD]).

    Local line layout: [hw halo lines | nown owned lines | he halo lines].  My west halo = the last lines my
    west neighbour owns; my east halo = the first lines my east neighbour owns.
    """
    ops = []
    first, last = hw, hw + nown - 1
    rw, re_, sf, sl = only if only is not None else (True, True, True, True)
    if rank > 0:
        w_eff = min(w, hw)
        if w_eff and rw:
            ops.append(("recv", lines[first - w_eff:first], rank - 1))
        if e and sf:   # the west neighbour's east halo (it has one, since it has an east neighbour: me)
            ops.append(("send", lines[first:first + min(e, HALO_E)], rank - 1))
    if rank < world - 1:
        e_eff = min(e, he)
        if e_eff and re_:
            ops.append(("recv", lines[last + 1:last + 1 + e_eff], rank + 1))
        if w and sl:
            ops.append(("send", lines[last + 1 - min(w, HALO_W):last + 1], rank + 1))
    return ops


class LocalComm:
    """All ranks in ONE process (single-GPU tests of the slab kernels): copies instead of messages."""

    def __init__(self, world):
        self.world = world

    def exchange_all(self, slabs, name, w=1, e=1, colour=None):
        for r in range(self.world - 1):
            a, b = slabs[r], slabs[r + 1]          # a is west of b
            la, lb = a.lines(name), b.lines(name)
            a_last, b_first = a.hw + a.nown - 1, b.hw
            # same decisions the two ranks would take independently over messages
            fa = colour_filter(a.lo, a.hi, *colour) if colour else (True,) * 4
            fb = colour_filter(b.lo, b.hi, *colour) if colour else (True,) * 4
            assert fa[3] == fb[0] and fa[1] == fb[2]   # sender and receiver agree
            we = min(w, b.hw)
            if we and fb[0]:
                lb[b_first - we:b_first].copy_(la[a_last + 1 - we:a_last + 1])
            ee = min(e, a.he)
            if ee and fa[1]:
                la[a_last + 1:a_last + 1 + ee].copy_(lb[b_first:b_first + ee])


# ----------------------------------------------------------------------------------------------------------
class Slab:
    """One rank's share of the global cavity problem: a CavityFVM slab context plus its halo bookkeeping."""

    def __init__(self, nicv, njcv, rank, world, *, reynolds=100.0, utop=1.0, rho=1000.0, iadv=3,
                 urf=(0.5, 0.5, 0.8), nswp=(3, 3, 3), device=-1, xlim=(0.0, 1.0), ylim=(0.0, 1.0)):
        self.rank, self.world = rank, world
        ni_g, nj = nicv + 2, njcv + 2
        self.ni_global, self.nj = ni_g, nj
        self.goff, self.ni, self.hw, self.he = slab_extent(ni_g, world, rank)
        self.lo, self.hi = partition(ni_g, world)[rank]
        self.nown = self.hi - self.lo + 1
        g = setgrid(nicv, njcv, xlim[0], xlim[1], ylim[0], ylim[1])      # host logic, O(ni+nj)
        sl = slice(self.goff, self.goff + self.ni)
        self.ctx = CavityFVM(nicv, njcv, iadv=iadv, order="redblack", urf=urf, nswp=nswp, device=device,
                             ni_local=self.ni, goff=self.goff, halo_w=self.hw, halo_e=self.he,
                             sew_nim1=float(g["sew"][ni_g - 2]))
        self.ctx.set_geometry(x=g["x"][sl], xu=g["xu"][sl], sew=g["sew"][sl], sewu=g["sewu"][sl],
                              y=g["y"], yv=g["yv"], sns=g["sns"], snsv=g["snsv"])
        # cavity_fvm.f90:534-538 on the local lines (halo lines included: den, vis are static)
        den = np.full((self.ni, nj), rho, order="F")
        vis = np.asfortranarray(den * utop * (xlim[1] - xlim[0]) / reynolds)
        u = np.zeros((self.ni, nj), order="F")
        gi = np.arange(self.goff + 1, self.goff + self.ni + 1)            # global 1-based line numbers
        u[gi >= 2, nj - 1] = utop
        self.ctx.upload("den", den)
        self.ctx.upload("vis", vis)
        self.ctx.upload("u", u)
        self._finish(nswp)

    def _finish(self, nswp):
        self.nswp = nswp
        self.ipref, self.jpref = self.ctx.ipref, self.ctx.jpref
        self.owner_of_pref = self.lo <= self.ipref <= self.hi
        self._lines = {}
        import torch
        self.ppref = torch.zeros(1, dtype=torch.float64, device="cuda")
        self.res = device_tensor(self.ctx.residual_device_ptr(), (8,))

    def lines(self, name):
        """Field `name` as a torch tensor [ni_local, LD] aliasing the context's device memory."""
        if name not in self._lines:
            base, stride, n = self.ctx.device_view(name)
            self._lines[name] = device_tensor(base, (n, stride))
        return self._lines[name]

    def ppref_element(self):
        """1-element view of pp(ipref,jpref) (owner rank only)."""
        L = self.ctx.layout
        j0 = self.jpref - 1
        q = (j0 % L["M"]) * L["T"] + j0 // L["M"]
        return self.lines("pp")[self.ipref - 1 - self.goff, q:q + 1]

    def owned(self, a):
        """Owned part of a downloaded local array."""
        return a[self.hw:self.hw + self.nown]


class TeachSlab(Slab):
    """One rank's i-slab of the TEACH backward-step problem (BASELINE.json configs[4]: 16384 x 4096 over 8 GPUs).
    `nicv, njcv` are ni-2, nj-2 of the global grid, as for the cavity slabs.  The outlet mass balance of modu
    (promod.f90:79-94) lives entirely on the last rank (it owns the lines ni-2 .. ni), the inlet flow is a host
    constant, so the stage list is the cavity's: no extra exchange."""

    def __init__(self, nicv, njcv, rank, world, *, device=-1, nswp=(3, 3, 5), turbulent=False, **kw):
        from .teach import TeachFVM
        self.rank, self.world = rank, world
        ni_g, nj = nicv + 2, njcv + 2
        self.ni_global, self.nj = ni_g, nj
        self.goff, self.ni, self.hw, self.he = slab_extent(ni_g, world, rank)
        self.lo, self.hi = partition(ni_g, world)[rank]
        self.nown = self.hi - self.lo + 1
        if world > 1 and rank == world - 1 and self.nown < 3:
            raise ValueError("the last TEACH slab must own the lines ni-2 .. ni (outlet mass balance)")
        self.ctx = TeachFVM(ni_g, nj, order="redblack", device=device, nswp=nswp, ni_local=self.ni, goff=self.goff,
                            halo_w=self.hw, halo_e=self.he, turbulent=turbulent, **kw)
        # k-epsilon + energy as shipped (control.f90:210-217): the wall arrays taun, yplusn are per line and stay with
        # the slab (the entry beside the slab is restated by the neighbour, teach_turb_kernels.cu), tauw, xplusw live
        # with the slab that owns line istep; vis and gamh are fields now and get a halo exchange after props
        self.turb = self.ctx.model if turbulent else None
        self._finish(nswp)

    def wall_owned(self, name):
        """Owned part of a per-line wall array (taun, yplusn)."""
        return self.ctx.wall_download(name)[self.hw:self.hw + self.nown]


class BuoyancySlab(Slab):
    """One rank's i-slab of the Boussinesq cavity (general family, BASELINE.json configs[3]).  `nicv, njcv` = ni-2,
    nj-2 of the global (square) grid.  Adds the temperature equation to the stage list (`solve_t`)."""

    solve_t = True

    def __init__(self, nicv, njcv, rank, world, *, device=-1, nswp=(3, 3, 3, 3), **kw):
        from .general import BuoyantCavity
        self.rank, self.world = rank, world
        ni_g, nj = nicv + 2, njcv + 2
        self.ni_global, self.nj = ni_g, nj
        self.goff, self.ni, self.hw, self.he = slab_extent(ni_g, world, rank)
        self.lo, self.hi = partition(ni_g, world)[rank]
        self.nown = self.hi - self.lo + 1
        self.ctx = BuoyantCavity(ni_g, nj, order="redblack", device=device, nsweep=nswp, ni_local=self.ni, goff=self.goff,
                                 halo_w=self.hw, halo_e=self.he, **kw)
        self._finish(nswp)


# stages of one outer iteration: ("compute", fn(slab)), ("halo", field, w, e, colour filter), ("halos", ...),
# ("sweep", field, istart, jstart, colour, first) = colour pass + halo exchange of that colour's edge lines, ("ppref",)
def iteration_stages(nswp, solve_t=False, turb=None):
    st = [("halos", (("u", 1, 1), ("v", 1, 1), ("p", 1, 0))),
          ("compute", lambda s: s.ctx.assemble_u()), ("halo", "du", 0, 1, None)]
    for name, istart, jstart, n, first in (("u", 3, 2, nswp[0], False), ("v", 2, 3, nswp[1], False), ("pp", 2, 2, nswp[2], True)):
        if name == "v":
            st.append(("compute", lambda s: s.ctx.assemble_v()))
        if name == "pp":
            st.append(("compute", lambda s: s.ctx.assemble_p()))
        for k in range(n):
            for colour in (0, 1):
                # colour pass + exchange of the lines of this colour at the slab edges (only they moved)
                st.append(("sweep", name, istart, jstart, colour, first and k == 0))
    st += [("ppref",), ("compute", lambda s: s.ctx.correct(s.ppref.data_ptr()))]
    if solve_t:     # calct (buoyancy_cav.f90:358-433): reads the corrected u(i+1), t(i-1 .. i+1)
        st += [("halos", (("u", 0, 1), ("t", 1, 1))), ("compute", lambda s: s.ctx.assemble_t())]
        for k in range(nswp[3]):
            for colour in (0, 1):
                st.append(("sweep", "t", 2, 2, colour, False))
    if turb is not None:   # control.f90:213-217: calcte, calced, calct, props after calcp (TEACH, k-epsilon + energy)
        # the scalar equations read the corrected u(i+1), v(i-1), v(i+1); te, ed, t halos follow their own sweeps
        st.append(("halos", (("u", 0, 1), ("v", 1, 1))))
        for on, fn, name, n in ((turb.incalk, lambda s: s.ctx.assemble_te(), "te", turb.nswpk),
                                (turb.incald, lambda s: s.ctx.assemble_ed(), "ed", turb.nswpd),
                                (turb.incalt, lambda s: s.ctx.assemble_t(), "t", turb.nswpt)):
            if on:
                st.append(("compute", fn))
                for k in range(n):
                    for colour in (0, 1):
                        st.append(("sweep", name, 2, 2, colour, False))
        if turb.inpro:
            st += [("compute", lambda s: s.ctx.props()), ("halos", (("vis", 1, 1), ("gamh", 1, 1)))]
    return st


def peer_stages(stages):
    """The stage list of `iteration_stages` for slabs whose neighbours are attached over peer memory: grouped NCCL
    send/recv stages become one cfd_halo_push each, and the colour passes of an equation (with the exchange after each)
    become one cfd_sweeps launch that hands the edge lines over from inside the kernel."""
    out, k = [], 0
    while k < len(stages):
        st = stages[k]
        if st[0] == "sweep":
            name, istart, jstart, _, first = st[1:]
            n = 0
            while k < len(stages) and stages[k][0] == "sweep" and stages[k][1] == name:
                n += 1
                k += 1
            out.append(("sweeps", name, istart, jstart, n // 2, first))
            continue
        if st[0] == "ppref":
            out.append(st)
            out.append(("correct",))                     # the correction reads the broadcast value from the mailbox
            k += 2                                       # (skips the NCCL path's correct stage that follows)
            continue
        if st[0] == "halo":
            out.append(("push", ((st[1], st[2], st[3]),)))
        elif st[0] == "halos":
            out.append(("push", tuple(st[1])))
        else:
            out.append(st)
        k += 1
    return out


class _E2EMixin:
    def e2e(self, steps):
        """bench.py's end-to-end leg: per step upload u,v,p slabs from pinned host memory, one outer
        iteration, download u,v,p and the residuals; max over ranks of the wall time."""
        import time
        torch = self.torch
        ni, nj = self.slab.ni, self.slab.nj
        host = {n: torch.empty((nj, ni), dtype=torch.float64, pin_memory=True) for n in ("u", "v", "p")}
        for n in host:
            self.ctx.download_ptr(n, host[n].data_ptr(), ni)
        fb = ni * nj * 8
        cells = self._global_cells()

        def leg(full_download):
            t0 = 0.0
            for it in range(steps + 1):
                if it == 1:
                    torch.cuda.synchronize()
                    self.dist.barrier()
                    t0 = time.perf_counter()
                for n in host:
                    self.ctx.upload_ptr(n, host[n].data_ptr(), ni)
                self.iterate(1)                               # all-reduced residuals come back to the host
                if full_download:
                    for n in host:
                        self.ctx.download_ptr(n, host[n].data_ptr(), ni)
            torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
            self.dist.all_reduce(dt, op=self.dist.ReduceOp.MAX)
            return float(dt.item())

        dt, dt_full = leg(False), leg(True)
        return {"value": cells * steps / dt / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": 3 * fb * self.world,
                "d2h_bytes_per_step": 64 * self.world, "steps": steps, "ms_per_step": dt / steps * 1e3,
                "what": "per step and rank: cfd_upload(u,v,p slab) from pinned host arrays, one slab-decomposed outer "
                        "iteration with halo exchange, all-reduced residuals to the host",
                "roundtrip": {"value": cells * steps / dt_full / 1e9, "unit": "GLUPS",
                              "h2d_bytes_per_step": 3 * fb * self.world,
                              "d2h_bytes_per_step": (3 * fb + 64) * self.world, "ms_per_step": dt_full / steps * 1e3,
                              "what": "same, plus cfd_download(u,v,p slab) of the whole fields every step"}}



class SlabCavity(_E2EMixin):
    """SPMD driver: one Slab per process over torch.distributed (bench.py, torchrun)."""

    def __init__(self, nicv, njcv, *, reynolds=1000.0, stream=None, slab_cls=None, **kw):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.comm = TorchDistComm()
        self.rank, self.world = self.comm.rank, self.comm.world
        if slab_cls is None:
            slab_cls, kw = Slab, dict(kw, reynolds=reynolds)
        self.slab = slab_cls(nicv, njcv, self.rank, self.world, device=torch.cuda.current_device(), **kw)
        self.ctx = self.slab.ctx
        if stream is not None:
            self.ctx.set_stream(stream.cuda_stream)
        self.stages = iteration_stages(self.slab.nswp, getattr(self.slab, "solve_t", False), getattr(self.slab, "turb", None))
        self.peer = self._attach_peers()
        self.pstages = peer_stages(self.stages) if self.peer else None
        self.overlap = os.environ.get("CFD_B200_OVERLAP", "0") != "0"   # measured: no gain at N=2
        self._comm_stream = None
        self.pref_owner = [r for r, (lo, hi) in enumerate(partition(nicv + 2, self.world)) if lo <= self.slab.ipref <= hi][0]
        # static fields: halos filled at upload (den, vis are uniform here); general case: exchange once
        s = self.slab
        self.comm.exchange(halo_ops(s.lines("den"), s.rank, s.world, s.hw, s.he, s.nown, HALO_W, HALO_E))
        self.comm.exchange(halo_ops(s.lines("vis"), s.rank, s.world, s.hw, s.he, s.nown, 1, 1))
        if getattr(s, "turb", None) is not None:        # the first props ran on the owned lines only
            self.comm.exchange(halo_ops(s.lines("gamh"), s.rank, s.world, s.hw, s.he, s.nown, 1, 1))

    def _global_cells(self):
        return (self.slab.ni_global - 2) * (self.slab.nj - 2)

    def _attach_peers(self):
        """Map the neighbouring slabs' fields and flag blocks into this process (CUDA IPC over NVLink peer access), so
        that halo lines travel as plain stores from inside the kernels instead of NCCL messages issued by the host
        between kernels.  All ranks decide together: if any mapping fails everywhere falls back to the NCCL stages.
        CFD_B200_PEER=0 forces the NCCL path."""
        torch, dist, s, ctx = self.torch, self.dist, self.slab, self.ctx
        if self.world == 1:
            return False
        names = ("u", "v", "p", "pp", "du") + (("t",) if getattr(s, "solve_t", False) else ())
        if getattr(s, "turb", None) is not None:
            names += ("te", "ed", "t", "vis", "gamh")
        mine, ok = None, os.environ.get("CFD_B200_PEER", "1") != "0"
        if ok:
            try:
                mine = {"ni": s.ni, "hw": s.hw, "he": s.he, "host": os.uname().nodename,
                        "h": {n: ctx.ipc_export(n) for n in names + ("flags",)}}
            except Exception as e:                          # noqa: BLE001
                self.peer_error, ok = repr(e), False
        every = [None] * self.world
        dist.all_gather_object(every, mine if ok else None)
        ok = ok and all(x is not None and x["host"] == every[0]["host"] for x in every)
        if ok:
            try:
                opened = {}                                 # (rank, what) -> device pointer: every handle is opened once

                def peer_ptr(r, n):
                    if (r, n) not in opened:
                        opened[(r, n)] = ctx.ipc_open(every[r]["h"][n])
                    return opened[(r, n)]

                for side, r in ((0, self.rank - 1), (1, self.rank + 1)):
                    if 0 <= r < self.world:
                        for n in names + ("flags",):
                            ctx.peer_attach(side, n, peer_ptr(r, n), every[r]["ni"], every[r]["hw"], every[r]["he"])
                ok = ctx.sweeps_single_launch()             # short lines keep the pass-by-pass solver, which needs the NCCL stages
                if ok and s.owner_of_pref:                  # the owner of pp(ipref,jpref) reaches every rank's mailbox
                    for r in range(self.world):
                        ctx.bcast_attach(r, None if r == self.rank else peer_ptr(r, "flags"), every[r]["ni"])
            except Exception as e:                          # noqa: BLE001
                self.peer_error, ok = repr(e), False
        agree = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(agree, op=dist.ReduceOp.MIN)
        if int(agree.item()) != 1:
            ctx.peer_detach()
            return False
        torch.cuda.synchronize()
        dist.barrier()                                      # nobody pushes before everybody has attached
        return True

    def _iterate_peer(self, n=1):
        s = self.slab
        for _ in range(n):
            for st in self.pstages:
                if st[0] == "correct":
                    s.ctx.correct(self._ppref_dev)
                elif st[0] == "compute":
                    st[1](s)
                elif st[0] == "sweeps":
                    s.ctx.sweeps(st[2], st[3], st[1], st[4], first=st[5])
                elif st[0] == "push":
                    s.ctx.halo_push(st[1])
                elif st[0] == "ppref":
                    # pp(ipref,jpref) from its owner into every rank's mailbox (peer stores), then the correction
                    self._ppref_dev = s.ctx.scalar_bcast(s.owner_of_pref, "pp", s.ipref - s.goff, s.jpref)
                else:
                    raise ValueError(st[0])

    def enable_graph(self, stream):
        """Capture one outer iteration (kernels + NCCL halo exchanges) into a CUDA graph and replay it from then
        on: one launch per iteration instead of ~85 host calls.  Falls back to eager issue if capture fails."""
        torch = self.torch
        try:
            self._iterate_eager(2)                      # communicators, attributes, allocations: all warm
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                self._iterate_eager(1)
            torch.cuda.synchronize()
            self.graph = g
        except Exception as e:                          # noqa: BLE001
            self.graph = None
            self.graph_error = repr(e)
            torch.cuda.synchronize()
        return self.graph is not None

    def iterate_async(self, n=1):
        if getattr(self, "graph", None) is not None:
            for _ in range(n):
                self.graph.replay()
            return
        self._iterate_eager(n)

    def _iterate_eager(self, n=1):
        if self.peer:
            return self._iterate_peer(n)
        s = self.slab
        for _ in range(n):
            for st in self.stages:
                if st[0] == "compute":
                    st[1](s)
                elif st[0] == "sweep":
                    self._sweep_and_exchange(*st[1:])
                elif st[0] == "halo":
                    only = colour_filter(s.lo, s.hi, *st[4]) if st[4] else None
                    self.comm.exchange(halo_ops(s.lines(st[1]), s.rank, s.world, s.hw, s.he, s.nown, st[2], st[3], only))
                elif st[0] == "halos":
                    ops = []
                    for name, w, e in st[1]:
                        ops += halo_ops(s.lines(name), s.rank, s.world, s.hw, s.he, s.nown, w, e)
                    self.comm.exchange(ops)
                else:
                    if s.owner_of_pref:
                        s.ppref.copy_(s.ppref_element())
                    self.comm.broadcast(s.ppref, self.pref_owner)

    def _sweep_and_exchange(self, name, istart, jstart, colour, first):
        """One colour pass with the halo exchange hidden behind it: the slab-edge lines of the colour are solved
        first (cfd_sweep_part 1), their exchange runs on a side stream while the other lines are solved (part 2);
        the compute stream then waits for the exchange before anything reads the halos."""
        s, torch = self.slab, self.torch
        only = colour_filter(s.lo, s.hi, istart, colour)
        ops = halo_ops(s.lines(name), s.rank, s.world, s.hw, s.he, s.nown, 1, 1, only)
        if not self.overlap or not ops:
            s.ctx.sweep(istart, jstart, name, colour, first=first)
            self.comm.exchange(ops)
            return
        main = torch.cuda.current_stream()
        if self._comm_stream is None:
            self._comm_stream = torch.cuda.Stream(priority=-1)   # its two small launches must not queue behind the bulk
            self._ev = [torch.cuda.Event(), torch.cuda.Event()]
        side = self._comm_stream
        # edge lines + their exchange on the side stream, concurrently with the other lines on the compute stream
        # (the two launches touch disjoint lines of this colour and only read lines of the other colour)
        self._ev[0].record(main)
        side.wait_event(self._ev[0])
        s.ctx.set_stream(side.cuda_stream)
        s.ctx.sweep_part(istart, jstart, name, colour, first=first, part=1)
        s.ctx.set_stream(main.cuda_stream)
        with torch.cuda.stream(side):
            self.comm.exchange(ops)
            self._ev[1].record(side)
        s.ctx.sweep_part(istart, jstart, name, colour, first=first, part=2)
        main.wait_event(self._ev[1])

    def residuals(self):
        r = self.slab.res.clone()
        self.comm.allreduce_sum(r)
        return tuple(r.cpu().tolist())

    def iterate(self, n=1):
        self.iterate_async(n)
        return self.residuals()[:3]

    def gather_wall(self, name):
        """taun / yplusn along the top wall on rank 0 (None elsewhere): owned entries of every slab in line order."""
        out = [None] * self.world if self.rank == 0 else None
        self.dist.gather_object(self.slab.wall_owned(name), out, dst=0)
        return np.concatenate(out) if self.rank == 0 else None

    def gather_field(self, name):
        """Global (ni,nj) array on rank 0 (None elsewhere): owned lines of every slab, in order."""
        a = np.ascontiguousarray(self.slab.owned(self.ctx.download(name)))
        out = [None] * self.world if self.rank == 0 else None
        self.dist.gather_object(a, out, dst=0)
        return np.asfortranarray(np.concatenate(out, axis=0)) if self.rank == 0 else None

class LocalSlabCavity:
    """All slabs in one process on one GPU (tests): same stages, LocalComm copies instead of messages."""

    def __init__(self, nicv, njcv, world, slab_cls=None, **kw):
        self.world = world
        self.slabs = [(slab_cls or Slab)(nicv, njcv, r, world, **kw) for r in range(world)]
        self.comm = LocalComm(world)
        self.stages = iteration_stages(self.slabs[0].nswp, getattr(self.slabs[0], "solve_t", False),
                                       getattr(self.slabs[0], "turb", None))
        self.comm.exchange_all(self.slabs, "den", HALO_W, HALO_E)
        self.comm.exchange_all(self.slabs, "vis", 1, 1)
        if getattr(self.slabs[0], "turb", None) is not None:
            self.comm.exchange_all(self.slabs, "gamh", 1, 1)

    def iterate(self, n=1):
        import torch
        for _ in range(n):
            for st in self.stages:
                if st[0] == "compute":
                    for s in self.slabs:
                        st[1](s)
                elif st[0] == "sweep":       # same split as the SPMD driver: edge lines, exchange, the other lines
                    name, istart, jstart, colour, first = st[1:]
                    for s in self.slabs:
                        s.ctx.sweep_part(istart, jstart, name, colour, first=first, part=1)
                    self.comm.exchange_all(self.slabs, name, 1, 1, (istart, colour))
                    for s in self.slabs:
                        s.ctx.sweep_part(istart, jstart, name, colour, first=first, part=2)
                elif st[0] == "halo":
                    self.comm.exchange_all(self.slabs, st[1], st[2], st[3], st[4])
                elif st[0] == "halos":
                    for name, w, e in st[1]:
                        self.comm.exchange_all(self.slabs, name, w, e)
                else:
                    owner = [s for s in self.slabs if s.owner_of_pref][0]
                    for s in self.slabs:
                        s.ppref.copy_(owner.ppref_element())
        torch.cuda.synchronize()
        res = sum(s.res.clone() for s in self.slabs)
        return tuple(res.cpu().tolist())[:3]

    def gather_field(self, name):
        return np.asfortranarray(np.concatenate([s.owned(s.ctx.download(name)) for s in self.slabs], axis=0))

    def residuals(self):
        """All residual slots (resoru, resorv, resorm, resort, resork, resore, ...), summed over the slabs."""
        res = sum(s.res.clone() for s in self.slabs)
        return tuple(res.cpu().tolist())

    def gather_wall(self, name):
        """taun / yplusn along the top wall: owned entries of every slab in line order."""
        return np.concatenate([s.wall_owned(name) for s in self.slabs])

    def close(self):
        for s in self.slabs:
            s.ctx.close()


# ----------------------------------------------------------------------------------------------------------
# 2-D FTCS heat conduction: j-slabs (rows are contiguous in the column-major layout), one halo row per side
# and step.  The update loop is fdm1d_explicit_heat_cond.f90:130-139 in 2-D.
# ----------------------------------------------------------------------------------------------------------
def heat_slab_extent(ny_global: int, world: int, rank: int, halo: int = 1):
    """(j_goff, ny_local, halo_s, halo_n): owned rows plus `halo` rows towards each neighbour (1 for the explicit scheme;
    the implicit one needs cfd_heat_implicit_halo rows)."""
    lo, hi = partition(ny_global, world)[rank]
    if world > 1 and hi - lo + 1 < halo:
        raise ValueError(f"slab of rank {rank} owns {hi - lo + 1} rows; the halo is {halo} rows deep")
    hs = halo if rank > 0 else 0
    hn = halo if rank < world - 1 else 0
    return lo - 1 - hs, (hi - lo + 1) + hs + hn, hs, hn


def implicit_halo_rows(nx, ny_global, **kw):
    """Halo depth of a row slab for implicit (LOD) steps of this problem: cfd_heat_implicit_halo at its Fo_y."""
    import ctypes as C
    from . import _lib
    dx = (kw.get("xb", 1.0) - kw.get("xa", 0.0)) / (nx - 1)
    dy = (kw.get("yb", 1.0) - kw.get("ya", 0.0)) / (ny_global - 1)
    k, rho, cp = kw.get("k", 5.0), kw.get("rho", 1.0), kw.get("cp", 1.0)
    dt = kw.get("dt")
    if dt is None:
        dt = kw.get("fo", 0.2) * dx * dx / (k / rho / cp)
    rows = C.c_int()
    rc = _lib.load().cfd_heat_implicit_halo(k * dt / rho / cp / (dy * dy), C.byref(rows))
    if rc:
        raise RuntimeError("cfd_heat_implicit_halo failed")
    return rows.value


class HeatSlab:
    def __init__(self, nx, ny_global, rank, world, device=-1, t0=None, halo=1, **kw):
        from .heat import Heat2D
        self.rank, self.world, self.nx, self.ny_global = rank, world, nx, ny_global
        self.j_goff, self.ny, self.hs, self.hn = heat_slab_extent(ny_global, world, rank, halo)
        self.nown = self.ny - self.hs - self.hn
        dx = (kw.get("xb", 1.0) - kw.get("xa", 0.0)) / (nx - 1)
        k, rho, cp = kw.get("k", 5.0), kw.get("rho", 1.0), kw.get("cp", 1.0)
        if kw.get("dt") is None:
            kw["dt"] = kw.pop("fo", 0.2) * dx * dx / (k / rho / cp)
        self.h = Heat2D(nx, self.ny, device=device, j_goff=self.j_goff, ny_global=ny_global, halo_s=self.hs,
                        halo_n=self.hn, **kw)
        if t0 is not None:
            self.h.upload(np.asfortranarray(t0[:, self.j_goff:self.j_goff + self.ny]))

    def rows(self):
        """Current field as a torch tensor [ny_local, nx] (rows contiguous) aliasing device memory."""
        base, stride, n = self.h.device_view()
        return device_tensor(base, (n, stride))

    def halo_ops(self):
        """Every halo row from the neighbour's owned rows (one row for the explicit scheme, the whole warm-up depth for
        the implicit one)."""
        r, ops = self.rows(), []
        first, last = self.hs, self.hs + self.nown - 1
        if self.rank > 0:
            ops += [("recv", r[first - self.hs:first], self.rank - 1), ("send", r[first:first + self.hs], self.rank - 1)]
        if self.rank < self.world - 1:
            ops += [("recv", r[last + 1:last + 1 + self.hn], self.rank + 1),
                    ("send", r[last + 1 - self.hn:last + 1], self.rank + 1)]
        return ops

    def owned(self, a):
        return a[:, self.hs:self.hs + self.nown]


class SlabHeat:
    """SPMD driver over torch.distributed: step = FTCS update of the owned rows, then halo-row exchange."""

    def __init__(self, nx, ny_global, *, stream=None, **kw):
        import torch
        self.torch = torch
        self.comm = TorchDistComm()
        self.slab = HeatSlab(nx, ny_global, self.comm.rank, self.comm.world, device=torch.cuda.current_device(), **kw)
        if stream is not None:
            self.slab.h.set_stream(stream.cuda_stream)

    def step(self, n=1):
        for _ in range(n):
            self.slab.h.step(1)
            self.comm.exchange(self.slab.halo_ops())

    def implicit_step(self, n=1):
        """Implicit (LOD) steps; the slabs must have been built with halo=implicit_halo_rows(...).  Per step: the x- and
        y-sweeps on the local rows, then the refresh of the halo rows (whose results the sweeps leave inexact)."""
        for _ in range(n):
            self.slab.h.implicit_step(1)
            self.comm.exchange(self.slab.halo_ops())


class LocalSlabHeat:
    """All slabs in one process on one GPU (tests)."""

    def __init__(self, nx, ny_global, world, t0, **kw):
        self.slabs = [HeatSlab(nx, ny_global, r, world, t0=t0, **dict(kw)) for r in range(world)]

    def implicit_step(self, n=1):
        for _ in range(n):
            for s in self.slabs:
                s.h.implicit_step(1)
            self._exchange()

    def _exchange(self):
        rows = [s.rows() for s in self.slabs]
        for r in range(len(self.slabs) - 1):
            a, b = self.slabs[r], self.slabs[r + 1]
            a_last = a.hs + a.nown - 1
            rows[r + 1][b.hs - b.hs:b.hs].copy_(rows[r][a_last + 1 - b.hs:a_last + 1])
            rows[r][a_last + 1:a_last + 1 + a.hn].copy_(rows[r + 1][b.hs:b.hs + a.hn])

    def step(self, n=1):
        for _ in range(n):
            for s in self.slabs:
                s.h.step(1)
            rows = [s.rows() for s in self.slabs]
            for r in range(len(self.slabs) - 1):
                a, b = self.slabs[r], self.slabs[r + 1]
                rows[r + 1][b.hs - 1:b.hs].copy_(rows[r][a.hs + a.nown - 1:a.hs + a.nown])
                rows[r][a.hs + a.nown:a.hs + a.nown + 1].copy_(rows[r + 1][b.hs:b.hs + 1])

    def gather(self):
        return np.asfortranarray(np.concatenate([s.owned(s.h.download()) for s in self.slabs], axis=1))

    def close(self):
        for s in self.slabs:
            s.h.close()


# ----------------------------------------------------------------------------------------------------------
# Cavity family cut along j (second variant; north_star: "lines along the slab direction are solved
# with a partitioned Thomas (Wang/SPIKE) method whose small reduced system is all-gathered").
#
# Every rank holds ALL lines but only rows joff+1 .. joff+nj of each, so each tridiagonal system of lisolv
# (cavity_fvm.f90:62-80) is cut into `world` pieces.  Per colour pass of a sweep:
#     cfd_sweep            piece solution y with zero values at the cut ends            (local, 8 doubles/cell)
#     cfd_spike_pack       first/last y of every piece -> [nlines, 2]                   (tiny)
#     all_gather           -> [world, nlines, 2]                                        (NCCL)
#     cfd_spike_finish     reduced solve per line + x = y + w*x_s + z*x_n, halo rows    (local, 4 doubles/cell)
# and once per assembly cfd_spike_basis (w, z: two matrix-only solves) + all_gather of [ni, 4].  Red/black line
# colouring needs NO exchange here (all lines are local); halo ROWS of u, v, p move once per outer iteration,
# of dv once before the p' assembly.
# ----------------------------------------------------------------------------------------------------------
JHALO_S, JHALO_N = 2, 1      # calcv reads den(i, j-2) (cavity_fvm.f90:215); everything else needs one row


def jslab_extent(nj_global: int, world: int, rank: int):
    """(joff, nj_local, halo_s, halo_n) of rank's local arrays: owned rows plus halo rows, clipped at the walls."""
    lo, hi = partition(nj_global, world)[rank]
    hs = min(JHALO_S, lo - 1) if rank > 0 else 0
    hn = min(JHALO_N, nj_global - hi) if rank < world - 1 else 0
    if world > 1 and hi - lo + 1 < 3:
        raise ValueError(f"j-slab of rank {rank} owns {hi - lo + 1} rows; every slab needs at least 3")
    return lo - 1 - hs, (hi - lo + 1) + hs + hn, hs, hn


class JSlab:
    """One rank's rows of the global cavity problem plus the device buffers of the partition method."""

    def __init__(self, nicv, njcv, rank, world, *, reynolds=100.0, utop=1.0, rho=1000.0, iadv=3,
                 urf=(0.5, 0.5, 0.8), nswp=(3, 3, 3), device=-1, xlim=(0.0, 1.0), ylim=(0.0, 1.0)):
        import torch
        self.torch = torch
        self.rank, self.world = rank, world
        ni, nj_g = nicv + 2, njcv + 2
        self.ni, self.nj_global = ni, nj_g
        self.joff, self.nj, self.hs, self.hn = jslab_extent(nj_g, world, rank)
        self.lo, self.hi = partition(nj_g, world)[rank]
        self.nown = self.hi - self.lo + 1
        g = setgrid(nicv, njcv, xlim[0], xlim[1], ylim[0], ylim[1])
        sl = slice(self.joff, self.joff + self.nj)
        self.ctx = CavityFVM(nicv, njcv, iadv=iadv, order="redblack", urf=urf, nswp=nswp, device=device,
                             nj_local=self.nj, joff=self.joff, halo_s=self.hs, halo_n=self.hn,
                             sns_njm1=float(g["sns"][nj_g - 2]))
        self.ctx.set_geometry(x=g["x"], xu=g["xu"], sew=g["sew"], sewu=g["sewu"],
                              y=g["y"][sl], yv=g["yv"][sl], sns=g["sns"][sl], snsv=g["snsv"][sl])
        den = np.full((ni, self.nj), rho, order="F")
        vis = np.asfortranarray(den * utop * (xlim[1] - xlim[0]) / reynolds)
        u = np.zeros((ni, self.nj), order="F")
        jl = nj_g - 1 - self.joff                                      # local index of the lid row
        if 0 <= jl < self.nj:
            u[1:, jl] = utop
        self.ctx.upload("den", den)
        self.ctx.upload("vis", vis)
        self.ctx.upload("u", u)
        self.nswp = nswp
        self.ipref, self.jpref = self.ctx.ipref, self.ctx.jpref
        self.owner_of_pref = self.lo <= self.jpref <= self.hi
        self.ppref = torch.zeros(1, dtype=torch.float64, device="cuda")
        self.res = device_tensor(self.ctx.residual_device_ptr(), (8,))
        f64 = dict(dtype=torch.float64, device="cuda")
        self.wz_local = torch.zeros((ni, 4), **f64)
        self.wz_all = torch.zeros((world, ni, 4), **f64)
        self._y = {}
        self.row_send = {k: torch.zeros((2, ni), **f64) for k in ("s", "n")}     # to the south / north neighbour
        self.row_recv = {k: torch.zeros((2, ni), **f64) for k in ("s", "n")}
        self._pp_lines = None

    def y_buffers(self, nl):
        if nl not in self._y:
            t = self.torch
            self._y[nl] = (t.zeros((nl, 2), dtype=t.float64, device="cuda"),
                           t.zeros((self.world, nl, 2), dtype=t.float64, device="cuda"))
        return self._y[nl]

    def ppref_element(self):
        if self._pp_lines is None:
            base, stride, n = self.ctx.device_view("pp")
            self._pp_lines = device_tensor(base, (n, stride))
        L = self.ctx.layout
        j0 = self.jpref - 1 - self.joff
        q = (j0 % L["M"]) * L["T"] + j0 // L["M"]
        return self._pp_lines[self.ipref - 1, q:q + 1]

    def owned(self, a):
        return a[:, self.hs:self.hs + self.nown]


def _nlines(ni, istart, colour):
    return len(range(istart + colour, ni, 2))          # global lines istart..ni-1 with (i-istart)%2 == colour


class _JSlabEngine:
    """Stage sequence of one outer iteration over a list of slabs (one per process under torch.distributed, all of
    them in one process for single-GPU tests); `gather` and `rows` do the two kinds of communication."""

    def _equation(self, name, istart, jstart, nsw, first):
        slabs = self.slabs
        for s in slabs:
            s.ctx.spike_basis(istart, jstart)
            s.ctx.spike_pack(istart, jstart, name, 0, 1, s.wz_local.data_ptr())
        self.gather([(s.wz_local, s.wz_all) for s in slabs])
        for k in range(nsw):
            for colour in (0, 1):
                nl = _nlines(slabs[0].ni, istart, colour)
                if nl == 0:
                    continue
                for s in slabs:
                    yl, _ = s.y_buffers(nl)
                    s.ctx.sweep(istart, jstart, name, colour, first=(first and k == 0))
                    s.ctx.spike_pack(istart, jstart, name, colour, 0, yl.data_ptr())
                self.gather([s.y_buffers(nl) for s in slabs])
                for s in slabs:
                    _, ya = s.y_buffers(nl)
                    s.ctx.spike_finish(istart, jstart, name, colour, s.wz_all.data_ptr(), ya.data_ptr(), s.rank, s.world)

    def _iterate_once(self):
        slabs = self.slabs
        nswp = slabs[0].nswp
        self.rows((("u", 1, 1), ("v", 1, 1), ("p", 1, 0)))
        for s in slabs:
            s.ctx.assemble_u()
        self._equation("u", 3, 2, nswp[0], False)
        for s in slabs:
            s.ctx.assemble_v()
        self._equation("v", 2, 3, nswp[1], False)
        self.rows((("dv", 0, 1),))
        for s in slabs:
            s.ctx.assemble_p()
        self._equation("pp", 2, 2, nswp[2], True)
        self.bcast_ppref()
        for s in slabs:
            s.ctx.correct(s.ppref.data_ptr())


class LocalJSlabCavity(_JSlabEngine):
    """All j-slabs in one process on one GPU (tests): copies instead of messages."""

    def __init__(self, nicv, njcv, world, **kw):
        self.world = world
        self.slabs = [JSlab(nicv, njcv, r, world, **kw) for r in range(world)]

    def gather(self, pairs):
        for _, dst in pairs:
            for r, (src, _) in enumerate(pairs):
                dst[r].copy_(src)

    def rows(self, spec):
        for name, ns, nn in spec:
            for r in range(self.world - 1):
                a, b = self.slabs[r], self.slabs[r + 1]          # a is south of b
                n_s = min(ns, b.hs)                               # b's south halo <- a's last owned rows
                if n_s:
                    buf = b.row_recv["s"][:n_s]
                    a.ctx.rows_pack(name, a.hs + a.nown - n_s + 1, n_s, buf.data_ptr())
                    b.ctx.rows_unpack(name, b.hs - n_s + 1, n_s, buf.data_ptr())
                n_n = min(nn, a.hn)                               # a's north halo <- b's first owned rows
                if n_n:
                    buf = a.row_recv["n"][:n_n]
                    b.ctx.rows_pack(name, b.hs + 1, n_n, buf.data_ptr())
                    a.ctx.rows_unpack(name, a.hs + a.nown + 1, n_n, buf.data_ptr())

    def bcast_ppref(self):
        owner = [s for s in self.slabs if s.owner_of_pref][0]
        for s in self.slabs:
            s.ppref.copy_(owner.ppref_element())

    def iterate(self, n=1):
        import torch
        for _ in range(n):
            self._iterate_once()
        torch.cuda.synchronize()
        res = sum(s.res.clone() for s in self.slabs)
        return tuple(res.cpu().tolist())[:3]

    def gather_field(self, name):
        return np.asfortranarray(np.concatenate([s.owned(s.ctx.download(name)) for s in self.slabs], axis=1))

    def close(self):
        for s in self.slabs:
            s.ctx.close()


class JSlabCavity(_JSlabEngine, _E2EMixin):
    """SPMD driver: one JSlab per process over torch.distributed (NCCL all_gather + send/recv of halo rows)."""

    def __init__(self, nicv, njcv, *, reynolds=1000.0, stream=None, **kw):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.comm = TorchDistComm()
        self.rank, self.world = self.comm.rank, self.comm.world
        self.slab = JSlab(nicv, njcv, self.rank, self.world, reynolds=reynolds, device=torch.cuda.current_device(), **kw)
        self.slabs = [self.slab]
        self.ctx = self.slab.ctx
        if stream is not None:
            self.ctx.set_stream(stream.cuda_stream)
        self.pref_owner = [r for r, (lo, hi) in enumerate(partition(njcv + 2, self.world)) if lo <= self.slab.jpref <= hi][0]

    def gather(self, pairs):
        for src, dst in pairs:
            self.comm.allgather(dst, src)

    def rows(self, spec):
        s, ops, todo = self.slab, [], []
        for k, (name, ns, nn) in enumerate(spec):
            if s.rank > 0:                                        # south neighbour: my first owned rows go to its north halo
                if nn:
                    buf = self._rowbuf("ss", k)[:1]
                    s.ctx.rows_pack(name, s.hs + 1, 1, buf.data_ptr())
                    ops.append(("send", buf, s.rank - 1))
                n_s = min(ns, s.hs)
                if n_s:
                    buf = self._rowbuf("rs", k)[:n_s]
                    ops.append(("recv", buf, s.rank - 1))
                    todo.append((name, s.hs - n_s + 1, n_s, buf))
            if s.rank < s.world - 1:                              # north neighbour
                if ns:
                    n_s = min(ns, JHALO_S)
                    buf = self._rowbuf("sn", k)[:n_s]
                    s.ctx.rows_pack(name, s.hs + s.nown - n_s + 1, n_s, buf.data_ptr())
                    ops.append(("send", buf, s.rank + 1))
                n_n = min(nn, s.hn)
                if n_n:
                    buf = self._rowbuf("rn", k)[:n_n]
                    ops.append(("recv", buf, s.rank + 1))
                    todo.append((name, s.hs + s.nown + 1, n_n, buf))
        self.comm.exchange(ops)
        for name, j1, n, buf in todo:
            s.ctx.rows_unpack(name, j1, n, buf.data_ptr())

    def _rowbuf(self, kind, k):
        key = (kind, k)
        if not hasattr(self, "_rb"):
            self._rb = {}
        if key not in self._rb:
            t = self.torch
            self._rb[key] = t.zeros((2, self.slab.ni), dtype=t.float64, device=getattr(self, "_device", "cuda"))
        return self._rb[key]

    def _global_cells(self):
        return (self.slab.ni - 2) * (self.slab.nj_global - 2)

    def bcast_ppref(self):
        s = self.slab
        if s.owner_of_pref:
            s.ppref.copy_(s.ppref_element())
        self.comm.broadcast(s.ppref, self.pref_owner)

    def iterate_async(self, n=1):
        for _ in range(n):
            self._iterate_once()

    def residuals(self):
        r = self.slab.res.clone()
        self.comm.allreduce_sum(r)
        return tuple(r.cpu().tolist())

    def iterate(self, n=1):
        self.iterate_async(n)
        return self.residuals()[:3]

    def gather_field(self, name):
        a = np.ascontiguousarray(self.slab.owned(self.ctx.download(name)))
        out = [None] * self.world if self.rank == 0 else None
        self.dist.gather_object(a, out, dst=0)
        return np.asfortranarray(np.concatenate(out, axis=1)) if self.rank == 0 else None
