"""Host-side mirror of the reference's "general" SIMPLE programs over the C ABI: modules vars / user / serv and
`program main` of cylinder-heat-conduction-2d/cylin_cond.f90:42-312 and buoyancy-driven-cavity-2d/buoyancy_cav.f90.

`grid()` (buoyancy_cav.f90:28-120) is O(ni+nj) host logic and is restated here; the `user` module's hooks
(control / initial / gamsor / boundary, cylin_cond.f90:83-169) become: static fields uploaded once (gam per
equation, per-volume sources) plus enumerated device hooks (`cfd_set_hook`).  All field arithmetic -- calcu,
calcv, calcp, calct, tdma -- runs in libcfd_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import CfdConfig, CfdHookDesc, check
from .cavity import CavityFVM

GEOM_I = ("x", "xu", "xdif", "sew", "sewu", "xcvi", "xcvip", "fx", "fxm")
GEOM_J = ("y", "yv", "ydif", "sns", "snsv", "ycvr", "ycvrs", "arx", "arxj", "arxjp", "r", "rmn", "sx", "sxmn",
          "fv", "fvp", "fy", "fym")


def grid(ni, nj, xstart, xend, ystart, yend, rstart, icrd):
    """grid() of buoyancy_cav.f90:28-120, statement by statement (1-based arrays held in index-shifted numpy)."""
    g = {n: np.zeros(ni + 1) for n in GEOM_I}
    g.update({n: np.zeros(nj + 1) for n in GEOM_J})
    x, xu, xdif, sew, sewu, xcvi, xcvip, fx, fxm = (g[n] for n in GEOM_I)
    (y, yv, ydif, sns, snsv, ycvr, ycvrs, arx, arxj, arxjp, r, rmn, sx, sxmn, fv, fvp, fy, fym) = (g[n] for n in GEOM_J)
    nim1, nim2, njm1, njm2 = ni - 1, ni - 2, nj - 1, nj - 2
    dx = (xend - xstart) / (ni - 2)
    xu[2] = xstart
    for i in range(3, ni + 1):
        xu[i] = xu[i - 1] + dx
    for i in range(2, nim1 + 1):
        sew[i] = xu[i + 1] - xu[i]
    x[1] = xu[2]
    for i in range(2, nim1 + 1):
        x[i] = 0.5 * (xu[i] + xu[i + 1])
    x[ni] = xu[ni]
    for i in range(2, ni + 1):
        xdif[i] = x[i] - x[i - 1]
    for i in range(3, nim1 + 1):
        sewu[i] = xdif[i]
    sewu[3] = sewu[3] + xdif[2]
    sewu[nim1] = sewu[nim1] + xdif[ni]
    for i in range(3, nim2 + 1):
        xcvi[i] = 0.5 * sew[i]
        xcvip[i] = xcvi[i]
    xcvip[2] = sew[2]
    xcvi[nim1] = sew[nim1]
    dy = (yend - ystart) / (nj - 2)
    yv[2] = ystart
    for j in range(3, nj + 1):
        yv[j] = yv[j - 1] + dy
    for j in range(2, njm1 + 1):
        sns[j] = yv[j + 1] - yv[j]
    y[1] = yv[2]
    for j in range(2, njm1 + 1):
        y[j] = 0.5 * (yv[j] + yv[j + 1])
    y[nj] = yv[nj]
    for j in range(2, nj + 1):
        ydif[j] = y[j] - y[j - 1]
    for j in range(3, njm1 + 1):
        snsv[j] = ydif[j]
    snsv[3] = snsv[3] + ydif[2]
    snsv[njm1] = snsv[njm1] + ydif[nj]
    r[1] = rstart
    if icrd == 1:
        rmn[1:] = 1.0
        r[1:] = 1.0
    else:
        for j in range(2, nj + 1):
            r[j] = r[j - 1] + ydif[j]
        rmn[2] = r[1]
        for j in range(3, nj + 1):
            rmn[j] = rmn[j - 1] + sns[j - 1]
    sx[1:] = 1.0
    sxmn[1:] = 1.0
    if icrd == 3:
        sx[1:] = r[1:]
        sxmn[2:] = rmn[2:]
    for j in range(2, njm1 + 1):
        ycvr[j] = r[j] * sns[j]
        arx[j] = ycvr[j]
    if icrd == 3:
        arx[1:] = sns[1:]
    for j in range(4, njm2 + 1):
        ycvrs[j] = 0.5 * (r[j] + r[j - 1]) * ydif[j]
    ycvrs[3] = 0.5 * (r[3] + r[1]) * snsv[3]
    ycvrs[njm1] = 0.5 * (r[nj] + r[njm2]) * snsv[njm1]
    for j in range(3, njm2 + 1):
        if icrd == 2:
            arxj[j] = 0.25 * (1.0 + rmn[j] / r[j]) * arx[j]
            arxjp[j] = arx[j] - arxj[j]
        else:
            arxj[j] = 0.5 * arx[j]
            arxjp[j] = arxj[j]
    arxjp[2] = arx[2]
    arxj[njm1] = arx[njm1]
    for j in range(3, njm2 + 1):
        fv[j] = arxjp[j] / arx[j]
        fvp[j] = 1.0 - fv[j]
    for i in range(2, ni + 1):
        fx[i] = 0.5 * sew[i - 1] / xdif[i]
        fxm[i] = 1.0 - fx[i]
    for j in range(2, nj + 1):
        fy[j] = 0.5 * sns[j - 1] / ydif[j]
        fym[j] = 1.0 - fy[j]
    return {k: v[1:].copy() for k, v in g.items()}


class GeneralFVM(CavityFVM):
    """One GPU context of CFD_FAMILY_GENERAL.  Attribute names follow module vars (cylin_cond.f90:42-71)."""

    def __init__(self, ni, nj, *, icrd=1, iadv=3, order="redblack", urf=(0.3, 0.3, 0.9, 1.0), nsweep=(3, 3, 3, 3),
                 lsolve_flow=True, lsolve_t=True, ipref=1, jpref=1, xstart=0.0, xend=1.0, ystart=0.0, yend=1.0,
                 rstart=1.0, device=-1, tdma_m=0, ni_local=None, goff=0, halo_w=0, halo_e=0):
        self.lib = _lib.load()
        nl = ni if ni_local is None else ni_local           # i-slab: lines goff+1 .. goff+nl of the global ni x nj grid
        self.ni, self.nj, self.icrd, self.iadv = nl, nj, icrd, iadv
        self.ni_global, self.nicv, self.njcv = ni, ni - 2, nj - 2
        self.ipref, self.jpref = ipref, jpref
        self.imon, self.jmon = ni // 2, nj // 2
        self.goff = goff
        self.lsolve_flow, self.lsolve_t = lsolve_flow, lsolve_t
        cfg = CfdConfig()
        cfg.family = _lib.FAMILY_GENERAL
        cfg.ni, cfg.nj, cfg.iadv = nl, nj, iadv
        cfg.goff, cfg.halo_w, cfg.halo_e = goff, halo_w, halo_e
        cfg.ni_global = ni if ni_local is not None else 0
        cfg.order = _lib.ORDER_SEQUENTIAL if order in ("sequential", "seq", 1) else _lib.ORDER_REDBLACK
        for k in range(4):
            cfg.nswp[k] = nsweep[k]
            cfg.urf[k] = urf[k]
        cfg.ipref, cfg.jpref, cfg.device, cfg.tdma_m = ipref, jpref, device, tdma_m
        cfg.solve_flow, cfg.solve_t = int(lsolve_flow), int(lsolve_t)
        self._cfg = cfg
        self._h = C.c_void_p()
        check(self.lib.cfd_create(C.byref(self._h), C.byref(cfg)))
        self.resor = [0.0] * 4
        g = grid(ni, nj, xstart, xend, ystart, yend, rstart, icrd)       # global 1-D arrays, O(ni+nj) host logic
        self.geometry_global = g
        if ni_local is not None:
            g = {k: (v[goff:goff + nl].copy() if k in GEOM_I else v) for k, v in g.items()}
        for k, v in g.items():
            setattr(self, k, v)
        self.set_geometry(**g)

    def calct(self) -> float:
        r = C.c_double()
        check(self.lib.cfd_calct(self._h, C.byref(r)))
        self.resor[3] = r.value
        return r.value

    def assemble_t(self):
        check(self.lib.cfd_assemble_t(self._h))

    def set_hook(self, eq: int, kind: int, *params):
        d = CfdHookDesc(kind=kind)
        for k, v in enumerate(params):
            d.params[k] = v
        check(self.lib.cfd_set_hook(self._h, eq, C.byref(d)))

    def iterate(self, n: int = 1):
        res = (C.c_double * 8)()
        check(self.lib.cfd_iterate(self._h, n, res))
        self.resor = list(res[:4])
        return tuple(res[:4])

    def run(self, max_iter=19999, tol=1.0e-10, check_every=1):
        """program main (cylin_cond.f90:295-307) with the stop rule of output() (:183); returns iterations done.
        The `boundary()` hook of both shipped user modules does nothing until the last pass, so it is applied by
        the caller after `run` (see CylinderConduction.finish / BuoyantCavity.finish)."""
        it = 0
        while it < max_iter:
            n = min(check_every, max_iter - it)
            res = self.iterate(n)
            it += n
            if max(res) < tol or it >= max_iter - 1:
                if it < max_iter:          # the llast pass
                    self.iterate(1)
                    it += 1
                break
        return it


class CylinderConduction(GeneralFVM):
    """The conduction case exactly as shipped: module user of cylin_cond.f90:73-169."""

    def __init__(self, ni=52, nj=52, **kw):
        super().__init__(ni, nj, icrd=2, iadv=kw.pop("iadv", 1), lsolve_flow=False, lsolve_t=True,
                         xstart=0.0, xend=2.0, ystart=0.0, yend=1.0, rstart=1.0, **kw)
        self.ibnd = 2
        self.is1, self.is2, self.js2 = ni // 2 - 10, ni // 2 + 10, nj
        self.js1 = self.js2 - 20
        self.gam0, self.gam1, self.qin, self.htc, self.t0, self.tf = 3.0, 1.0, 50.0, 50.0, 20.0, 10.0
        # initial(): cylin_cond.f90:117-134 (+ den_ini = cp_ini = 1, grid() :118-119)
        t = np.full((ni, nj), self.t0, order="F")
        t[0, :] = self.t0 * (1.0 + self.y)
        one = np.ones((ni, nj), order="F")
        self.upload("t", t)
        self.upload("den", one)
        self.upload("cp", one)
        # gamsor(): cylin_cond.f90:149-169 -- static, so evaluated once on the host and uploaded
        gam = np.full((ni, nj), self.gam0, order="F")
        su = np.zeros((ni, nj), order="F")
        sp = np.zeros((ni, nj), order="F")
        a, b, c, d = self.is1 - 1, self.is2, self.js1 - 1, self.js2
        gam[a:b, c:d] = self.gam1
        su[a:b, c:d] = 200.0
        sp[a:b, c:d] = -0.1
        nim1, njm1 = ni - 1, nj - 1
        gam[1:nim1, nj - 1] = 0.0
        for i in range(1, nim1):
            su[i, 1] = su[i, 1] + self.r[0] / self.arx[1] * self.tf / (self.ydif[1] / gam[i, 1] + 1.0 / self.htc)
            sp[i, 1] = sp[i, 1] - self.r[0] / self.arx[1] / (self.ydif[1] / gam[i, 1] + 1.0 / self.htc)
            gam[i, 0] = 0.0
        su[nim1 - 1, 1:njm1] = su[nim1 - 1, 1:njm1] + self.qin / self.sew[nim1 - 1]
        gam[ni - 1, 1:njm1] = 0.0
        self.gam = gam
        self.upload("gamt", gam)
        self.upload("susrc", su)
        self.upload("spsrc", sp)
        self.set_hook(_lib.EQ_T, _lib.HOOK_FIELD_SOURCES)

    def finish(self):
        """boundary() on the last pass (cylin_cond.f90:140-147) and the wall heat flux of output() (:187-190)."""
        t = self.download("t")
        ni, nj = self.ni, self.nj
        t[1:ni - 1, nj - 1] = t[1:ni - 1, nj - 2]
        t[1:ni - 1, 0] = (self.gam0 * t[1:ni - 1, 1] + self.htc * self.ydif[1] * self.tf) / (self.gam0 + self.htc * self.ydif[1])
        t[ni - 1, 1:nj - 1] = t[ni - 2, 1:nj - 1] + self.qin * self.xdif[ni - 1] / self.gam0
        return t


class BuoyantCavity(GeneralFVM):
    """Boussinesq cavity (BASELINE.json configs[3]).  The reference's own `user` module for buoyancy_cav.f90 is
    missing from its repository; this one is authored by the build (same choices as oracle/general_oracle.c):
    unit square, T = 1 west wall, T = 0 east wall, adiabatic north/south, rho = cp = 1, nu = Pr*alpha,
    rho*g*beta = Ra*nu*alpha, Tref = 0.5, hybrid scheme, urf .3/.3/.9/1."""

    def __init__(self, ni, nj, *, ra=1.0e6, pr=0.71, alpha=2.3e-5, **kw):
        super().__init__(ni, nj, icrd=1, iadv=kw.pop("iadv", 3), lsolve_flow=True, lsolve_t=True,
                         ipref=ni // 2, jpref=nj // 2, **kw)
        nu = pr * alpha
        self.visc, self.cond, self.rgbeta, self.tref = nu, alpha, ra * nu * alpha, 0.5
        nl, gi = self.ni, np.arange(self.goff + 1, self.goff + self.ni + 1)      # local lines and their global numbers
        one = np.ones((nl, nj), order="F")
        t = np.full((nl, nj), self.tref, order="F")
        t[gi == 1, :] = 1.0
        t[gi == ni, :] = 0.0
        self.upload("t", t)
        self.upload("den", one)
        self.upload("cp", one)
        self.upload("vis", np.full((nl, nj), self.visc, order="F"))
        gamt = np.full((nl, nj), self.cond, order="F")
        inner = (gi >= 2) & (gi <= ni - 1)
        gamt[inner, nj - 1] = 0.0
        gamt[inner, 0] = 0.0
        self.upload("gamt", gamt)
        self.set_hook(_lib.EQ_V, _lib.HOOK_BOUSSINESQ, self.rgbeta, self.tref)

    def finish(self):
        t = self.download("t")
        t[1:self.ni - 1, self.nj - 1] = t[1:self.ni - 1, self.nj - 2]
        t[1:self.ni - 1, 0] = t[1:self.ni - 1, 1]
        return t
