"""Host-side mirror of the TEACH-T sudden-expansion program (backward-step-flow-2d/: control.f90, general.f90)
over the C ABI, laminar subset (calcu, calcv, calcp with the modu / modv hooks; k-epsilon, energy and props are
outside this subset).  Grid generation (control.f90:72-90), `init` (general.f90:69-140) and the initial fields
(control.f90:148-179, 197-203) are O(ni+nj)/one-off host logic restated here; every per-iteration array
operation runs in libcfd_b200.so.  The reference computes in default REAL (FP32); this path is FP64.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import CfdConfig, CfdTurbulence, WALL_ID, check
from .cavity import CavityFVM


def teach_grid(ni, nj, rlarge=0.0124, epsx=1.05):
    """x(1:ni), y(1:nj) of control.f90:72-90 (geometric stretching epsx in x, uniform in y)."""
    nim1 = ni - 1
    x, y = np.zeros(ni + 1), np.zeros(nj + 1)
    altot = 16.0 * rlarge
    if epsx == 1.0:
        sumx = 0.5 + float(ni - 3) + 0.5
    else:
        sumx = 0.5 * epsx ** (ni - 4) + (epsx ** (ni - 3) - 1.0) / (epsx - 1.0) + 0.5
    dx = altot / sumx
    x[1] = -0.5 * dx
    x[2] = -x[1]
    for i in range(3, nim1 + 1):
        x[i] = x[i - 1] + dx
        dx = epsx * dx
    x[ni] = x[nim1] - x[ni - 2] + x[nim1]
    dy = rlarge / float(nj - 2)
    y[1] = -0.5 * dy
    for j in range(2, nj + 1):
        y[j] = y[j - 1] + dy
    return x[1:].copy(), y[1:].copy()


def teach_init(x, y, indcos):
    """init (general.f90:69-140): cell sizes and radii from the node coordinates."""
    ni, nj = len(x), len(y)
    nim1, njm1 = ni - 1, nj - 1
    X, Y = np.concatenate(([0.0], x)), np.concatenate(([0.0], y))      # 1-based
    zi, zj = (lambda: np.zeros(ni + 2)), (lambda: np.zeros(nj + 2))
    r, rv, rcv, yv, sns, snsv, dynp, dyps, dynpv, dypsv = (zj() for _ in range(10))
    xu, sew, sewu, dxep, dxpw, dxepu, dxpwu = (zi() for _ in range(7))
    for j in range(1, nj + 1):
        r[j] = Y[j] if indcos != 1 else 1.0
    for i in range(1, nim1 + 1):
        dxep[i] = X[i + 1] - X[i]
        dxpw[i + 1] = dxep[i]
    for j in range(1, njm1 + 1):
        dynp[j] = Y[j + 1] - Y[j]
        dyps[j + 1] = dynp[j]
    for i in range(2, nim1 + 1):
        sew[i] = 0.5 * (dxep[i] + dxpw[i])
    for j in range(2, njm1 + 1):
        sns[j] = 0.5 * (dynp[j] + dyps[j])
    for i in range(2, ni + 1):
        xu[i] = 0.5 * (X[i] + X[i - 1])
    for i in range(2, nim1 + 1):
        dxepu[i] = xu[i + 1] - xu[i]
        dxpwu[i + 1] = dxepu[i]
    for i in range(3, nim1 + 1):
        sewu[i] = 0.5 * (dxepu[i] + dxpwu[i])
    for j in range(2, nj + 1):
        rv[j] = 0.5 * (r[j] + r[j - 1])
        rcv[j] = 0.5 * (rv[j] + rv[j - 1])
        yv[j] = 0.5 * (Y[j] + Y[j - 1])
    for j in range(2, njm1 + 1):
        dynpv[j] = yv[j + 1] - yv[j]
        dypsv[j + 1] = dynpv[j]
    for j in range(3, njm1 + 1):
        snsv[j] = 0.5 * (dynpv[j] + dypsv[j])
    gi = dict(x=X, xu=xu, sew=sew, sewu=sewu, dxep=dxep, dxpw=dxpw, dxepu=dxepu, dxpwu=dxpwu)
    gj = dict(y=Y, yv=yv, r=r, rv=rv, rcv=rcv, sns=sns, snsv=snsv, dynp=dynp, dyps=dyps, dynpv=dynpv, dypsv=dypsv)
    out = {k: v[1:ni + 1].copy() for k, v in gi.items()}
    out.update({k: v[1:nj + 1].copy() for k, v in gj.items()})
    return out


class TeachFVM(CavityFVM):
    """One GPU context of CFD_FAMILY_TEACH with the shipped case's literals (control.f90:15-141), laminar.

    `ni_local / goff / halo_w / halo_e` make it an i-slab of the global ni x nj problem (lines goff+1 .. goff+ni_local,
    red/black order only): geometry and initial fields are cut from the global 1-D arrays / per-line formulas."""

    def __init__(self, ni=100, nj=100, *, order="redblack", epsx=1.05, indcos=2, rlarge=0.0124, uin=0.5,
                 densit=1000.0, viscos=0.8e-3, urf=(0.5, 0.5, 1.0), nswp=(3, 3, 5), jstep=None, device=-1, tdma_m=0,
                 ni_local=None, goff=0, halo_w=0, halo_e=0, turbulent=False, urf_turb=(0.7, 0.7, 1.0, 0.7),
                 nswp_turb=(3, 3, 3), tin=0.0, turbin=0.03, alamda=0.005, twalln=100.0, prandl=7.0):
        self.lib = _lib.load()
        nl = ni if ni_local is None else ni_local
        self.ni, self.nj = nl, nj
        self.ni_global, self.nicv, self.njcv, self.goff = ni, ni - 2, nj - 2, goff
        self.istep = 2
        self.jstep = jstep if jstep is not None else max(2, int(round(0.08 * nj)))
        self.ipref, self.jpref = 2, 2
        self.imon, self.jmon = min(10, ni - 1), min(10, nj - 1)
        x, y = teach_grid(ni, nj, rlarge, epsx)
        g = teach_init(x, y, indcos)                                   # global 1-D arrays (host logic, O(ni+nj))
        self.geometry_global = g
        sl = slice(goff, goff + nl)
        gl = {k: (v[sl].copy() if len(v) == ni and k in _BY_I else v) for k, v in g.items()}
        for k, v in gl.items():
            setattr(self, k, v)
        # initial fields and inlet mass flow (control.f90:148-179, 197-203), local lines only
        den = np.full((nl, nj), densit, order="F")
        vis = np.full((nl, nj), viscos, order="F")
        u = np.zeros((nl, nj), order="F")
        js, jstp1 = self.jstep, self.jstep + 1
        yv, rv = g["yv"], g["rv"]
        for il in range(nl):
            i = il + 1 + goff                                           # global 1-based line
            if i == 2:
                u[il, 1:js] = uin
            if i >= 2:
                jfin = nj if i >= self.istep + 1 else jstp1
                factor = (yv[jstp1 - 1] * rv[jstp1 - 1]) / (yv[jfin - 1] * rv[jfin - 1])
                u[il, 1:jfin - 1] = uin * factor
        flowin = 0.0                                                    # inlet plane: den uniform, u(2,j) as set above
        jfin2 = nj if 2 >= self.istep + 1 else jstp1
        u2 = np.zeros(nj)
        u2[1:jfin2 - 1] = uin * (yv[jstp1 - 1] * rv[jstp1 - 1]) / (yv[jfin2 - 1] * rv[jfin2 - 1])
        for j in range(2, nj):
            arden = 0.5 * (densit + densit) * g["r"][j - 1] * g["sns"][j - 1]
            flowin = flowin + arden * u2[j - 1]
        self.flowin = flowin
        cfg = CfdConfig()
        cfg.family = _lib.FAMILY_TEACH
        cfg.ni, cfg.nj, cfg.iadv = nl, nj, 3
        cfg.order = _lib.ORDER_SEQUENTIAL if order in ("sequential", "seq", 1) else _lib.ORDER_REDBLACK
        for k in range(3):
            cfg.nswp[k] = nswp[k]
            cfg.urf[k] = urf[k]
        cfg.ipref, cfg.jpref, cfg.device, cfg.tdma_m = 2, 2, device, tdma_m
        cfg.indcos, cfg.istep, cfg.jstep = indcos, self.istep, self.jstep
        cfg.viscos, cfg.flowin, cfg.great = viscos, flowin, 1.0e30
        cfg.goff = goff
        cfg.ni_global = ni if ni_local is not None else 0
        cfg.halo_w, cfg.halo_e = halo_w, halo_e
        self._cfg = cfg
        self._h = C.c_void_p()
        check(self.lib.cfd_create(C.byref(self._h), C.byref(cfg)))
        self.set_geometry(**gl)
        self.upload("den", den)
        self.upload("vis", vis)
        self.upload("u", u)
        self.resoru = self.resorv = self.resorm = 0.0
        self.turbulent = False
        if turbulent:
            self._init_turbulence(uin, rlarge, urf_turb, nswp_turb, tin, turbin, alamda, twalln, prandl)

    # -- k-epsilon / energy (control.f90:105-179): constants, initial k, eps, T, wall y+, first props ------------
    def _init_turbulence(self, uin, rlarge, urf_turb, nswp_turb, tin, turbin, alamda, twalln, prandl):
        ni, nj, js = self.ni, self.nj, self.jstep
        m = CfdTurbulence()
        m.cmu, m.cd, m.c1, m.c2, m.cappa, m.elog = 0.09, 1.00, 1.44, 1.92, .4187, 9.793
        m.pred = m.cappa * m.cappa / (m.c2 - m.c1) / (m.cmu ** .5)
        m.prte, m.prandt, m.prandl = 1.0, 0.9, prandl
        pfun = m.prandl / m.prandt
        m.pfun = 9.24 * (pfun ** 0.75 - 1.0) * (1.0 + 0.28 * np.exp(-0.007 * pfun))
        m.twalln = twalln
        m.urfk, m.urfe, m.urft, m.urfvis = urf_turb
        m.nswpk, m.nswpd, m.nswpt = nswp_turb
        m.incalk = m.incald = m.incalt = m.inpro = 1
        self.model = m
        check(self.lib.cfd_set_turbulence(self._h, C.byref(m)))
        self.turbulent = True
        tein = turbin * (uin * uin)
        edin = tein ** 1.5 / (alamda * rlarge)
        # local lines of an i-slab: il = 0 .. ni-1 is the global 1-based line il + 1 + goff
        goff, NI = self.goff, self.ni_global
        te, ed, t = (np.zeros((ni, nj), order="F") for _ in range(3))
        yplusn, xplusw = np.zeros(ni), np.zeros(nj)
        for il in range(ni):
            i = il + 1 + goff
            if i == 1:
                te[il, 1:js] = tein
                ed[il, 1:js] = edin
                t[il, 1:js] = tin
                t[il, js - 1:nj] = tin
            else:
                jfin = nj if i >= self.istep + 1 else js + 1
                te[il, 1:jfin - 1] = tein
                ed[il, 1:jfin - 1] = edin
            if 2 <= i <= NI - 1:
                t[il, nj - 1] = twalln
                yplusn[il] = 11.0
        xplusw[js - 1:nj] = 11.0
        xplusw[js - 1] = 0.0
        self.upload("te", te)
        self.upload("ed", ed)
        self.upload("t", t)
        self.wall_upload("yplusn", yplusn)
        self.wall_upload("xplusw", xplusw)
        self.props()      # i-slabs: the halo lines of vis, gamh follow by the driver's first exchange (dist.py)

    def wall_upload(self, name, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        check(self.lib.cfd_wall_upload(self._h, WALL_ID[name], a.ctypes.data_as(C.c_void_p), a.size))

    def wall_download(self, name):
        a = np.zeros(self.ni if name in ("taun", "yplusn") else self.nj)
        check(self.lib.cfd_wall_download(self._h, WALL_ID[name], a.ctypes.data_as(C.c_void_p), a.size))
        return a

    def props(self):
        check(self.lib.cfd_props(self._h))

    def assemble_te(self):
        check(self.lib.cfd_assemble_te(self._h))

    def assemble_ed(self):
        check(self.lib.cfd_assemble_ed(self._h))

    def assemble_t(self):
        """calct chapters 1-3 (calc.f90:309-385) + modt, without the sweeps (the slab driver interleaves the exchange)."""
        check(self.lib.cfd_assemble_t(self._h))

    def calcte(self):
        r = C.c_double()
        check(self.lib.cfd_calcte(self._h, C.byref(r)))
        return r.value

    def calced(self):
        r = C.c_double()
        check(self.lib.cfd_calced(self._h, C.byref(r)))
        return r.value

    def calct(self):
        r = C.c_double()
        check(self.lib.cfd_calct(self._h, C.byref(r)))
        return r.value

    def wall_coefficients(self, rlarge=0.0124, rsmall=None, uin=0.5, densit=1000.0, tin=0.0, cpp=4.190e3, tk=6.0e-1):
        """Shear-stress coefficient and Nusselt number along the large-duct wall (control.f90:243-272), from the
        device's taun / yplusn and the fields te, t, den; returns rows (i, xu/h, ssc, nusselt, x/h), i = istep..nim1."""
        m = self.model
        ni, nj = self.ni, self.nj
        rsmall = rsmall if rsmall is not None else 0.5 * (self.y[self.jstep - 1] + self.y[self.jstep])   # control.f90:91
        ularge = uin * (rsmall / rlarge) ** 2
        taun, yplusn = self.wall_download("taun"), self.wall_download("yplusn")
        j = nj - 1                                                          # njm1
        # only the wall-adjacent row of the three fields is needed: fetched by the device (no whole-field download)
        te, t, den = ({j - 1: self.extract_row(n, j)} for n in ("te", "t", "den"))
        te, t, den = (_Row(a) for a in (te, t, den))
        xlref, tref, delt = rlarge - rsmall, tin, abs(tin - m.twalln)
        dyn = self.yv[nj - 1] - self.y[nj - 2]
        rows = []
        for i in range(self.istep, ni):                                     # istep .. nim1
            ssc = taun[i - 1] / (densit * ularge * ularge)
            xuh = self.xu[i - 1] / (rlarge - rsmall)
            xh = self.x[i - 1] / xlref
            if yplusn[i - 1] <= 11.63:
                hfluxn = tk * abs(m.twalln - t[i - 1, j - 1]) / dyn
            else:
                uplusn = np.log(m.elog * yplusn[i - 1]) / m.cappa
                hfluxn = den[i - 1, j - 1] * cpp * np.sqrt(te[i - 1, j - 1]) * (tref - t[i - 1, j - 1]) / m.prandt / (uplusn + m.pfun)
            hfluxn = abs(hfluxn)
            rows.append((i, xuh, ssc, hfluxn * xlref / (tk * delt), xh))
        return rows

    def iterate_turb(self, n=1):
        """n passes of control.f90:210-217; returns (resoru, resorv, resorm, resort, resork, resore)."""
        res = (C.c_double * 8)()
        check(self.lib.cfd_iterate(self._h, n, res))
        return res[0], res[1], res[2], res[3], res[4], res[5]


class _Row:
    """a[i, j] on a single extracted row j (wall_coefficients reads row njm1 only)."""

    def __init__(self, rows):
        self.rows = rows

    def __getitem__(self, ij):
        return self.rows[ij[1]][ij[0]]


_BY_I = ("x", "xu", "sew", "sewu", "dxep", "dxpw", "dxepu", "dxpwu")
