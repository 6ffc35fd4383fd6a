"""ctypes front-end to the CPU oracle (oracle/*.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package (cfd_b200/) never imports this module.

The oracle restates /root/reference/lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90 and the heat-conduction
stencils (see the headers of cavity_oracle.c / heat_oracle.c for the file:line map).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")
_dp = C.POINTER(C.c_double)


def build(force: bool = False) -> None:
    """Compile the two oracle libraries with the committed Makefile (gcc only)."""
    need = force or not all(
        os.path.exists(os.path.join(_BUILD, n)) for n in ("liboracle.so", "liboracle_fast.so"))
    if not need:
        srcs = [os.path.join(_HERE, f) for f in ("cavity_oracle.c", "heat_oracle.c", "general_oracle.c",
                                                 "teach_oracle.c", "fdcavity_oracle.c", "Makefile")]
        newest = max(os.path.getmtime(s) for s in srcs if os.path.exists(s))
        need = newest > min(os.path.getmtime(os.path.join(_BUILD, n))
                            for n in ("liboracle.so", "liboracle_fast.so"))
    if need:
        subprocess.run(["make", "-C", _HERE, "-s", "-B"], check=True)


class _CavityStruct(C.Structure):
    _fields_ = ([("ni", C.c_int), ("nj", C.c_int), ("iadv", C.c_int),
                 ("nswpu", C.c_int), ("nswpv", C.c_int), ("nswpp", C.c_int),
                 ("ipref", C.c_int), ("jpref", C.c_int), ("order", C.c_int),
                 ("urfu", C.c_double), ("urfv", C.c_double), ("urfp", C.c_double),
                 ("resoru", C.c_double), ("resorv", C.c_double), ("resorm", C.c_double)]
                + [(n, _dp) for n in ("x", "y", "xu", "yv", "sew", "sns", "sewu", "snsv",
                                      "u", "v", "p", "pp", "vis", "den",
                                      "ae", "aw", "an", "as_", "ap", "su", "sp", "du", "dv")])


class _GeneralStruct(C.Structure):
    GEOM_I = ("x", "xu", "xdif", "sew", "sewu", "xcvi", "xcvip", "fx", "fxm")
    GEOM_J = ("y", "yv", "ydif", "sns", "snsv", "ycvr", "ycvrs", "arx", "arxj", "arxjp", "r", "rmn", "sx", "sxmn",
              "fv", "fvp", "fy", "fym")
    FIELDS = ("u", "v", "p", "pp", "du", "dv", "t", "den", "gam", "cp", "ae", "aw", "an", "as_", "ap", "su", "sp")
    _fields_ = ([(n, C.c_int) for n in ("ni", "nj", "icrd", "iadv", "ipref", "jpref", "order", "user")]
                + [("lsolve", C.c_int * 4), ("nsweep", C.c_int * 4), ("urf", C.c_double * 4), ("resor", C.c_double * 4)]
                + [(n, C.c_double) for n in ("xstart", "xend", "ystart", "yend", "rstart")]
                + [(n, C.c_int) for n in ("ibnd", "is1", "is2", "js1", "js2")]
                + [(n, C.c_double) for n in ("t0", "gam0", "gam1", "htc", "tf", "qin", "visc", "cond", "rgbeta", "tref")]
                + [("llast", C.c_int)]
                + [(n, _dp) for n in GEOM_I + GEOM_J + FIELDS])


def _teach_struct(real):
    rp = C.POINTER(real)

    class _S(C.Structure):
        G1 = ("x", "y", "xu", "yv", "r", "rv", "rcv", "sew", "sns", "sewu", "snsv",
              "dxep", "dxpw", "dynp", "dyps", "dxepu", "dxpwu", "dynpv", "dypsv", "yplusn", "xplusw", "taun", "tauw")
        F2 = ("u", "v", "p", "pp", "du", "dv", "den", "vis", "te", "ae", "aw", "an", "as_", "ap", "su", "sp")
        _fields_ = ([(n, C.c_int) for n in ("ni", "nj", "it", "indcos", "istep", "jstep", "ipref", "jpref", "order",
                                             "nswpu", "nswpv", "nswpp")]
                    + [(n, real) for n in ("urfu", "urfv", "urfp", "great", "viscos", "densit", "flowin",
                                           "resoru", "resorv", "resorm")]
                    + [(n, rp) for n in G1 + F2])
    return _S


_TeachF64, _TeachF32 = _teach_struct(C.c_double), _teach_struct(C.c_float)


def _teach_turb_struct(real, base):
    rp = C.POINTER(real)

    class _S(C.Structure):
        F2 = ("ed", "t", "gen", "gamh", "sukd", "spkd")
        _fields_ = ([("b", C.POINTER(base))] + [(n, C.c_int) for n in ("nswpk", "nswpd", "nswpt")]
                    + [(n, real) for n in ("urfk", "urfe", "urft", "urfvis", "cmu", "cd", "c1", "c2", "cappa", "elog",
                                           "pred", "prte", "prandl", "prandt", "pfun", "twalln", "resork", "resore", "resort")]
                    + [(n, rp) for n in F2])
    return _S


_TeachTurbF64, _TeachTurbF32 = _teach_turb_struct(C.c_double, _TeachF64), _teach_turb_struct(C.c_float, _TeachF32)

_libs: dict = {}


def lib(fast: bool = False) -> C.CDLL:
    key = "fast" if fast else "ref"
    if key not in _libs:
        build()
        path = os.path.join(_BUILD, "liboracle_fast.so" if fast else "liboracle.so")
        L = C.CDLL(path)
        L.orc_cavity_run.restype = C.c_int
        L.orc_cavity_run.argtypes = [C.POINTER(_CavityStruct), C.c_int, C.c_double, C.c_int, C.c_int, _dp]
        L.orc_setgrid.argtypes = [C.c_int, C.c_int] + [C.c_double] * 4 + [_dp] * 8
        L.orc_lisolv.argtypes = [C.POINTER(_CavityStruct), C.c_int, C.c_int, _dp]
        for n in ("orc_calcu", "orc_calcv", "orc_calcp", "orc_assemble_u", "orc_assemble_v",
                  "orc_assemble_p", "orc_correct"):
            getattr(L, n).argtypes = [C.POINTER(_CavityStruct)]
            getattr(L, n).restype = None
        L.orc_centred_vel.argtypes = [C.c_int, C.c_int, _dp, _dp, _dp, _dp]
        L.orc_corner_p.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _dp]
        L.orc_heat2d_ftcs.argtypes = [C.c_double] * 14 + [_dp, _dp, C.c_int, C.c_int, C.c_int]
        L.orc_heat2d_ftcs.restype = None
        L.orc_heat2d_lod.argtypes = [C.c_double] * 14 + [_dp, C.c_int, C.c_int, C.c_int]
        L.orc_heat2d_lod.restype = None
        L.orc_heat1d_explicit.argtypes = ([C.c_double] * 7 + [_dp] + [C.c_double] * 4 + [_dp, _dp, C.c_int, C.c_int])
        L.orc_heat1d_explicit.restype = C.c_int
        L.orc_heat1d_implicit.argtypes = L.orc_heat1d_explicit.argtypes
        L.orc_heat1d_implicit.restype = C.c_int
        L.orc_thomas.argtypes = [_dp] * 5 + [C.c_int]
        L.orc_thomas.restype = None
        L.orc_heat2d_coeffs.argtypes = [_dp] * 4 + [C.c_double] * 2 + [_dp] * 5 + [C.c_int, C.c_int]
        L.orc_heat2d_coeffs.restype = None
        L.orc_heat2d_apply.argtypes = [_dp] * 7 + [C.c_int, C.c_int]
        L.orc_heat2d_apply.restype = None
        L.orc_heat2d_ftcs_var.argtypes = [_dp] * 5 + [C.c_double] * 7 + [_dp, C.c_int, C.c_int, C.c_int]
        L.orc_heat2d_ftcs_var.restype = None
        L.orc_heat2d_gs.argtypes = [_dp] * 5 + [C.c_double] * 6 + [C.c_int, C.c_int, _dp, C.c_int, C.c_int, _dp]
        L.orc_heat2d_gs.restype = C.c_int
        for n in ("orc_gen_grid", "orc_gen_boundary", "orc_gen_initial", "orc_gen_assemble_u", "orc_gen_assemble_v",
                  "orc_gen_assemble_p", "orc_gen_assemble_t", "orc_gen_correct", "orc_gen_calcu", "orc_gen_calcv",
                  "orc_gen_calcp", "orc_gen_calct"):
            getattr(L, n).argtypes = [C.POINTER(_GeneralStruct)]
            getattr(L, n).restype = None
        L.orc_gen_tdma.argtypes = [C.POINTER(_GeneralStruct), C.c_int, C.c_int, _dp]
        L.orc_gen_tdma.restype = None
        L.orc_gen_run.argtypes = [C.POINTER(_GeneralStruct), C.c_int, C.c_double, C.c_int, C.c_int, _dp]
        L.orc_gen_run.restype = C.c_int
        _libs[key] = L
    return _libs[key]


def _p(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags["F_CONTIGUOUS"] or a.ndim == 1
    return a.ctypes.data_as(_dp)


QTHR_REF = float(np.float32(0.1))   # the reference compares x with the default-real literal 0.1


class Cavity:
    """State of `program cavity_fvm` (cavity_fvm.f90:509-552): fields as (ni,nj) Fortran-ordered arrays."""

    FIELDS = ("u", "v", "p", "pp", "vis", "den", "ae", "aw", "an", "as_", "ap", "su", "sp", "du", "dv")
    GEOM_I = ("x", "xu", "sew", "sewu")
    GEOM_J = ("y", "yv", "sns", "snsv")

    def __init__(self, nicv=10, njcv=10, reynolds=100.0, iadv=3, order=0, fast=False,
                 urf=(0.5, 0.5, 0.8), nswp=(3, 3, 3), utop=1.0, rho=1000.0,
                 xlim=(0.0, 1.0), ylim=(0.0, 1.0)):
        self.L = lib(fast)
        self.ni, self.nj = nicv + 2, njcv + 2
        ni, nj = self.ni, self.nj
        for n in self.GEOM_I:
            setattr(self, n, np.zeros(ni))
        for n in self.GEOM_J:
            setattr(self, n, np.zeros(nj))
        for n in self.FIELDS:
            setattr(self, n, np.zeros((ni, nj), order="F"))
        self.L.orc_setgrid(nicv, njcv, xlim[0], xlim[1], ylim[0], ylim[1],
                           _p(self.x), _p(self.y), _p(self.xu), _p(self.yv),
                           _p(self.sew), _p(self.sns), _p(self.sewu), _p(self.snsv))
        # cavity_fvm.f90:534-538
        self.den[:, :] = rho
        self.vis[:, :] = self.den * utop * (xlim[1] - xlim[0]) / reynolds
        self.u[1:, nj - 1] = utop
        self.ipref, self.jpref = ni // 2, nj // 2          # cavity_fvm.f90:549-552 (1-based)
        self.imon, self.jmon = self.ipref, self.jpref
        self.s = _CavityStruct(ni=ni, nj=nj, iadv=iadv, nswpu=nswp[0], nswpv=nswp[1], nswpp=nswp[2],
                               ipref=self.ipref, jpref=self.jpref, order=order,
                               urfu=urf[0], urfv=urf[1], urfp=urf[2])
        for n in self.GEOM_I + self.GEOM_J + self.FIELDS:
            setattr(self.s, n, _p(getattr(self, n)))

    def run(self, maxit=90000, tol=1.0e-10, history=False):
        hist = np.zeros((maxit, 5)) if history else None
        n = self.L.orc_cavity_run(C.byref(self.s), maxit, tol, self.imon, self.jmon,
                                  hist.ctypes.data_as(_dp) if history else None)
        return (n, hist[:n]) if history else n

    def calcu(self):
        self.L.orc_calcu(C.byref(self.s)); return self.s.resoru

    def calcv(self):
        self.L.orc_calcv(C.byref(self.s)); return self.s.resorv

    def calcp(self):
        self.L.orc_calcp(C.byref(self.s)); return self.s.resorm

    def assemble_u(self):
        self.L.orc_assemble_u(C.byref(self.s)); return self.s.resoru

    def assemble_v(self):
        self.L.orc_assemble_v(C.byref(self.s)); return self.s.resorv

    def assemble_p(self):
        self.L.orc_assemble_p(C.byref(self.s)); return self.s.resorm

    def correct(self):
        self.L.orc_correct(C.byref(self.s))

    def lisolv(self, istart, jstart, phi):
        self.L.orc_lisolv(C.byref(self.s), istart, jstart, _p(phi))

    def centred_vel(self):
        uc = np.zeros_like(self.u); vc = np.zeros_like(self.v)
        self.L.orc_centred_vel(self.ni, self.nj, _p(self.u), _p(self.v), _p(uc), _p(vc))
        return uc, vc

    @property
    def residuals(self):
        return self.s.resoru, self.s.resorv, self.s.resorm


def heat2d_ftcs(tc, nsteps, k=5.0, rho=1.0, cp=1.0, lim=(0.0, 1.0, 0.0, 1.0), tbc=(0.0, 0.0, 0.0, 0.0),
                dt=None, fo=0.2, qthr=QTHR_REF, qval=1.0e3, fast=False):
    """nsteps FTCS steps on tc (nx,ny) Fortran-ordered; returns the new field (tc is not modified)."""
    tc = np.array(tc, dtype=np.float64, order="F", copy=True)
    nx, ny = tc.shape
    dx = (lim[1] - lim[0]) / (nx - 1)
    if dt is None:
        dt = fo * dx * dx / (k / rho / cp)
    t = np.zeros_like(tc, order="F")
    lib(fast).orc_heat2d_ftcs(k, rho, cp, lim[0], lim[1], lim[2], lim[3], tbc[0], tbc[1], tbc[2], tbc[3],
                              dt, qthr, qval, _p(tc), _p(t), nx, ny, nsteps)
    return tc


def heat2d_lod(tc, nsteps, k=5.0, rho=1.0, cp=1.0, lim=(0.0, 1.0, 0.0, 1.0), tbc=(0.0, 0.0, 0.0, 0.0),
               dt=1.0e-4, qthr=QTHR_REF, qval=1.0e3, fast=False):
    tc = np.array(tc, dtype=np.float64, order="F", copy=True)
    nx, ny = tc.shape
    lib(fast).orc_heat2d_lod(k, rho, cp, lim[0], lim[1], lim[2], lim[3], tbc[0], tbc[1], tbc[2], tbc[3],
                             dt, qthr, qval, _p(tc), nx, ny, nsteps)
    return tc


class HeatVar:
    """The variable-conductivity 5-point operator of steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:162-190:
    `lam(x, y)` and `qdot(x, y)` are the program's functions LAMBDA and QDOT (vectorised callables); coefficient arrays
    as interior() computes them."""

    def __init__(self, nx, ny, lam, qdot, lim=(0.0, 0.03, 0.0, 0.02), fast=False):
        self.L, self.nx, self.ny = lib(fast), nx, ny
        self.dx, self.dy = (lim[1] - lim[0]) / (nx - 1), (lim[3] - lim[2]) / (ny - 1)
        x = lim[0] + np.arange(nx) * self.dx                              # xymesh (:112-129)
        y = lim[2] + np.arange(ny) * self.dy
        x[0], y[0] = lim[0], lim[2]
        self.x, self.y = x, y
        X, Y = x[:, None] + 0.0 * y[None, :], 0.0 * x[:, None] + y[None, :]
        xm, xp = np.zeros_like(X), np.zeros_like(X)
        ym, yp = np.zeros_like(X), np.zeros_like(X)
        xm[1:, :] = 0.5 * (x[1:] + x[:-1])[:, None]                        # :171-174: face mid-points
        xp[:-1, :] = 0.5 * (x[:-1] + x[1:])[:, None]
        ym[:, 1:] = 0.5 * (y[1:] + y[:-1])[None, :]
        yp[:, :-1] = 0.5 * (y[:-1] + y[1:])[None, :]
        f = lambda a: np.asfortranarray(np.broadcast_to(np.asarray(a, dtype=np.float64), X.shape).copy())
        self.lamw, self.lame = f(lam(xm, Y)), f(lam(xp, Y))
        self.lams, self.lamn = f(lam(X, ym)), f(lam(X, yp))
        self.q = f(qdot(X, Y))
        z = lambda: np.zeros((nx, ny), order="F")
        self.aw, self.ae, self.as_, self.an, self.ap = z(), z(), z(), z(), z()
        self.L.orc_heat2d_coeffs(_p(self.lamw), _p(self.lame), _p(self.lams), _p(self.lamn), self.dx, self.dy,
                                 _p(self.aw), _p(self.ae), _p(self.as_), _p(self.an), _p(self.ap), nx, ny)

    def _c(self):
        return [_p(a) for a in (self.aw, self.ae, self.as_, self.an, self.q)]

    def apply(self, t):
        t = np.asfortranarray(t, dtype=np.float64)
        r = np.zeros_like(t, order="F")
        self.L.orc_heat2d_apply(*self._c(), _p(t), _p(r), self.nx, self.ny)
        return r

    def ftcs(self, t, nsteps, dt, rho=1.0, cp=1.0, tbc=(0.0, 0.0, 0.0, 0.0)):
        t = np.array(t, dtype=np.float64, order="F", copy=True)
        self.L.orc_heat2d_ftcs_var(*self._c(), rho, cp, dt, tbc[0], tbc[1], tbc[2], tbc[3], _p(t), self.nx, self.ny, nsteps)
        return t

    def gauss_seidel(self, x0=None, tbc=(300.0, 300.0, 300.0, 300.0), rf=0.5, tol=1.0e-7, imax=500, order=0):
        """tbc = (tleft, tright, tbottom, ttop); returns (x, passes, last norm)."""
        x = np.zeros((self.nx, self.ny), order="F") if x0 is None else np.array(x0, dtype=np.float64, order="F", copy=True)
        d = C.c_double()
        n = self.L.orc_heat2d_gs(*self._c(), tbc[0], tbc[1], tbc[2], tbc[3], rf, tol, imax, order, _p(x), self.nx, self.ny,
                                 C.byref(d))
        return x, n, d.value


def thomas(a, b, c, d):
    a, b, c, d = (np.ascontiguousarray(v, dtype=np.float64) for v in (a, b, c, d))
    x = np.zeros_like(d)
    lib().orc_thomas(_p(a), _p(b), _p(c), _p(d), _p(x), len(d))
    return x


class General:
    """State of the general SIMPLE family (modules vars/user/serv, cylin_cond.f90:42-287; program main :289-312).

    user="cylinder": the conduction case exactly as shipped (cylin_cond.f90:83-169).
    user="buoyancy": the Boussinesq cavity authored by this build (the reference's own `user` module is missing).
    """

    def __init__(self, user="cylinder", ni=52, nj=52, order=0, fast=False, iadv=None, urf=None, nsweep=(3, 3, 3, 3),
                 ra=1.0e6, pr=0.71, alpha=2.3e-5):
        self.L = lib(fast)
        S = _GeneralStruct
        self.ni, self.nj = ni, nj
        for n in S.GEOM_I:
            setattr(self, n, np.zeros(ni))
        for n in S.GEOM_J:
            setattr(self, n, np.zeros(nj))
        for n in S.FIELDS:
            setattr(self, n, np.zeros((ni, nj), order="F"))
        s = S(ni=ni, nj=nj, order=order, ipref=1, jpref=1)
        if user == "cylinder":           # control(), cylin_cond.f90:83-115 and initial() :117-134
            s.user, s.icrd, s.iadv = 1, 2, 1 if iadv is None else iadv
            s.xstart, s.xend, s.ystart, s.yend, s.rstart = 0.0, 2.0, 0.0, 1.0, 1.0
            s.lsolve[:] = [0, 0, 0, 1]
            u4 = urf or (0.3, 0.3, 0.9, 1.0)
            s.ibnd = 2
            s.is1, s.is2, s.js2 = ni // 2 - 10, ni // 2 + 10, nj
            s.js1 = s.js2 - 20
            s.gam0, s.gam1, s.qin, s.htc, s.t0, s.tf = 3.0, 1.0, 50.0, 50.0, 20.0, 10.0
        else:                            # authored Boussinesq cavity (SURVEY 8d input 4)
            s.user, s.icrd, s.iadv = 2, 1, 3 if iadv is None else iadv
            s.xstart, s.xend, s.ystart, s.yend, s.rstart = 0.0, 1.0, 0.0, 1.0, 1.0
            s.lsolve[:] = [1, 0, 0, 1]
            u4 = urf or (0.3, 0.3, 0.9, 1.0)
            nu = pr * alpha
            s.visc, s.cond = nu, alpha              # rho = cp = 1
            s.rgbeta = ra * nu * alpha              # rho*g*beta = Ra*nu*alpha/(dT*L^3), dT = L = 1
            s.tref = 0.5
            s.ipref, s.jpref = ni // 2, nj // 2
        for k in range(4):
            s.urf[k] = u4[k]
            s.nsweep[k] = nsweep[k]
        self.imon, self.jmon = ni // 2, nj // 2
        for n in S.GEOM_I + S.GEOM_J + S.FIELDS:
            setattr(s, n, _p(getattr(self, n)))
        self.s = s
        self.L.orc_gen_grid(C.byref(s))
        self.L.orc_gen_initial(C.byref(s))

    def run(self, max_iter=19999, tol=1.0e-10, history=False):
        hist = np.zeros((max_iter, 5)) if history else None
        n = self.L.orc_gen_run(C.byref(self.s), max_iter, tol, self.imon, self.jmon,
                               hist.ctypes.data_as(_dp) if history else None)
        return (n, hist[:n]) if history else n

    def call(self, name, *a):
        getattr(self.L, "orc_gen_" + name)(C.byref(self.s), *a)

    def tdma(self, istart, jstart, phi):
        self.L.orc_gen_tdma(C.byref(self.s), istart, jstart, _p(phi))

    @property
    def resor(self):
        return tuple(self.s.resor)


class Teach:
    """Laminar subset of the TEACH-T sudden-expansion program (backward-step-flow-2d/: control.f90, general.f90,
    calc.f90, promod.f90).  dtype float64 = parity oracle of the CUDA path; float32 = the reference as written."""

    def __init__(self, ni=100, nj=100, it=None, dtype=np.float64, order=0, fast=False, epsx=1.05, indcos=2,
                 rlarge=0.0124, uin=0.5, densit=1000.0, viscos=0.8e-3, urf=(0.5, 0.5, 1.0), nswp=(3, 3, 5), jstep=None):
        self.L = lib(fast)
        self.dtype = np.dtype(dtype)
        f64 = self.dtype == np.float64
        S = _TeachF64 if f64 else _TeachF32
        self.sfx = "_f64" if f64 else "_f32"
        real = C.c_double if f64 else C.c_float
        rp = C.POINTER(real)
        it = it or ni
        self.ni, self.nj, self.it = ni, nj, it
        n1 = 2 * max(it, nj)
        for n in S.G1:
            setattr(self, n, np.zeros(n1, dtype=self.dtype))
        for n in S.F2:
            setattr(self, n, np.zeros((it, nj), dtype=self.dtype, order="F"))
        jstep = jstep if jstep is not None else max(2, int(round(0.08 * nj)))
        s = S(ni=ni, nj=nj, it=it, indcos=indcos, istep=2, jstep=jstep, ipref=2, jpref=2, order=order,
              nswpu=nswp[0], nswpv=nswp[1], nswpp=nswp[2], urfu=urf[0], urfv=urf[1], urfp=urf[2],
              great=1.0e30, viscos=viscos, densit=densit)
        for n in S.G1 + S.F2:
            setattr(s, n, getattr(self, n).ctypes.data_as(rp))
        self.s = s
        ptr = C.POINTER(S)
        for n in ("grid", "init", "initial", "assemble_u", "assemble_v", "assemble_p", "correct", "calcu", "calcv",
                  "calcp", "iterate", "lisolv"):
            fn = getattr(self.L, "orc_teach_" + n + self.sfx)
            fn.restype = None
            fn.argtypes = {"grid": [ptr, real, real], "initial": [ptr, real], "iterate": [ptr, C.c_int],
                           "lisolv": [ptr, C.c_int, C.c_int, rp]}.get(n, [ptr])
        self._call("grid", rlarge, epsx)
        self._call("init")
        self._call("initial", uin)
        self.jstep = jstep

    def _call(self, name, *a):
        getattr(self.L, "orc_teach_" + name + self.sfx)(C.byref(self.s), *a)

    def call(self, name):
        self._call(name)

    def iterate(self, n=1):
        self._call("iterate", n)
        return self.s.resoru, self.s.resorv, self.s.resorm

    def lisolv(self, istart, jstart, phi):
        real = C.c_double if self.dtype == np.float64 else C.c_float
        self._call("lisolv", istart, jstart, phi.ctypes.data_as(C.POINTER(real)))

    @property
    def residuals(self):
        return self.s.resoru, self.s.resorv, self.s.resorm


class TeachTurb(Teach):
    """The TEACH-T program as shipped: k-epsilon + energy on top of `Teach` (calcte, calced, calct, props and the
    wall-function hooks modte / moded / modt; teach_turb.inc)."""

    def __init__(self, ni=100, nj=100, *, urf_turb=(0.7, 0.7, 1.0, 0.7), nswp_turb=(3, 3, 3), tin=0.0, turbin=0.03,
                 alamda=0.005, rlarge=0.0124, uin=0.5, **kw):
        super().__init__(ni, nj, rlarge=rlarge, uin=uin, **kw)
        f64 = self.dtype == np.float64
        S = _TeachTurbF64 if f64 else _TeachTurbF32
        real = C.c_double if f64 else C.c_float
        rp = C.POINTER(real)
        for n in S.F2:
            setattr(self, n, np.zeros((self.it, self.nj), dtype=self.dtype, order="F"))
        q = S(b=C.pointer(self.s), nswpk=nswp_turb[0], nswpd=nswp_turb[1], nswpt=nswp_turb[2],
              urfk=urf_turb[0], urfe=urf_turb[1], urft=urf_turb[2], urfvis=urf_turb[3])
        for n in S.F2:
            setattr(q, n, getattr(self, n).ctypes.data_as(rp))
        self.q = q
        ptr = C.POINTER(S)
        for n in ("turb_initial", "props", "assemble_te", "assemble_ed", "assemble_t", "calcte", "calced", "calct",
                  "iterate_turb"):
            fn = getattr(self.L, "orc_teach_" + n + self.sfx)
            fn.restype = None
            fn.argtypes = {"turb_initial": [ptr] + [real] * 5, "iterate_turb": [ptr, C.c_int]}.get(n, [ptr])
        self._tcall("turb_initial", uin, tin, turbin, alamda, rlarge)

    def _tcall(self, name, *a):
        getattr(self.L, "orc_teach_" + name + self.sfx)(C.byref(self.q), *a)

    def tcall(self, name):
        self._tcall(name)

    def iterate_turb(self, n=1):
        self._tcall("iterate_turb", n)
        return (self.s.resoru, self.s.resorv, self.s.resorm, self.q.resort, self.q.resork, self.q.resore)


class _FdcStruct(C.Structure):
    _fields_ = ([("ni", C.c_int), ("nj", C.c_int)] + [(n, C.c_double) for n in ("re", "dt", "utop", "urfp")]
                + [(n, _dp) for n in ("u", "v", "p", "uo", "vo", "po", "pp", "su")] + [("sip_iters", C.c_int)])


class FDCavity:
    """Explicit finite-difference cavity (lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90:90-255);
    arrays (ni+1, nj+1) Fortran-ordered, the reference's smaller extents are their leading parts."""

    FIELDS = ("u", "v", "p", "uo", "vo", "po", "pp", "su")

    def __init__(self, ni=11, nj=11, re=10.0, dt=1.0e-4, utop=1.0, urfp=0.8, fast=False):
        self.L = lib(fast)
        self.ni, self.nj = ni, nj
        for n in self.FIELDS:
            setattr(self, n, np.zeros((ni + 1, nj + 1), order="F"))
        self.s = _FdcStruct(ni=ni, nj=nj, re=re, dt=dt, utop=utop, urfp=urfp)
        for n in self.FIELDS:
            setattr(self.s, n, _p(getattr(self, n)))
        ptr = C.POINTER(_FdcStruct)
        for n in ("init", "predict", "solve_pp"):
            getattr(self.L, "orc_fdc_" + n).argtypes = [ptr]
            getattr(self.L, "orc_fdc_" + n).restype = None
        for n in ("correct", "step"):
            getattr(self.L, "orc_fdc_" + n).argtypes = [ptr]
            getattr(self.L, "orc_fdc_" + n).restype = C.c_double
        self.L.orc_fdc_init(C.byref(self.s))

    def predict(self):
        self.L.orc_fdc_predict(C.byref(self.s))

    def solve_pp(self):
        self.L.orc_fdc_solve_pp(C.byref(self.s))
        return self.s.sip_iters

    def correct(self):
        return self.L.orc_fdc_correct(C.byref(self.s))

    def step(self):
        return self.L.orc_fdc_step(C.byref(self.s))


class _FdciStruct(C.Structure):
    _fields_ = ([("ni", C.c_int), ("nj", C.c_int), ("steady", C.c_int)]
                + [(n, C.c_double) for n in ("re", "dt", "utop", "urfu", "urfv", "urfp")]
                + [(n, _dp) for n in ("u", "v", "p", "pp", "ae", "aw", "an", "as_", "ap", "su")]
                + [("sip_iters", C.c_int * 3)])


class FDCavityImplicit:
    """Implicit-Euler finite-difference cavity (lid-driven-cavity-2d/fdm-implicit/fdm_implicit_cavity.f90:96-288)."""

    FIELDS = ("u", "v", "p", "pp", "ae", "aw", "an", "as_", "ap", "su")

    def __init__(self, ni=11, nj=11, re=10.0, dt=1.0e-4, utop=1.0, urf=(1.0, 1.0, 0.5), steady=False, fast=False):
        self.L = lib(fast)
        self.ni, self.nj = ni, nj
        for n in self.FIELDS:
            setattr(self, n, np.zeros((ni + 1, nj + 1), order="F"))
        self.s = _FdciStruct(ni=ni, nj=nj, steady=int(steady), re=re, dt=dt, utop=utop, urfu=urf[0], urfv=urf[1], urfp=urf[2])
        for n in self.FIELDS:
            setattr(self.s, n, _p(getattr(self, n)))
        ptr = C.POINTER(_FdciStruct)
        for n in ("init", "assemble_u", "assemble_v", "assemble_p"):
            getattr(self.L, "orc_fdci_" + n).argtypes = [ptr]
            getattr(self.L, "orc_fdci_" + n).restype = None
        self.L.orc_fdci_solve.argtypes = [ptr, C.c_int]
        self.L.orc_fdci_solve.restype = C.c_int
        for n in ("correct", "step"):
            getattr(self.L, "orc_fdci_" + n).argtypes = [ptr]
            getattr(self.L, "orc_fdci_" + n).restype = C.c_double
        self.L.orc_fdci_init(C.byref(self.s))

    def call(self, name):
        return getattr(self.L, "orc_fdci_" + name)(C.byref(self.s))

    def solve(self, which):
        return self.L.orc_fdci_solve(C.byref(self.s), which)

    def step(self):
        return self.L.orc_fdci_step(C.byref(self.s))


def sipsol_arr(aw, ae, as_, an, ap, su, phi, i2=None, j2=None, maxit=3000, alpha=0.80, tol=1.0e-6, fast=False):
    """orc_sipsol_arr: sipsol of fdm_implicit_cavity.f90:29-92 on array coefficients (the four neighbour arrays are negated
    inside, as the program's call does); phi (Fortran order) updated in place; returns the iterations used."""
    L = lib(fast)
    L.orc_sipsol_arr.restype = C.c_int
    L.orc_sipsol_arr.argtypes = [C.c_int] * 4 + [_dp] * 7 + [C.c_int, C.c_double, C.c_double]
    ld, nrow = phi.shape
    arrs = [np.asfortranarray(a, dtype=np.float64) for a in (aw, ae, as_, an, ap, su)]
    return L.orc_sipsol_arr(ld, nrow, ld if i2 is None else i2, nrow if j2 is None else j2, *(_p(a) for a in arrs), _p(phi),
                            maxit, alpha, tol)


# ---- scalar-transport demos (oracle/lin_oracle.c) ------------------------------------------------------------------
class _LinBase:
    COEF = ("aw", "ae", "as_", "an", "ap", "su")

    def _alloc(self, nicv, njcv, fast):
        self.L = lib(fast)
        self.nicv, self.njcv = nicv, njcv
        self.ni, self.nj = nicv + 2, njcv + 2
        for n, m in (("x", self.ni), ("y", self.nj), ("xc", self.ni), ("yc", self.nj), ("fx", self.ni), ("fy", self.nj)):
            setattr(self, n, np.zeros(m))
        for n in ("u", "v", "phi", "phio", "ae", "aw", "an", "as_", "ap", "su"):
            setattr(self, n, np.zeros((self.ni, self.nj), order="F"))
        self.L.orc_lin_grid.argtypes = [C.c_int, C.c_int] + [C.c_double] * 4 + [_dp] * 6
        self.L.orc_lin_grid.restype = None
        self.L.orc_lin_grid(nicv, njcv, 0.0, 1.0, 0.0, 1.0, *(_p(getattr(self, n)) for n in ("x", "y", "xc", "yc", "fx", "fy")))

    def _coef(self):
        return [_p(getattr(self, n)) for n in self.COEF]


class StagPoint(_LinBase):
    """fvm2d_stag_point.f90 / stag_point_unsteady.f90 restated (oracle/lin_oracle.c)."""

    def __init__(self, nicv=20, njcv=20, isch=1, isol=2, den=1.2, gam=0.1, order=0, fast=False):
        self._alloc(nicv, njcv, fast)
        self.isch, self.isol, self.den, self.gam, self.order = isch, isol, den, gam, order
        L = self.L
        L.orc_stag_setup.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double] + [_dp] * 15
        L.orc_stag_setup.restype = None
        L.orc_stag_setup(nicv, njcv, isch, den, gam, *(_p(getattr(self, n)) for n in
                         ("x", "y", "xc", "yc", "fx", "fy", "u", "v", "phi", "ae", "aw", "an", "as_", "ap", "su")))
        for n in ("orc_stag_wesweep", "orc_stag_snsweep"):
            getattr(L, n).argtypes = [C.c_int] * 3 + [_dp] * 7
            getattr(L, n).restype = None
        L.orc_stag_residual.argtypes = [C.c_int] * 2 + [_dp] * 7
        L.orc_stag_residual.restype = C.c_double
        L.orc_stag_solve.argtypes = [C.c_int] * 5 + [C.c_double] + [_dp] * 8
        L.orc_stag_solve.restype = C.c_int
        L.orc_stag_finish.argtypes = [C.c_int, C.c_int, C.c_double, _dp, _dp, _dp]
        L.orc_stag_finish.restype = C.c_double
        L.orc_stag_time_terms.argtypes = [C.c_int] * 4 + [C.c_double] * 3 + [_dp] * 15
        L.orc_stag_time_terms.restype = None
        L.orc_ustag_solve.argtypes = [C.c_int] * 5 + [C.c_double] + [_dp] * 8
        L.orc_ustag_solve.restype = C.c_int

    def wesweep(self):
        self.L.orc_stag_wesweep(self.ni, self.nj, self.order, *self._coef(), _p(self.phi))

    def snsweep(self):
        self.L.orc_stag_snsweep(self.ni, self.nj, self.order, *self._coef(), _p(self.phi))

    def residual(self):
        return self.L.orc_stag_residual(self.ni, self.nj, *self._coef(), _p(self.phi))

    def solve(self, maxit=1000, tol=1.0e-4):
        rsm = C.c_double()
        it = self.L.orc_stag_solve(self.ni, self.nj, self.order, 0 if self.isol == 1 else 1, maxit, tol, *self._coef(),
                                   _p(self.phi), C.byref(rsm))
        self.rsm = rsm.value
        return it

    def finish(self):
        return self.L.orc_stag_finish(self.ni, self.nj, self.gam, _p(self.y), _p(self.xc), _p(self.phi))

    def step(self, itsch, dt=1.0e-3):
        """One pass of the time loop of stag_point_unsteady.f90:583-700; returns the solver's iteration count."""
        self.phio[:, :] = self.phi
        self.L.orc_stag_time_terms(self.nicv, self.njcv, self.isch, itsch, self.den, self.gam, dt,
                                   *(_p(getattr(self, n)) for n in ("x", "y", "xc", "yc", "fx", "fy", "u", "v", "phio",
                                                                   "ae", "aw", "an", "as_", "ap", "su")))
        rsm = C.c_double()      # the unsteady file's own solver copies (all terms on the left)
        it = self.L.orc_ustag_solve(self.ni, self.nj, self.order, 0 if self.isol == 1 else 1, 1000, 1.0e-4, *self._coef(),
                                    _p(self.phi), C.byref(rsm))
        self.rsm = rsm.value
        return it


class Oblique(_LinBase):
    """fvm2d_oblique_conv.f90 restated (oracle/lin_oracle.c)."""

    def __init__(self, nicv=20, njcv=20, vel=2.0, east_north_bc=1, order=0, fast=False):
        self._alloc(nicv, njcv, fast)
        self.order, self.bc = order, east_north_bc
        L = self.L
        L.orc_oblique_setup.argtypes = [C.c_int, C.c_int, C.c_double, C.c_int] + [_dp] * 11
        L.orc_oblique_setup.restype = None
        L.orc_oblique_setup(nicv, njcv, vel, east_north_bc, *(_p(getattr(self, n)) for n in
                            ("x", "y", "u", "v", "phi", "ae", "aw", "an", "as_", "ap", "su")))
        L.orc_oblique_solve.argtypes = [C.c_int] * 4 + [C.c_double] + [_dp] * 8
        L.orc_oblique_solve.restype = C.c_int

    def solve(self, maxit=100, tol=1.0e-4):
        rsm = C.c_double()
        it = self.L.orc_oblique_solve(self.ni, self.nj, self.order, maxit, tol, *self._coef(), _p(self.phi), C.byref(rsm))
        self.rsm = rsm.value
        return it

    def finish(self):
        if self.bc == 1:
            self.phi[self.ni - 1, :] = self.phi[self.ni - 2, :]
            self.phi[:, self.nj - 1] = self.phi[:, self.nj - 2]
        return self.phi
