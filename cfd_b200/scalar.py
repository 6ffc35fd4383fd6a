"""Host-side mirror of the reference's scalar-transport demos on a 5-point stencil over the C ABI:

  StagnationPoint          stagnation-point-flow-2d/steady/fvm2d_stag_point.f90    (UDS / CDS; tdma_we / tdma_sn)
  StagnationPointUnsteady  stagnation-point-flow-2d/unsteady/stag_point_unsteady.f90 (explicit / implicit Euler / CN)
  ObliqueConvection        diagonal-convection-2d/fvm2d_oblique_conv.f90           (UDS step profile, two b.c. sets)

Grid, velocities and the steady coefficients are one-off O(N) set-up (the programs compute them once, before their
solver call) and are evaluated here with numpy in the reference's operation order; everything that repeats -- the
line sweeps, the residual sums of the convergence loops, the per-step time terms -- runs in libcfd_b200.so on the
coefficients uploaded into a context (`LinearSystem5`).  Field slots of the cavity-family context used as plain
storage: phi = "pp", phio = "p", steady ae/aw/an/as = "u"/"v"/"du"/"dv", ct = "den".
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import FIELD_ID, check
from .cavity import CavityFVM


def lin_grid(nicv, njcv, xmin=0.0, xmax=1.0, ymin=0.0, ymax=1.0):
    """Faces x, y, centres xc, yc and the CDS factors fx, fy (fvm2d_stag_point.f90:405-447; expansion factors 1)."""
    ni, nj = nicv + 2, njcv + 2
    x, y, xc, yc, fx, fy = (np.zeros(n) for n in (ni, nj, ni, nj, ni, nj))
    dx, dy = (xmax - xmin) / float(nicv), (ymax - ymin) / float(njcv)
    x[0] = xmin
    for i in range(1, ni - 1):
        x[i] = x[i - 1] + dx
    x[ni - 1] = x[ni - 2]
    xc[0] = x[0]
    xc[1:ni - 1] = 0.5 * (x[1:ni - 1] + x[0:ni - 2])
    xc[ni - 1] = x[ni - 2]
    y[0] = ymin
    for j in range(1, nj - 1):
        y[j] = y[j - 1] + dy
    y[nj - 1] = y[nj - 2]
    yc[0] = y[0]
    yc[1:nj - 1] = 0.5 * (y[1:nj - 1] + y[0:nj - 2])
    yc[nj - 1] = y[nj - 2]
    fx[1:ni - 1] = (x[1:ni - 1] - xc[1:ni - 1]) / (xc[2:ni] - xc[1:ni - 1])
    fy[1:nj - 1] = (y[1:nj - 1] - yc[1:nj - 1]) / (yc[2:nj] - yc[1:nj - 1])
    return x, y, xc, yc, fx, fy


class LinearSystem5:
    """ap*phi = ae*phiE + aw*phiW + an*phiN + as*phiS + su on a context that keeps the raw coefficient arrays."""

    def __init__(self, nicv, njcv, order="sequential", device=-1):
        self.ctx = CavityFVM(nicv, njcv, order=order, device=device, raw_coefficients=True)
        self.ni, self.nj = nicv + 2, njcv + 2
        self.lib, self._h = self.ctx.lib, self.ctx._h

    def upload_coefficients(self, ae, aw, an, as_, ap, su):
        for n, a in (("ae", ae), ("aw", aw), ("an", an), ("as", as_), ("ap", ap), ("su", su)):
            self.ctx.upload(n, a)

    def upload_phi(self, phi):
        self.ctx.upload("pp", phi)

    def phi(self):
        return self.ctx.download("pp")

    def wesweep(self):
        check(self.lib.cfd_lin_wesweep(self._h, FIELD_ID["pp"]))

    def snsweep(self):
        check(self.lib.cfd_lin_snsweep(self._h, FIELD_ID["pp"]))

    def residual(self):
        r = C.c_double()
        check(self.lib.cfd_lin_residual(self._h, FIELD_ID["pp"], C.byref(r)))
        return r.value

    def solve(self, direction="we", variant=0, maxit=1000, tol=1.0e-4):
        """tdma_we / tdma_sn (variant 0) or the oblique program's tdma_solver (variant 1): (iterations, rsm)."""
        it, rsm = C.c_int(), C.c_double()
        check(self.lib.cfd_lin_solve(self._h, FIELD_ID["pp"], 0 if direction == "we" else 1, variant, maxit, tol,
                                     C.byref(it), C.byref(rsm)))
        return it.value, rsm.value

    def close(self):
        self.ctx.close()


def _stag_coefficients(ni, nj, isch, den, gam, x, y, xc, yc, fx, fy, u, v):
    """fvm2d_stag_point.f90:458-490 (the same lines in stag_point_unsteady.f90:590-627), interior nodes."""
    ae, aw, an, as_ = (np.zeros((ni, nj), order="F") for _ in range(4))
    I, J = slice(1, ni - 1), slice(1, nj - 1)
    dyf = (y[1:nj - 1] - y[0:nj - 2])[None, :]
    dxf = (x[1:ni - 1] - x[0:ni - 2])[:, None]
    de = -gam * dyf / (xc[2:ni] - xc[1:ni - 1])[:, None]
    dw = -gam * dyf / (xc[1:ni - 1] - xc[0:ni - 2])[:, None]
    dn = -gam * dxf / (yc[2:nj] - yc[1:nj - 1])[None, :]
    ds = -gam * dxf / (yc[1:nj - 1] - yc[0:nj - 2])[None, :]
    ue, uw, vn, vs = u[1:ni - 1, J], u[0:ni - 2, J], v[I, 1:nj - 1], v[I, 0:nj - 2]
    ge, gw = den * ue * dyf, -den * uw * dyf
    gn, gs = den * vn * dxf, -den * vs * dxf
    if isch == 1:
        ce, cw, cn, cs = (np.minimum(g, 0.0) for g in (ge, gw, gn, gs))
    else:
        ce, cw = ge * fx[1:ni - 1][:, None], gw * (1.0 - fx[0:ni - 2])[:, None]
        cn, cs = gn * fy[1:nj - 1][None, :], gs * (1.0 - fy[0:nj - 2])[None, :]
    ae[I, J], aw[I, J], an[I, J], as_[I, J] = ce + de, cw + dw, cn + dn, cs + ds
    return ae, aw, an, as_


class StagnationPoint:
    """fvm2d_stag_point.f90: steady scalar transport in u = x, v = -y.  isch 1 UDS / 2 CDS, isol 1 tdma_we / 2 tdma_sn."""

    def __init__(self, nicv=20, njcv=20, isch=1, isol=2, den=1.2, gam=0.1, order="sequential", device=-1):
        ni, nj = nicv + 2, njcv + 2
        self.ni, self.nj, self.isol, self.gam = ni, nj, isol, gam
        self.x, self.y, self.xc, self.yc, self.fx, self.fy = lin_grid(nicv, njcv)
        x, y, yc = self.x, self.y, self.yc
        self.u = np.asfortranarray(np.repeat(x[:, None], nj, axis=1))
        self.v = np.asfortranarray(np.repeat(-y[None, :], ni, axis=0))
        phi = np.zeros((ni, nj), order="F")
        phi[0, 0:nj - 1] = 1.0 - (yc[0:nj - 1] - y[0]) / (y[nj - 2] - y[0])
        ae, aw, an, as_ = _stag_coefficients(ni, nj, isch, den, gam, x, y, self.xc, yc, self.fx, self.fy, self.u, self.v)
        ap, su = np.zeros((ni, nj), order="F"), np.zeros((ni, nj), order="F")
        I, J = slice(1, ni - 1), slice(1, nj - 1)
        ap[I, J] = (ae[I, J] + aw[I, J] + an[I, J] + as_[I, J])
        ap[ni - 2, J] = ap[ni - 2, J] - ae[ni - 2, J]          # east: outflow (:496-498)
        ae[ni - 2, J] = 0.0
        ap[I, 1] = ap[I, 1] - as_[I, 1]                        # south: symmetry (:503-505)
        as_[I, 1] = 0.0
        self.sys = LinearSystem5(nicv, njcv, order=order, device=device)
        self.sys.upload_coefficients(ae, aw, an, as_, ap, su)
        self.sys.upload_phi(phi)

    def solve(self):
        self.iters, self.rsm = self.sys.solve("we" if self.isol == 1 else "sn", variant=0, maxit=1000, tol=1.0e-4)
        return self.iters

    def finish(self):
        """Outlet / symmetry copies and the wall flux (fvm2d_stag_point.f90:518-527); returns (phi, fwall)."""
        ni, nj = self.ni, self.nj
        phi = self.sys.phi()
        phi[1:ni - 1, 0] = phi[1:ni - 1, 1]
        phi[ni - 1, :] = phi[ni - 2, :]
        fwall = 0.0
        for j in range(1, nj - 1):
            fwall = fwall + self.gam * (self.y[j] - self.y[j - 1]) * (phi[1, j] - phi[0, j]) / (self.xc[1] - self.xc[0])
        return phi, fwall

    def close(self):
        self.sys.close()


class StagnationPointUnsteady(StagnationPoint):
    """stag_point_unsteady.f90: the same problem marched in time; itsch 1 explicit, 2 implicit Euler, 3 Crank-Nicolson.
    The per-step coefficient update and the phio = phi copy run on the device."""

    def __init__(self, nicv=20, njcv=20, isch=2, itsch=2, isol=2, den=1.2, gam=0.1, dt=1.0e-3, order="sequential", device=-1):
        super().__init__(nicv, njcv, isch=isch, isol=isol, den=den, gam=gam, order=order, device=device)
        ni, nj = self.ni, self.nj
        self.itsch = itsch
        ae, aw, an, as_ = _stag_coefficients(ni, nj, isch, den, gam, self.x, self.y, self.xc, self.yc, self.fx, self.fy,
                                             self.u, self.v)
        ae[ni - 2, 1:nj - 1] = 0.0                              # (:629-630)
        as_[1:ni - 1, 1] = 0.0
        ct = np.zeros((ni, nj), order="F")
        dxf = (self.x[1:ni - 1] - self.x[0:ni - 2])[:, None]
        dyf = (self.y[1:nj - 1] - self.y[0:nj - 2])[None, :]
        ct[1:ni - 1, 1:nj - 1] = den / dt * dxf * dyf
        c = self.sys.ctx
        for n, a in (("u", ae), ("v", aw), ("du", an), ("dv", as_), ("den", ct)):
            c.upload(n, a)
        self._base = (C.c_int * 4)(FIELD_ID["u"], FIELD_ID["v"], FIELD_ID["du"], FIELD_ID["dv"])
        self.steps = 0

    def step(self, n=1):
        """n passes of the time loop (stag_point_unsteady.f90:583-700); returns the solver iterations of each step."""
        s, out = self.sys, []
        for _ in range(n):
            check(s.lib.cfd_field_copy(s._h, FIELD_ID["p"], FIELD_ID["pp"]))                       # phio = phi
            check(s.lib.cfd_lin_time_terms(s._h, self.itsch, FIELD_ID["p"], self._base, FIELD_ID["den"]))
            out.append(s.solve("we" if self.isol == 1 else "sn", variant=0, maxit=1000, tol=1.0e-4)[0])
            self.steps += 1
        return out


class ObliqueConvection:
    """fvm2d_oblique_conv.f90: UDS convection of a step profile in a uniform flow oblique to the grid.  Its solver keeps
    every term on the left (as*phiS + ap*phi + an*phiN = su - aw*phiW - ae*phiE): uploading the negated neighbour
    coefficients maps it onto LinearSystem5's form with identical bits."""

    def __init__(self, nicv=20, njcv=20, vel=2.0, east_north_bc=1, order="sequential", device=-1):
        ni, nj = nicv + 2, njcv + 2
        self.ni, self.nj, self.bc = ni, nj, east_north_bc
        self.x, self.y, self.xc, self.yc, _, _ = lin_grid(nicv, njcv)
        x, y = self.x, self.y
        phi = np.zeros((ni, nj), order="F")
        phi[0, 1:nj - 1] = 1.0
        phi[0, 0] = 0.5
        if east_north_bc == 2:
            phi[0:ni - 1, nj - 1] = 1.0
            phi[ni - 1, 0:nj - 1] = 0.0
            phi[ni - 1, nj - 1] = 0.5
        I, J = slice(1, ni - 1), slice(1, nj - 1)
        dyf = (y[1:nj - 1] - y[0:nj - 2])[None, :] * np.ones((ni - 2, 1))
        dxf = (x[1:ni - 1] - x[0:ni - 2])[:, None] * np.ones((1, nj - 2))
        ae, aw, an, as_, ap, su = (np.zeros((ni, nj), order="F") for _ in range(6))
        ae[I, J], aw[I, J] = np.minimum(vel * dyf, 0.0), np.minimum(-vel * dyf, 0.0)
        an[I, J], as_[I, J] = np.minimum(vel * dxf, 0.0), np.minimum(-vel * dxf, 0.0)
        ap[I, J] = -(ae[I, J] + aw[I, J] + an[I, J] + as_[I, J])
        su[1, J] = su[1, J] - aw[1, J] * phi[0, J]
        aw[1, J] = 0.0
        if east_north_bc == 1:
            ap[ni - 2, J] = ae[ni - 2, J] + ap[ni - 2, J]
        else:
            su[ni - 2, J] = su[ni - 2, J] - ae[ni - 2, J] * phi[ni - 1, J]
        ae[ni - 2, J] = 0.0
        if east_north_bc == 1:
            ap[I, nj - 2] = an[I, nj - 2] + ap[I, nj - 2]
        else:
            su[I, nj - 2] = su[I, nj - 2] - an[I, nj - 2] * phi[I, nj - 1]
        an[I, nj - 2] = 0.0
        su[I, 1] = su[I, 1] - as_[I, 1] * phi[I, 0]
        as_[I, 1] = 0.0
        self.sys = LinearSystem5(nicv, njcv, order=order, device=device)
        self.sys.upload_coefficients(-ae, -aw, -an, -as_, ap, su)
        self.sys.upload_phi(phi)

    def solve(self):
        self.iters, self.rsm = self.sys.solve("we", variant=1, maxit=100, tol=1.0e-4)
        return self.iters

    def finish(self):
        phi = self.sys.phi()
        if self.bc == 1:
            phi[self.ni - 1, :] = phi[self.ni - 2, :]
            phi[:, self.nj - 1] = phi[:, self.nj - 2]
        return phi

    def close(self):
        self.sys.close()
