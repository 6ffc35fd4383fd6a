#!/usr/bin/env python
"""Known-answer tables for the CPU oracle, from a SECOND restatement of the reference programs.

The reference (brli3/CFD, Fortran 90) ships no tests or expected outputs and the build image has no Fortran compiler, so
oracle/*.c cannot be checked against the reference's own binaries.  To take the single-author failure mode out -- one
transcription slip shared by oracle/*.c and the CUDA kernels would pass every GPU-vs-oracle test -- this file restates the
programs once more, written from the .f90 sources (cited per function), NOT from oracle/*.c, and in a different form:
whole-array numpy expressions over index windows instead of cell loops, 1-based arrays padded with an unused index 0,
and numba only for the recurrences (Thomas / SIP), written in P-Q form rather than the reference's a/c form.  It shares no
code with oracle/ or cfd_b200/.  tests/test_oracle_*.py assert the C oracle against the tables written here.

    python tools/gen_kat.py [cavity] [cylinder] [teach] [fdcavity] [stagnation] [buoyancy]     (default: all)

writes tests/golden/<family>_kat.json.  Floating-point sums are accumulated in a different order than the reference's
loops (numpy pairwise sums), so the tables pin residual sums to ~1e-12 relative, fields and iteration counts exactly to the
extent that the iteration is insensitive to that (the counts below are reproduced by the C oracle).
"""
from __future__ import annotations

import json
import os
import sys
import time

import numba
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def W(lo, hi, d=0):
    """Index window lo..hi (inclusive, 1-based) shifted by d: the numpy slice that addresses a(lo+d : hi+d)."""
    return slice(lo + d, hi + d + 1)


# =====================================================================================================================
# Line solver shared by the SIMPLE programs: lisolv (lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90:49-82),
# lisolv (backward-step-flow-2d/general.f90:294-327), tdma (buoyancy-driven-cavity-2d/buoyancy_cav.f90:471-505).
# W->E over lines i = istart..ni-1; each line is the tridiagonal system
#     ap*phi(j) = an*phi(j+1) + as*phi(j-1) + [ae*phi(i+1,j) + aw*phi(i-1,j) + su]       j = jstart..nj-1
# with phi(jstart-1), phi(nj) given.  Thomas in P-Q form: phi(j) = P(j)*phi(j+1) + Q(j).
# =====================================================================================================================
@numba.njit(cache=True)
def line_sweep(istart, jstart, ni, nj, phi, ap, an, as_, ae, aw, su):
    P = np.zeros(nj + 2)
    Q = np.zeros(nj + 2)
    for i in range(istart, ni):
        P[jstart - 1] = 0.0
        Q[jstart - 1] = phi[i, jstart - 1]
        for j in range(jstart, nj):
            rhs = ae[i, j] * phi[i + 1, j] + aw[i, j] * phi[i - 1, j] + su[i, j]
            inv = 1.0 / (ap[i, j] - as_[i, j] * P[j - 1])
            P[j] = an[i, j] * inv
            Q[j] = (rhs + as_[i, j] * Q[j - 1]) * inv
        for j in range(nj - 1, jstart - 1, -1):
            phi[i, j] = P[j] * phi[i, j + 1] + Q[j]


# =====================================================================================================================
# 1. lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90
# =====================================================================================================================
class CavityRef:
    """program cavity_fvm (:509-566) with setgrid (:366-414), calcu (:92-192), calcv (:194-291), calcp (:293-355)."""

    def __init__(self, nicv, njcv, reynolds, iadv=3, urf=(0.5, 0.5, 0.8), nswp=(3, 3, 3)):
        ni, nj = nicv + 2, njcv + 2
        self.ni, self.nj, self.iadv, self.urf, self.nswp = ni, nj, iadv, urf, nswp
        z1 = lambda n: np.zeros(n + 2)
        self.x, self.xu, self.sew, self.sewu = z1(ni), z1(ni), z1(ni), z1(ni)
        self.y, self.yv, self.sns, self.snsv = z1(nj), z1(nj), z1(nj), z1(nj)
        # setgrid (:366-414), xstart = ystart = 0, xend = yend = 1
        dx, dy = 1.0 / nicv, 1.0 / njcv
        self.xu[1] = self.xu[2] = 0.0
        for i in range(3, ni + 1):
            self.xu[i] = self.xu[i - 1] + dx
        self.x[1] = self.xu[2]
        self.x[W(2, ni - 1)] = 0.5 * (self.xu[W(2, ni - 1)] + self.xu[W(2, ni - 1, 1)])
        self.sew[W(2, ni - 1)] = self.xu[W(2, ni - 1, 1)] - self.xu[W(2, ni - 1)]
        self.x[ni] = self.xu[ni]
        self.sewu[W(3, ni - 1)] = self.x[W(3, ni - 1)] - self.x[W(3, ni - 1, -1)]
        self.yv[1] = self.yv[2] = 0.0
        for j in range(3, nj + 1):
            self.yv[j] = self.yv[j - 1] + dy
        self.y[1] = self.yv[2]
        self.y[W(2, nj - 1)] = 0.5 * (self.yv[W(2, nj - 1)] + self.yv[W(2, nj - 1, 1)])
        self.sns[W(2, nj - 1)] = self.yv[W(2, nj - 1, 1)] - self.yv[W(2, nj - 1)]
        self.y[nj] = self.yv[nj]
        self.snsv[W(3, nj - 1)] = self.y[W(3, nj - 1)] - self.y[W(3, nj - 1, -1)]
        z2 = lambda: np.zeros((ni + 2, nj + 2))
        for n in ("u", "v", "p", "pp", "du", "dv", "ae", "aw", "an", "as_", "ap", "su"):
            setattr(self, n, z2())
        self.den = np.full((ni + 2, nj + 2), 1000.0)                       # :536
        self.vis = self.den * 1.0 * 1.0 / reynolds                         # :537
        self.u[2:ni + 1, nj] = 1.0                                         # :538  u(2:,nj) = utop
        self.ipref, self.jpref = ni // 2, nj // 2                          # :549-552
        self.imon, self.jmon = self.ipref, self.jpref
        self.res = [0.0, 0.0, 0.0]

    def _scheme(self, c_e, c_w, c_n, c_s, d_e, d_w, d_n, d_s):
        if self.iadv == 1:                                                 # :141-145
            return d_e - 0.5 * c_e, d_w + 0.5 * c_w, d_n - 0.5 * c_n, d_s + 0.5 * c_s
        if self.iadv == 2:                                                 # :146-150
            return (d_e + np.maximum(0.0, -c_e), d_w + np.maximum(0.0, c_w),
                    d_n + np.maximum(0.0, -c_n), d_s + np.maximum(0.0, c_s))
        return (np.maximum(np.abs(0.5 * c_e), d_e) - 0.5 * c_e, np.maximum(np.abs(0.5 * c_w), d_w) + 0.5 * c_w,
                np.maximum(np.abs(0.5 * c_n), d_n) - 0.5 * c_n, np.maximum(np.abs(0.5 * c_s), d_s) + 0.5 * c_s)

    def calcu(self):
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        u, v, p, den, vis = self.u, self.v, self.p, self.den, self.vis
        I = lambda d=0: W(3, nim1, d)
        J = lambda d=0: W(2, njm1, d)
        col = lambda a, d=0: a[I(d)][:, None]            # by-i geometry as a column
        row = lambda a, d=0: a[J(d)][None, :]            # by-j geometry as a row
        a_ns, a_ew = col(self.sewu), row(self.sns)       # arean = areas = sewu(i); areae = areaw = sns(j)   (:106-110)
        vol = a_ns * a_ew
        g_ee = 0.5 * (den[I(), J()] + den[I(1), J()]) * u[I(1), J()]                      # :112-118
        g_ww = 0.5 * (den[I(-2), J()] + den[I(-1), J()]) * u[I(-1), J()]
        g_ne = 0.5 * (den[I(), J()] + den[I(), J(1)]) * v[I(), J(1)]
        g_nw = 0.5 * (den[I(-1), J()] + den[I(-1), J(1)]) * v[I(-1), J(1)]
        g_se = 0.5 * (den[I(), J(-1)] + den[I(), J()]) * v[I(), J()]
        g_sw = 0.5 * (den[I(-1), J(-1)] + den[I(-1), J()]) * v[I(-1), J()]
        g_p = 0.5 * (den[I(-1), J()] + den[I(), J()]) * u[I(), J()]
        c_e, c_w = 0.5 * (g_p + g_ee) * a_ew, 0.5 * (g_ww + g_p) * a_ew                   # :120-128
        c_n, c_s = 0.5 * (g_nw + g_ne) * a_ns, 0.5 * (g_sw + g_se) * a_ns
        smp = c_e - c_w + c_n - c_s
        cp = np.maximum(smp, 0.0)
        v_e, v_w = vis[I(), J()], vis[I(-1), J()]                                         # :131-135
        v_n = 0.25 * (vis[I(-1), J()] + vis[I(-1), J(1)] + vis[I(), J(1)] + vis[I(), J()])
        v_s = 0.25 * (vis[I(-1), J(-1)] + vis[I(-1), J()] + vis[I(), J()] + vis[I(), J(-1)])
        d_e = v_e * a_ew / (col(self.xu, 1) - col(self.xu))                               # :137-140
        d_w = v_w * a_ew / (col(self.xu) - col(self.xu, -1))
        d_n = v_n * a_ns / (row(self.y, 1) - row(self.y))
        d_s = v_s * a_ns / (row(self.y) - row(self.y, -1))
        ae, aw, an, as_ = self._scheme(c_e, c_w, c_n, c_s, d_e, d_w, d_n, d_s)
        su = (p[I(-1), J()] - p[I(), J()]) * 0.5 * (a_ew + a_ew)                          # :158
        dudxe = (u[I(1), J()] - u[I(), J()]) / col(self.sew)                              # :160-163
        dudxw = (u[I(), J()] - u[I(-1), J()]) / col(self.sew, -1)
        dvdxn = (v[I(), J(1)] - v[I(-1), J(1)]) / a_ns
        dvdxs = (v[I(), J()] - v[I(-1), J()]) / a_ns
        su = su + ((v_e * dudxe - v_w * dudxw) / a_ns + (v_n * dvdxn - v_s * dvdxs) / a_ew) * vol
        su = su + cp * u[I(), J()]
        sp = -cp
        # second loop (:174-187); du uses the scalars areae, areaw LEFT OVER from the first loop's last cell: sns(njm1)
        ap = ae + aw + an + as_ - sp
        r = ap * u[I(), J()] - ae * u[I(1), J()] - aw * u[I(-1), J()] - an * u[I(), J(1)] - as_ * u[I(), J(-1)] - su
        self.res[0] = float(np.abs(r).sum())
        ap = ap / self.urf[0]
        su = su + (1 - self.urf[0]) * ap * u[I(), J()]
        stale = self.sns[njm1]
        self.du[I(), J()] = 0.5 * (stale + stale) / ap
        for name, val in (("ae", ae), ("aw", aw), ("an", an), ("as_", as_), ("ap", ap), ("su", su)):
            getattr(self, name)[I(), J()] = val
        for _ in range(self.nswp[0]):
            line_sweep(3, 2, ni, nj, u, self.ap, self.an, self.as_, self.ae, self.aw, self.su)

    def calcv(self):
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        u, v, p, den, vis = self.u, self.v, self.p, self.den, self.vis
        I = lambda d=0: W(2, nim1, d)
        J = lambda d=0: W(3, njm1, d)
        col = lambda a, d=0: a[I(d)][:, None]
        row = lambda a, d=0: a[J(d)][None, :]
        a_ew, a_ns = row(self.snsv), col(self.sew)       # areae = areaw = snsv(j); arean = areas = sew(i)   (:207-211)
        vol = a_ns * a_ew
        g_ne = 0.5 * (den[I(), J()] + den[I(1), J()]) * u[I(1), J()]                      # :213-219
        g_nw = 0.5 * (den[I(-1), J()] + den[I(), J()]) * u[I(), J()]
        g_nn = 0.5 * (den[I(), J()] + den[I(), J(1)]) * v[I(), J(1)]
        g_se = 0.5 * (den[I(), J(-1)] + den[I(1), J(-1)]) * u[I(1), J(-1)]
        g_sw = 0.5 * (den[I(-1), J(-1)] + den[I(), J(-1)]) * u[I(), J(-1)]
        g_ss = 0.5 * (den[I(), J(-2)] + den[I(), J(-1)]) * v[I(), J(-1)]
        g_p = 0.5 * (den[I(), J(-1)] + den[I(), J()]) * v[I(), J()]
        c_e, c_w = 0.5 * (g_ne + g_se) * a_ew, 0.5 * (g_nw + g_sw) * a_ew                 # :221-229
        c_n, c_s = 0.5 * (g_p + g_nn) * a_ns, 0.5 * (g_ss + g_p) * a_ns
        smp = c_e - c_w + c_n - c_s
        cp = np.maximum(smp, 0.0)
        v_e = 0.25 * (vis[I(), J(-1)] + vis[I(), J()] + vis[I(1), J()] + vis[I(1), J(-1)])   # :233-236
        v_w = 0.25 * (vis[I(-1), J(-1)] + vis[I(-1), J()] + vis[I(), J()] + vis[I(), J(-1)])
        v_n, v_s = vis[I(), J()], vis[I(), J(-1)]
        d_e = v_e * a_ew / (col(self.x, 1) - col(self.x))                                 # :238-241
        d_w = v_w * a_ew / (col(self.x) - col(self.x, -1))
        d_n = v_n * a_ns / (row(self.yv, 1) - row(self.yv))
        d_s = v_s * a_ns / (row(self.yv) - row(self.yv, -1))
        ae, aw, an, as_ = self._scheme(c_e, c_w, c_n, c_s, d_e, d_w, d_n, d_s)
        su = (p[I(), J(-1)] - p[I(), J()]) * 0.5 * (a_ns + a_ns)                          # :258
        dudye = (u[I(1), J()] - u[I(1), J(-1)]) / a_ew                                    # :259-262
        dudyw = (u[I(), J()] - u[I(), J(-1)]) / a_ew
        dvdyn = (v[I(), J(1)] - v[I(), J()]) / row(self.sns)
        dvdys = (v[I(), J()] - v[I(), J(-1)]) / row(self.sns, -1)
        su = su + (v_e * dudye - v_w * dudyw) / a_ns * vol + (v_n * dvdyn - v_s * dvdys) / a_ew * vol   # :263-264
        sp = -cp
        su = su + cp * v[I(), J()]
        ap = ae + aw + an + as_ - sp                                                      # :272-283
        r = ap * v[I(), J()] - ae * v[I(1), J()] - aw * v[I(-1), J()] - an * v[I(), J(1)] - as_ * v[I(), J(-1)] - su
        self.res[1] = float(np.abs(r).sum())
        ap = ap / self.urf[1]
        su = su + (1.0 - self.urf[1]) * ap * v[I(), J()]
        stale = self.sew[nim1]                            # arean, areas left over from the first loop's last cell
        self.dv[I(), J()] = 0.5 * (stale + stale) / ap
        for name, val in (("ae", ae), ("aw", aw), ("an", an), ("as_", as_), ("ap", ap), ("su", su)):
            getattr(self, name)[I(), J()] = val
        for _ in range(self.nswp[1]):
            line_sweep(2, 3, ni, nj, v, self.ap, self.an, self.as_, self.ae, self.aw, self.su)

    def calcp(self):
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        u, v, p, den, du, dv, pp = self.u, self.v, self.p, self.den, self.du, self.dv, self.pp
        I = lambda d=0: W(2, nim1, d)
        J = lambda d=0: W(2, njm1, d)
        a_ew, a_ns = self.sns[J()][None, :], self.sew[I()][:, None]                        # :304-307
        d_e = 0.5 * (den[I(), J()] + den[I(1), J()])                                       # :309-312
        d_w = 0.5 * (den[I(-1), J()] + den[I(), J()])
        d_n = 0.5 * (den[I(), J()] + den[I(), J(1)])
        d_s = 0.5 * (den[I(), J(-1)] + den[I(), J()])
        ae, aw = d_e * a_ew * du[I(1), J()], d_w * a_ew * du[I(), J()]                    # :313-316
        an, as_ = d_n * a_ns * dv[I(), J(1)], d_s * a_ns * dv[I(), J()]
        smp = (d_e * u[I(1), J()]) * a_ew - (d_w * u[I(), J()]) * a_ew + (d_n * v[I(), J(1)]) * a_ns \
            - (d_s * v[I(), J()]) * a_ns                                                   # :317-325
        self.res[2] = float(np.abs(smp).sum())
        for name, val in (("ae", ae), ("aw", aw), ("an", an), ("as_", as_), ("ap", ae + aw + an + as_ - 0.0), ("su", -smp)):
            getattr(self, name)[I(), J()] = val
        pp[:, :] = 0.0                                                                     # :342
        for _ in range(self.nswp[2]):
            line_sweep(2, 2, ni, nj, pp, self.ap, self.an, self.as_, self.ae, self.aw, self.su)
        ppref = pp[self.ipref, self.jpref]                                                 # :346-354
        p[I(), J()] += self.urf[2] * (pp[I(), J()] - ppref)
        Iu = W(3, nim1)
        u[Iu, J()] += du[Iu, J()] * (pp[W(3, nim1, -1), J()] - pp[Iu, J()])
        Jv = W(3, njm1)
        v[I(), Jv] += dv[I(), Jv] * (pp[I(), W(3, njm1, -1)] - pp[I(), Jv])

    def run(self, maxit=90000, tol=1.0e-10, record=()):
        hist = {}
        for it in range(1, maxit + 1):
            self.calcu()
            self.calcv()
            self.calcp()
            if it in record:
                hist[it] = {"res": list(self.res), "u_mon": float(self.u[self.imon, self.jmon]),
                            "v_mon": float(self.v[self.imon, self.jmon])}
            if max(self.res) < tol:                                                        # :565
                break
        return it, hist

    def uc_mon(self):
        return float(0.5 * (self.u[self.imon, self.jmon] + self.u[self.imon + 1, self.jmon]))   # centred_vel :428

    def sums(self):
        ni, nj = self.ni, self.nj
        return [float(np.abs(a[1:ni + 1, 1:nj + 1]).sum()) for a in (self.u, self.v, self.p)]


def gen_cavity():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90, "
                      "independent of oracle/cavity_oracle.c); hybrid scheme, rho 1000, utop 1, urf .5/.5/.8, nswp 3/3/3, "
                      "stop rule max(res) < 1e-10 (cavity_fvm.f90:565)",
           "case": "SURVEY.md section 8c table"}
    c = CavityRef(10, 10, 100.0)
    n, hist = c.run(record=(1, 2, 3, 10, 100))
    out["cv10_re100"] = {"iters": n, "imon": c.imon, "jmon": c.jmon,
                         "history": {str(k): [*v["res"], v["u_mon"], v["v_mon"]] for k, v in hist.items()},
                         "final_res": list(c.res), "final_uv": [float(c.u[c.imon, c.jmon]), float(c.v[c.imon, c.jmon])],
                         "uc_mon": c.uc_mon(), "sum_abs": c.sums()}
    for iadv in (1, 2):
        c = CavityRef(10, 10, 100.0, iadv=iadv)
        n, hist = c.run(maxit=20, tol=0.0, record=(1, 20))
        out[f"cv10_re100_iadv{iadv}"] = {"history": {str(k): [*v["res"], v["u_mon"], v["v_mon"]] for k, v in hist.items()}}
    c = CavityRef(32, 32, 100.0)
    n, _ = c.run()
    uc = 0.5 * (c.u[c.imon, 1:c.nj + 1] + c.u[c.imon + 1, 1:c.nj + 1])
    out["cv32_re100"] = {"iters": n, "min_centreline_uc": float(uc.min()), "sum_abs": c.sums()}
    c = CavityRef(64, 64, 100.0)
    n, _ = c.run()
    out["cv64_re100"] = {"iters": n, "imon": c.imon, "jmon": c.jmon, "final_res": list(c.res),
                         "final_uv": [float(c.u[c.imon, c.jmon]), float(c.v[c.imon, c.jmon])], "uc_mon": c.uc_mon()}
    c = CavityRef(128, 128, 1000.0)
    n, hist = c.run(record=(1000,))
    out["cv128_re1000"] = {"iters": n, "imon": c.imon, "jmon": c.jmon, "iter1000": hist[1000]["res"],
                           "final_uv": [float(c.u[c.imon, c.jmon]), float(c.v[c.imon, c.jmon])], "uc_mon": c.uc_mon()}
    return out



# =====================================================================================================================
# 2. the `vars` programs: cylinder-heat-conduction-2d/cylin_cond.f90 (complete, as shipped) and
#    buoyancy-driven-cavity-2d/buoyancy_cav.f90 (grid, calcu, calcv, calcp, calct, tdma; its `user` module is missing
#    from the reference -- the Boussinesq set-up below repeats the choices documented in DESIGN.md / SURVEY 8a).
#    Face-based: every equation is assembled from whole arrays of east-face and north-face (flow, diff) pairs, the
#    reference's scatter  aw(i+1,j) = acof + max(0,flow); ae(i,j) = aw(i+1,j) - flow  becoming array shifts.
# =====================================================================================================================
def dapec(flow, diff, iadv):
    """DAPEC (cylin_cond.f90:245-269), elementwise."""
    af = np.abs(flow)
    if iadv == 1:
        a = diff - 0.5 * af
    elif iadv == 3:
        a = np.maximum(0.0, diff - 0.5 * af)
    elif iadv == 4:
        with np.errstate(divide="ignore", invalid="ignore"):
            a = np.maximum(0.0, (1.0 - 0.1 * np.abs(flow / diff)) ** 5) * diff
    else:
        a = diff + 0.0 * af
    return np.where(flow == 0.0, diff, a)


def arimean(a, b, w):           # cylin_cond.f90:271-277
    return a * (1 - w) + b * w


def harmean(a, b, w):           # cylin_cond.f90:279-285 (0/0 where both conductivities are zero: never used there)
    with np.errstate(divide="ignore", invalid="ignore"):
        return a * b / (a * (1.0 - w) + b * w)


class GeneralRef:
    def __init__(self, ni, nj, icrd, iadv, lim, urf, nsweep):
        self.ni, self.nj, self.icrd, self.iadv, self.urf, self.nsweep = ni, nj, icrd, iadv, urf, nsweep
        xstart, xend, ystart, yend, rstart = lim
        nim1, nim2, njm1, njm2 = ni - 1, ni - 2, nj - 1, nj - 2
        zi, zj = (lambda: np.zeros(ni + 2)), (lambda: np.zeros(nj + 2))
        x, xu, xdif, sew, sewu, xcvi, xcvip, fx, fxm = (zi() for _ in range(9))
        (y, yv, ydif, sns, snsv, ycvr, ycvrs, arx, arxj, arxjp, r, rmn, sx, sxmn, fv, fvp, fy, fym) = (zj() for _ in range(18))
        # grid (buoyancy_cav.f90:28-120)
        dx = (xend - xstart) / (ni - 2)
        xu[2] = xstart
        for i in range(3, ni + 1):
            xu[i] = xu[i - 1] + dx
        sew[W(2, nim1)] = xu[W(3, ni)] - xu[W(2, nim1)]
        x[1] = xu[2]
        x[W(2, nim1)] = 0.5 * (xu[W(2, nim1)] + xu[W(3, ni)])
        x[ni] = xu[ni]
        xdif[W(2, ni)] = x[W(2, ni)] - x[W(1, nim1)]
        sewu[W(3, nim1)] = xdif[W(3, nim1)]
        sewu[3] += xdif[2]
        sewu[nim1] += xdif[ni]
        xcvi[W(3, nim2)] = 0.5 * sew[W(3, nim2)]
        xcvip[W(3, nim2)] = xcvi[W(3, nim2)]
        xcvip[2] = sew[2]
        xcvi[nim1] = sew[nim1]
        dy = (yend - ystart) / (nj - 2)
        yv[2] = ystart
        for j in range(3, nj + 1):
            yv[j] = yv[j - 1] + dy
        sns[W(2, njm1)] = yv[W(3, nj)] - yv[W(2, njm1)]
        y[1] = yv[2]
        y[W(2, njm1)] = 0.5 * (yv[W(2, njm1)] + yv[W(3, nj)])
        y[nj] = yv[nj]
        ydif[W(2, nj)] = y[W(2, nj)] - y[W(1, njm1)]
        snsv[W(3, njm1)] = ydif[W(3, njm1)]
        snsv[3] += ydif[2]
        snsv[njm1] += ydif[nj]
        r[1] = rstart
        if icrd == 1:
            rmn[1:nj + 1] = 1.0
            r[1:nj + 1] = 1.0
        else:
            for j in range(2, nj + 1):
                r[j] = r[j - 1] + ydif[j]
            rmn[2] = r[1]
            for j in range(3, nj + 1):
                rmn[j] = rmn[j - 1] + sns[j - 1]
        sx[1:nj + 1] = 1.0
        sxmn[1:nj + 1] = 1.0
        if icrd == 3:
            sx[1:nj + 1] = r[1:nj + 1]
            sxmn[2:nj + 1] = rmn[2:nj + 1]
        ycvr[W(2, njm1)] = r[W(2, njm1)] * sns[W(2, njm1)]
        arx[W(2, njm1)] = ycvr[W(2, njm1)]
        if icrd == 3:
            arx[1:nj + 1] = sns[1:nj + 1]
        ycvrs[W(4, njm2)] = 0.5 * (r[W(4, njm2)] + r[W(3, njm2 - 1)]) * ydif[W(4, njm2)]
        ycvrs[3] = 0.5 * (r[3] + r[1]) * snsv[3]
        ycvrs[njm1] = 0.5 * (r[nj] + r[njm2]) * snsv[njm1]
        if icrd == 2:
            arxj[W(3, njm2)] = 0.25 * (1.0 + rmn[W(3, njm2)] / r[W(3, njm2)]) * arx[W(3, njm2)]
            arxjp[W(3, njm2)] = arx[W(3, njm2)] - arxj[W(3, njm2)]
        else:
            arxj[W(3, njm2)] = 0.5 * arx[W(3, njm2)]
            arxjp[W(3, njm2)] = arxj[W(3, njm2)]
        arxjp[2] = arx[2]
        arxj[njm1] = arx[njm1]
        fv[W(3, njm2)] = arxjp[W(3, njm2)] / arx[W(3, njm2)]
        fvp[W(3, njm2)] = 1.0 - fv[W(3, njm2)]
        fx[W(2, ni)] = 0.5 * sew[W(1, nim1)] / xdif[W(2, ni)]
        fxm[W(2, ni)] = 1.0 - fx[W(2, ni)]
        fy[W(2, nj)] = 0.5 * sns[W(1, njm1)] / ydif[W(2, nj)]
        fym[W(2, nj)] = 1.0 - fy[W(2, nj)]
        for k, v in dict(x=x, xu=xu, xdif=xdif, sew=sew, sewu=sewu, xcvi=xcvi, xcvip=xcvip, fx=fx, y=y, yv=yv, ydif=ydif,
                         sns=sns, snsv=snsv, ycvr=ycvr, ycvrs=ycvrs, arx=arx, arxj=arxj, arxjp=arxjp, r=r, rmn=rmn, sx=sx,
                         sxmn=sxmn, fv=fv, fvp=fvp, fy=fy).items():
            setattr(self, k, v)
        z2 = lambda: np.zeros((ni + 2, nj + 2))
        for n in ("u", "v", "p", "pp", "du", "dv", "t", "ae", "aw", "an", "as_", "ap", "su"):
            setattr(self, n, z2())
        self.den = np.ones((ni + 2, nj + 2))            # den_ini = cp_ini = 1 (cylin_cond.f90:118-119 in grid())
        self.cp = np.ones((ni + 2, nj + 2))
        self.resor = [0.0, 0.0, 0.0, 0.0]
        self.ipref = self.jpref = 1

    def _store(self, I, J, ae, aw, an, as_, ap, su):
        for name, val in (("ae", ae), ("aw", aw), ("an", an), ("as_", as_), ("ap", ap), ("su", su)):
            getattr(self, name)[I, J] = val

    # ---- calct (buoyancy_cav.f90:358-433) ---------------------------------------------------------------------------
    def calct(self, gam, su_vol, sp_vol):
        ni, nj, nim1, njm1, iadv = self.ni, self.nj, self.ni - 1, self.nj - 1, self.iadv
        u, v, t = self.u, self.v, self.t
        den = self.den * self.cp                                            # :372
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        arx_, sx_ = row(self.arx), row(self.sx)
        # east faces of cells i = 2..nim1 (the last one is the boundary face, :388-391 with den(nj,j) as written)
        fe = arimean(den[I(), J()], den[I(1), J()], col(self.fx, 1)) * u[I(1), J()] * arx_
        de = harmean(gam[I(), J()], gam[I(1), J()], col(self.fx, 1)) * arx_ / (col(self.xdif, 1) * sx_)
        fe[-1, :] = den[nj, J()] * u[ni, J()] * self.arx[J()]
        de[-1, :] = gam[ni, J()] * self.arx[J()] / (self.xdif[ni] * self.sx[J()])
        aw_next = dapec(fe, de, iadv) + np.maximum(0.0, fe)                 # aw(i+1,j)
        ae = aw_next - fe
        fw0 = den[1, J()] * u[2, J()] * self.arx[J()]                       # west boundary face (:381-385)
        dw0 = gam[1, J()] * self.arx[J()] / (self.xdif[2] * self.sx[J()])
        aw = np.vstack([(dapec(fw0, dw0, iadv) + np.maximum(0.0, fw0))[None, :], aw_next[:-1, :]])
        # north faces of cells j = 2..njm1 (:400-412)
        a_n = row(self.rmn, 1) * col(self.sew)
        fn = arimean(den[I(), J()], den[I(), J(1)], row(self.fy, 1)) * v[I(), J(1)] * a_n
        dn = harmean(gam[I(), J()], gam[I(), J(1)], row(self.fy, 1)) * a_n / row(self.ydif, 1)
        fn[:, -1] = den[I(), nj] * v[I(), nj] * a_n[:, -1]
        dn[:, -1] = gam[I(), nj] * a_n[:, -1] / self.ydif[nj]
        as_next = dapec(fn, dn, iadv) + np.maximum(0.0, fn)
        an = as_next - fn
        a_s = self.r[1] * self.sew[I()]                                     # south boundary face (:373-379)
        fs0 = den[I(), 1] * v[I(), 2] * a_s
        ds0 = gam[I(), 1] * a_s / self.ydif[2]
        as_ = np.hstack([(dapec(fs0, ds0, iadv) + np.maximum(0.0, fs0))[:, None], as_next[:, :-1]])
        vol = row(self.ycvr) * col(self.sew)
        sp, su = sp_vol[I(), J()] * vol, su_vol[I(), J()] * vol
        ap = ae + aw + an + as_ - sp
        rq = ap * t[I(), J()] - ae * t[I(1), J()] - aw * t[I(-1), J()] - an * t[I(), J(1)] - as_ * t[I(), J(-1)] - su
        self.resor[3] = float(np.abs(rq).sum())
        ap = ap / self.urf[3]
        su = su + (1.0 - self.urf[3]) * ap * t[I(), J()]
        self._store(I(), J(), ae, aw, an, as_, ap, su)
        for _ in range(self.nsweep[3]):
            line_sweep(2, 2, ni, nj, t, self.ap, self.an, self.as_, self.ae, self.aw, self.su)

    # ---- calcu (buoyancy_cav.f90:122-198); the i loops run to njm1 as written (ni == nj required) ---------------------
    def calcu(self, gam):
        ni, nj, nim1, njm1, iadv = self.ni, self.nj, self.ni - 1, self.nj - 1, self.iadv
        u, v, p, den = self.u, self.v, self.p, self.den
        I, J = (lambda d=0: W(3, njm1, d)), (lambda d=0: W(2, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        arx_, sx_ = row(self.arx), row(self.sx)
        fl = u[I(), J()] * arimean(den[I(-1), J()], den[I(), J()], col(self.fx))            # :151-158
        flp = u[I(1), J()] * arimean(den[I(), J()], den[I(1), J()], col(self.fx, 1))
        fe = arx_ * 0.5 * (fl + flp)
        de = arx_ * gam[I(), J()] / (col(self.sew) * sx_)
        fe[-1, :] = self.arx[J()] * u[ni, J()] * den[ni, J()]
        de[-1, :] = self.arx[J()] * gam[ni, J()] / (self.sew[nim1] * self.sx[J()])
        aw_next = dapec(fe, de, iadv) + np.maximum(0.0, fe)
        ae = aw_next - fe
        fw0 = self.arx[J()] * u[2, J()] * den[1, J()]                                        # :146-149
        dw0 = self.arx[J()] * gam[1, J()] / (self.sew[2] * self.sx[J()])
        aw = np.vstack([(dapec(fw0, dw0, iadv) + np.maximum(0.0, fw0))[None, :], aw_next[:-1, :]])
        xc, xcm = col(self.xcvi), col(self.xcvip, -1)
        fn = row(self.rmn, 1) * (xc * v[I(), J(1)] * arimean(den[I(), J()], den[I(), J(1)], row(self.fy, 1))
                                 + xcm * v[I(-1), J(1)] * arimean(den[I(-1), J()], den[I(-1), J(1)], row(self.fy, 1)))
        dn = row(self.rmn, 1) * (harmean(gam[I(), J()], gam[I(), J(1)], row(self.fy, 1)) * xc
                                 + harmean(gam[I(-1), J()], gam[I(-1), J(1)], row(self.fy, 1)) * xcm) / row(self.ydif, 1)
        fn[:, -1] = self.rmn[nj] * (xc[:, 0] * v[I(), nj] * den[I(), nj] + xcm[:, 0] * v[I(-1), nj] * den[I(-1), nj])   # :163-166
        dn[:, -1] = self.r[nj] * (xc[:, 0] * gam[I(), nj] + xcm[:, 0] * gam[I(-1), nj]) / self.ydif[nj]
        as_next = dapec(fn, dn, iadv) + np.maximum(0.0, fn)
        an = as_next - fn
        fs0 = self.r[1] * (xc[:, 0] * v[I(), 2] * den[I(), 1] + xcm[:, 0] * v[I(-1), 2] * den[I(-1), 1])     # :135-142
        ds0 = self.r[1] * (xc[:, 0] * gam[I(), 1] + xcm[:, 0] * gam[I(-1), 1]) / self.ydif[2]
        as_ = np.hstack([(dapec(fs0, ds0, iadv) + np.maximum(0.0, fs0))[:, None], as_next[:, :-1]])
        vol = row(self.ycvr) * col(self.sewu)
        du = vol / (col(self.xdif) * sx_)
        su = 0.0 * vol + du * (p[I(-1), J()] - p[I(), J()])                                   # no u sources in either user module
        ap = ae + aw + an + as_
        rq = ap * u[I(), J()] - ae * u[I(1), J()] - aw * u[I(-1), J()] - an * u[I(), J(1)] - as_ * u[I(), J(-1)] - su
        self.resor[0] = float(np.abs(rq).sum())
        ap = ap / self.urf[0]
        su = su + (1 - self.urf[0]) * ap * u[I(), J()]
        self.du[I(), J()] = du / ap
        self._store(I(), J(), ae, aw, an, as_, ap, su)
        for _ in range(self.nsweep[0]):
            line_sweep(3, 2, ni, nj, u, self.ap, self.an, self.as_, self.ae, self.aw, self.su)

    # ---- calcv (buoyancy_cav.f90:200-277) -------------------------------------------------------------------------------
    def calcv(self, gam, su_vol):
        ni, nj, nim1, njm1, iadv = self.ni, self.nj, self.ni - 1, self.nj - 1, self.iadv
        u, v, p, den = self.u, self.v, self.p, self.den
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(3, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        aj, ajm = row(self.arxj), row(self.arxjp, -1)
        fe = aj * u[I(1), J()] * arimean(den[I(), J()], den[I(1), J()], col(self.fx, 1)) \
            + ajm * u[I(1), J(-1)] * arimean(den[I(), J(-1)], den[I(1), J(-1)], col(self.fx, 1))      # :235-239
        de = (harmean(gam[I(), J()], gam[I(1), J()], col(self.fx, 1)) * aj
              + harmean(gam[I(), J(-1)], gam[I(1), J(-1)], col(self.fx, 1)) * ajm) / (col(self.xdif, 1) * row(self.sxmn))
        fe[-1, :] = aj[0] * u[ni, J()] * den[ni, J()] + ajm[0] * u[ni, J(-1)] * den[ni, J(-1)]        # :231-233
        de[-1, :] = (aj[0] * gam[ni, J()] + ajm[0] * gam[ni, J(-1)]) / (self.xdif[ni] * self.sxmn[J()])
        aw_next = dapec(fe, de, iadv) + np.maximum(0.0, fe)
        ae = aw_next - fe
        fw0 = aj[0] * u[2, J()] * den[1, J()] + ajm[0] * u[2, J(-1)] * den[1, J(-1)]                  # :223-228
        dw0 = (aj[0] * gam[1, J()] + ajm[0] * gam[1, J(-1)]) / (self.xdif[2] * self.sxmn[J()])
        aw = np.vstack([(dapec(fw0, dw0, iadv) + np.maximum(0.0, fw0))[None, :], aw_next[:-1, :]])
        sew_ = col(self.sew)
        fl = v[I(), J()] * row(self.rmn) * sew_ * arimean(den[I(), J(-1)], den[I(), J()], row(self.fy))   # :250-254
        flp = v[I(), J(1)] * row(self.rmn, 1) * sew_ * arimean(den[I(), J()], den[I(), J(1)], row(self.fy, 1))
        fn = arimean(fl, flp, row(self.fvp))
        dn = (row(self.r) * sew_) * gam[I(), J()] / row(self.sns)
        a_top = self.r[nj] * self.sew[I()]                                                           # :245-248
        fn[:, -1] = a_top * v[I(), nj] * den[I(), nj]
        dn[:, -1] = a_top * gam[I(), nj] / self.sns[njm1]
        as_next = dapec(fn, dn, iadv) + np.maximum(0.0, fn)
        an = as_next - fn
        a_bot = self.r[1] * self.sew[I()]                                                            # :212-219, `* sns(2)` as written
        fs0 = a_bot * v[I(), 2] * den[I(), 1]
        ds0 = a_bot * gam[I(), 1] * self.sns[2]
        as_ = np.hstack([(dapec(fs0, ds0, iadv) + np.maximum(0.0, fs0))[:, None], as_next[:, :-1]])
        vol = row(self.ycvrs) * sew_
        su = su_vol[I(), J()] * vol
        dv = vol / row(self.ydif)
        su = su + dv * (p[I(), J(-1)] - p[I(), J()])
        ap = ae + aw + an + as_
        rq = ap * v[I(), J()] - ae * v[I(1), J()] - aw * v[I(-1), J()] - an * v[I(), J(1)] - as_ * v[I(), J(-1)] - su
        self.resor[1] = float(np.abs(rq).sum())
        ap = ap / self.urf[1]
        su = su + (1.0 - self.urf[1]) * ap * v[I(), J()]
        self.dv[I(), J()] = dv / ap
        self._store(I(), J(), ae, aw, an, as_, ap, su)
        for _ in range(self.nsweep[1]):
            line_sweep(2, 3, ni, nj, v, self.ap, self.an, self.as_, self.ae, self.aw, self.su)

    # ---- calcp (buoyancy_cav.f90:279-356) -------------------------------------------------------------------------------
    def calcp(self):
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        u, v, p, den, du, dv, pp = self.u, self.v, self.p, self.den, self.du, self.dv, self.pp
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        arx_ = row(self.arx)
        ce = den[I(1), J()] * u[I(1), J()] * arx_                           # east mass flow; den(i+1,j) as written (:319)
        ae = arimean(den[I(), J()], den[I(1), J()], col(self.fx, 1)) * arx_ * du[I(1), J()]
        ae[-1, :] = 0.0                                                     # :313-316
        cw = np.vstack([(den[1, J()] * u[2, J()] * self.arx[J()])[None, :], ce[:-1, :]])        # :308-311 and the scatter :322
        aw = np.vstack([np.zeros((1, njm1 - 1)), ae[:-1, :]])
        a_n = row(self.rmn, 1) * col(self.sew)
        dnn = arimean(den[I(), J()], den[I(), J(1)], row(self.fy, 1))
        cn = dnn * v[I(), J(1)] * a_n
        an = dnn * a_n * dv[I(), J(1)]
        cn[:, -1] = den[I(), nj] * v[I(), nj] * a_n[:, -1]                  # :326-329
        an[:, -1] = 0.0
        cs = np.hstack([(den[I(), 1] * v[I(), 2] * (self.rmn[2] * self.sew[I()]))[:, None], cn[:, :-1]])   # :300-305, :335
        as_ = np.hstack([np.zeros((nim1 - 1, 1)), an[:, :-1]])
        su = ((0.0 + cs) + cw) - ce - cn
        self.resor[2] = float(np.abs(su).sum())
        self._store(I(), J(), ae, aw, an, as_, ae + aw + an + as_, su)
        pp[I(), J()] = 0.0
        for _ in range(self.nsweep[2]):
            line_sweep(2, 2, ni, nj, pp, self.ap, self.an, self.as_, self.ae, self.aw, self.su)
        ppref = pp[self.ipref, self.jpref]
        p[I(), J()] += self.urf[2] * (pp[I(), J()] - ppref)
        Iu, Jv = W(3, nim1), W(3, njm1)
        u[Iu, J()] += du[Iu, J()] * (pp[W(3, nim1, -1), J()] - pp[Iu, J()])
        v[I(), Jv] += dv[I(), Jv] * (pp[I(), W(3, njm1, -1)] - pp[I(), Jv])


def gen_cylinder():
    """cylin_cond.f90 exactly as shipped: control (:83-115), initial (:117-134), gamsor (:149-169), boundary (:140-147),
    program main (:289-312), wall flux of output (:187-190)."""
    ni = nj = 52
    g = GeneralRef(ni, nj, 2, 1, (0.0, 2.0, 0.0, 1.0, 1.0), (0.3, 0.3, 0.9, 1.0), (3, 3, 3, 3))
    nim1, njm1 = ni - 1, nj - 1
    is1, is2, js2 = ni // 2 - 10, ni // 2 + 10, nj
    js1 = js2 - 20
    gam0, gam1, qin, htc, t0, tf = 3.0, 1.0, 50.0, 50.0, 20.0, 10.0
    g.t[2:ni + 1, 1:nj + 1] = t0
    g.t[1, 1:nj + 1] = t0 * (1.0 + g.y[1:nj + 1])
    gam = np.full((ni + 2, nj + 2), gam0)
    su, sp = np.zeros((ni + 2, nj + 2)), np.zeros((ni + 2, nj + 2))
    gam[is1:is2 + 1, js1:js2 + 1] = gam1
    su[is1:is2 + 1, js1:js2 + 1] = 200.0
    sp[is1:is2 + 1, js1:js2 + 1] = -0.1
    gam[2:nim1 + 1, nj] = 0.0
    for i in range(2, nim1 + 1):
        su[i, 2] += g.r[1] / g.arx[2] * tf / (g.ydif[2] / gam[i, 2] + 1.0 / htc)
        sp[i, 2] -= g.r[1] / g.arx[2] / (g.ydif[2] / gam[i, 2] + 1.0 / htc)
        gam[i, 1] = 0.0
    su[nim1, 2:njm1 + 1] += qin / g.sew[nim1]
    gam[ni, 2:njm1 + 1] = 0.0
    imon = jmon = 26
    hist, max_iter, llast, it = {}, 19999, False, 0
    prev = None
    for it in range(1, max_iter + 1):
        if llast:                                       # boundary() acts on the last pass only (ibnd = 2)
            g.t[2:nim1 + 1, nj] = g.t[2:nim1 + 1, njm1]
            g.t[2:nim1 + 1, 1] = (gam0 * g.t[2:nim1 + 1, 2] + htc * g.ydif[2] * tf) / (gam0 + htc * g.ydif[2])
            g.t[ni, 2:njm1 + 1] = g.t[nim1, 2:njm1 + 1] + qin * g.xdif[ni] / gam0
        g.calct(gam, su, sp)
        if it in (1, 2, 10, 100, 1000, 2000, 3000):
            hist[str(it)] = [g.resor[3], float(g.t[imon, jmon])]
        if llast:
            break
        if g.resor[3] < 1.0e-10 or it == max_iter - 1:
            llast = True
        # the iteration becomes bitwise stationary well before max_iter: stop restating once it has
        cur = (g.resor[3], float(g.t[imon, jmon]))
        if it > 3000 and cur == prev and not llast:
            stationary_at = it
            it = max_iter - 1                          # jump to the reference's last-but-one pass
            llast = True
            continue
        prev = cur
    qwall = sum(gam[1, j] * (g.y[j] - g.y[j - 1]) * (g.t[1, j] - g.t[2, j]) / (g.x[2] - g.x[1]) for j in range(2, njm1 + 1))
    T = g.t[1:ni + 1, 1:nj + 1]
    return {"_source": "tools/gen_kat.py (face-based numpy restatement of the complete program "
                       "cylinder-heat-conduction-2d/cylin_cond.f90 as shipped; independent of oracle/general_oracle.c)",
            "history": hist, "iterations": 19999, "final_tres": g.resor[3], "final_t_mon": float(g.t[imon, jmon]),
            "qwall_west": float(qwall), "sum_abs_t": float(np.abs(T).sum()), "t_max": float(T.max()), "t_min": float(T.min())}


def gen_buoyancy():
    """buoyancy_cav.f90 (calcu, calcv, calcp, calct, tdma, grid) with the Boussinesq `user` module this build authored
    (the reference's is missing): unit square, T = 1 / 0 on the west / east wall, adiabatic north / south, rho = cp = 1,
    nu = Pr*alpha, rho*g*beta = Ra*nu*alpha, Tref = 0.5, Ra = 1e6, Pr = 0.71, alpha = 2.3e-5, urf .3/.3/.9/1, nsweep 3."""
    out = {"_source": "tools/gen_kat.py (face-based numpy restatement of buoyancy-driven-cavity-2d/buoyancy_cav.f90 "
                      "calcu/calcv/calcp/calct/tdma/grid; user module authored by this build, DESIGN.md section 7)"}
    for n, iadv, its in ((34, 3, (1, 2, 10, 100)), (34, 4, (1, 5)), (34, 1, (1, 5)), (34, 2, (1, 5)), (66, 3, (1, 20))):
        ra, pr, alpha = 1.0e6, 0.71, 2.3e-5
        visc, cond, tref = pr * alpha, alpha, 0.5
        rgbeta = ra * visc * alpha
        g = GeneralRef(n, n, 1, iadv, (0.0, 1.0, 0.0, 1.0, 1.0), (0.3, 0.3, 0.9, 1.0), (3, 3, 3, 3))
        g.ipref = g.jpref = n // 2
        imon = jmon = n // 2
        g.t[1:n + 1, 1:n + 1] = tref
        g.t[1, 1:n + 1] = 1.0
        g.t[n, 1:n + 1] = 0.0
        gmom = np.full((n + 2, n + 2), visc)
        gt = np.full((n + 2, n + 2), cond)
        gt[2:n, n] = 0.0
        gt[2:n, 1] = 0.0
        zero = np.zeros((n + 2, n + 2))
        hist = {}
        for it in range(1, max(its) + 1):
            g.calcu(gmom)
            sv = np.zeros((n + 2, n + 2))
            Iv, Jv = W(2, n - 1), W(3, n - 1)
            sv[Iv, Jv] = rgbeta * (arimean(g.t[Iv, W(3, n - 1, -1)], g.t[Iv, Jv], g.fy[Jv][None, :]) - tref)
            g.calcv(gmom, sv)
            g.calcp()
            g.calct(gt, zero, zero)
            if it in its:
                hist[str(it)] = {"resor": list(g.resor), "u_mon": float(g.u[imon, jmon]), "v_mon": float(g.v[imon, jmon]),
                                 "p_mon": float(g.p[imon + 1, jmon + 2]), "t_mon": float(g.t[imon, jmon]),
                                 # t without its adiabatic boundary rows: the program fills them on its last pass only
                                 "sum_abs": [float(np.abs(a[1:n + 1, 1:n + 1]).sum()) for a in (g.u, g.v, g.p)]
                                 + [float(np.abs(g.t[1:n + 1, 2:n]).sum())]}
        out[f"n{n}_iadv{iadv}"] = {"ni": n, "nj": n, "iadv": iadv, "imon": imon, "jmon": jmon, "history": hist}
    return out



# =====================================================================================================================
# 3. backward-step-flow-2d/ (TEACH-T), laminar subset: contro (control.f90:15-204), init (general.f90:69-161),
#    calcu / calcv / calcp (calc.f90:20-305), modu / modv (promod.f90:26-133), lisolv (general.f90:294-327).
#    Double precision (the reference is default REAL; the C oracle and the CUDA path are FP64).
# =====================================================================================================================
class TeachRef:
    def __init__(self, ni, nj, jstep, epsx=1.05, indcos=2, rlarge=0.0124, uin=0.5, densit=1000.0, viscos=0.8e-3,
                 urf=(0.5, 0.5, 1.0), nswp=(3, 3, 5)):
        self.ni, self.nj, self.urf, self.nswp, self.viscos, self.indcos = ni, nj, urf, nswp, viscos, indcos
        nim1, njm1 = ni - 1, nj - 1
        self.istep, self.jstep, self.great = 2, jstep, 1.0e30
        istp1, jstp1 = 3, jstep + 1
        zi, zj = (lambda: np.zeros(ni + 2)), (lambda: np.zeros(nj + 2))
        x, y = zi(), zj()
        altot = 16.0 * rlarge                                              # control.f90:72-90
        sumx = 0.5 * epsx ** (ni - 4) + (epsx ** (ni - 3) - 1.0) / (epsx - 1.0) + 0.5 if epsx != 1.0 else float(ni - 2)
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
        # init (general.f90:69-140)
        r = y.copy() if indcos != 1 else np.concatenate(([0.0], np.ones(nj), [0.0]))
        dxep, dxpw, dynp, dyps = zi(), zi(), zj(), zj()
        dxep[W(1, nim1)] = x[W(2, ni)] - x[W(1, nim1)]
        dxpw[W(2, ni)] = dxep[W(1, nim1)]
        dynp[W(1, njm1)] = y[W(2, nj)] - y[W(1, njm1)]
        dyps[W(2, nj)] = dynp[W(1, njm1)]
        sew, sns = zi(), zj()
        sew[W(2, nim1)] = 0.5 * (dxep[W(2, nim1)] + dxpw[W(2, nim1)])
        sns[W(2, njm1)] = 0.5 * (dynp[W(2, njm1)] + dyps[W(2, njm1)])
        xu = zi()
        xu[W(2, ni)] = 0.5 * (x[W(2, ni)] + x[W(1, nim1)])
        dxepu, dxpwu, sewu = zi(), zi(), zi()
        dxepu[W(2, nim1)] = xu[W(3, ni)] - xu[W(2, nim1)]
        dxpwu[W(3, ni)] = dxepu[W(2, nim1)]
        sewu[W(3, nim1)] = 0.5 * (dxepu[W(3, nim1)] + dxpwu[W(3, nim1)])
        rv, rcv, yv = zj(), zj(), zj()
        rv[W(2, nj)] = 0.5 * (r[W(2, nj)] + r[W(1, njm1)])
        for j in range(2, nj + 1):
            rcv[j] = 0.5 * (rv[j] + rv[j - 1])
        yv[W(2, nj)] = 0.5 * (y[W(2, nj)] + y[W(1, njm1)])
        dynpv, dypsv, snsv = zj(), zj(), zj()
        dynpv[W(2, njm1)] = yv[W(3, nj)] - yv[W(2, njm1)]
        dypsv[W(3, nj)] = dynpv[W(2, njm1)]
        snsv[W(3, njm1)] = 0.5 * (dynpv[W(3, njm1)] + dypsv[W(3, njm1)])
        for k, v in dict(x=x, y=y, r=r, rv=rv, rcv=rcv, yv=yv, xu=xu, sew=sew, sns=sns, sewu=sewu, snsv=snsv, dxep=dxep,
                         dxpw=dxpw, dynp=dynp, dyps=dyps, dxepu=dxepu, dxpwu=dxpwu, dynpv=dynpv, dypsv=dypsv).items():
            setattr(self, k, v)
        z2 = lambda: np.zeros((ni + 2, nj + 2))
        for n in ("u", "v", "p", "pp", "du", "dv", "ae", "aw", "an", "as_", "ap", "su", "sp"):
            setattr(self, n, z2())
        self.den = np.full((ni + 2, nj + 2), densit)
        self.vis = np.full((ni + 2, nj + 2), viscos)
        u = self.u                                                         # control.f90:148-168
        u[2, W(2, jstep)] = uin
        jfin = jstp1
        for i in range(2, ni + 1):
            if i >= istp1:
                jfin = nj
            factor = (yv[jstp1] * rv[jstp1]) / (yv[jfin] * rv[jfin])
            u[i, W(2, jfin - 1)] = uin * factor
        J = W(2, njm1)                                                     # control.f90:197-203
        arden = 0.5 * (self.den[1, J] + self.den[2, J]) * r[J] * sns[J]
        self.flowin = float((arden * u[2, J]).sum())
        self.xmonin = float((arden * u[2, J] * u[2, J]).sum())
        self.res = [0.0, 0.0, 0.0]

    @staticmethod
    def _hybrid(c, d, sign):
        return np.maximum(np.abs(0.5 * c), d) + sign * 0.5 * c

    def _tmult_n(self):
        return self.viscos / (self.yv[self.nj] - self.y[self.nj - 1])

    def _tmult_w(self):
        return self.viscos / (self.x[self.istep] - self.xu[self.istep])

    def calcu(self):
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        u, v, p, den, vis = self.u, self.v, self.p, self.den, self.vis
        I, J = (lambda d=0: W(3, nim1, d)), (lambda d=0: W(2, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        a_n, a_s = row(self.rv, 1) * col(self.sewu), row(self.rv) * col(self.sewu)            # calc.f90:32-35
        a_ew = row(self.r) * row(self.sns)
        vol = row(self.r) * col(self.sewu) * row(self.sns)
        gn = 0.5 * (den[I(), J(1)] + den[I(), J()]) * v[I(), J(1)]                            # :39-46
        gnw = 0.5 * (den[I(-1), J()] + den[I(-1), J(1)]) * v[I(-1), J(1)]
        gs = 0.5 * (den[I(), J(-1)] + den[I(), J()]) * v[I(), J()]
        gsw = 0.5 * (den[I(-1), J()] + den[I(-1), J(-1)]) * v[I(-1), J()]
        ge = 0.5 * (den[I(1), J()] + den[I(), J()]) * u[I(1), J()]
        gp = 0.5 * (den[I(), J()] + den[I(-1), J()]) * u[I(), J()]
        gw = 0.5 * (den[I(-1), J()] + den[I(-2), J()]) * u[I(-1), J()]
        cn, cs = 0.5 * (gn + gnw) * a_n, 0.5 * (gs + gsw) * a_s                               # :48-51
        ce, cw = 0.5 * (ge + gp) * a_ew, 0.5 * (gp + gw) * a_ew
        visn = 0.25 * (vis[I(), J()] + vis[I(), J(1)] + vis[I(-1), J()] + vis[I(-1), J(1)])   # :54-59
        viss = 0.25 * (vis[I(), J()] + vis[I(), J(-1)] + vis[I(-1), J()] + vis[I(-1), J(-1)])
        dn, ds = visn * a_n / row(self.dynp), viss * a_s / row(self.dyps)
        de, dw = vis[I(), J()] * a_ew / col(self.dxepu), vis[I(-1), J()] * a_ew / col(self.dxpwu)
        cp = np.maximum(0.0, cn - cs + ce - cw)                                               # :63-66
        an, as_ = self._hybrid(cn, dn, -1), self._hybrid(cs, ds, +1)                          # :69-72
        ae, aw = self._hybrid(ce, de, -1), self._hybrid(cw, dw, +1)
        du = a_ew + 0.0 * vol
        su = cp * u[I(), J()] + du * (p[I(-1), J()] - p[I(), J()])                            # :76
        sp = -cp
        dudxp = (u[I(1), J()] - u[I(), J()]) / col(self.sew)                                  # :81-84
        dudxm = (u[I(), J()] - u[I(-1), J()]) / col(self.sew, -1)
        su = (vis[I(), J()] * dudxp - vis[I(-1), J()] * dudxm) / col(self.sewu) * vol + su
        dvdxp = row(self.rv, 1) * (v[I(), J(1)] - v[I(-1), J(1)]) / col(self.dxep)            # :86-94 (gamp = visn, gamm = viss)
        dvdxm = row(self.rv) * (v[I(), J()] - v[I(-1), J()]) / col(self.dxep)
        su = su + (visn * dvdxp - viss * dvdxm) / row(self.sns) / row(self.r) * vol
        for name, val in (("ae", ae), ("aw", aw), ("an", an), ("as_", as_), ("su", su), ("sp", sp), ("du", du)):
            getattr(self, name)[I(), J()] = val
        # modu (promod.f90:26-95); laminar runs: yplusn = 11 <= 11.63  ->  tmult = viscos / yp (TurbTeachRef: wall law)
        Iw = W(3, nim1)
        tmult = self._tmult_n()
        self.sp[Iw, njm1] -= tmult * self.sewu[Iw] * self.rv[nj]
        self.an[Iw, njm1] = 0.0
        self.aw[3, W(self.jstep + 1, njm1)] = 0.0
        u[1:ni + 1, 1] = u[1:ni + 1, 2]
        self.as_[1:ni + 1, 2] = 0.0
        Jo = W(2, njm1)
        arden = 0.5 * (den[nim1, Jo] + den[nim1 - 1, Jo]) * self.r[Jo] * self.sns[Jo]
        uinc = (self.flowin - float((arden * u[nim1, Jo]).sum())) / float(arden.sum())
        u[ni, Jo] = u[nim1, Jo] + uinc
        # chapter 3 (calc.f90:103-124)
        ae, aw, an, as_, su, sp = (getattr(self, n)[I(), J()] for n in ("ae", "aw", "an", "as_", "su", "sp"))
        ap = an + as_ + ae + aw - sp
        du = self.du[I(), J()] / ap
        rs = an * u[I(), J(1)] + as_ * u[I(), J(-1)] + ae * u[I(1), J()] + aw * u[I(-1), J()] - ap * u[I(), J()] + su
        self.res[0] = float(np.abs(rs).sum())
        ap = ap / self.urf[0]
        self.su[I(), J()] = su + (1.0 - self.urf[0]) * ap * u[I(), J()]
        self.ap[I(), J()] = ap
        self.du[I(), J()] = du * self.urf[0]
        for _ in range(self.nswp[0]):
            line_sweep(3, 2, ni, nj, u, self.ap, self.an, self.as_, self.ae, self.aw, self.su)

    def calcv(self):
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        u, v, p, den, vis = self.u, self.v, self.p, self.den, self.vis
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(3, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        a_n, a_s = row(self.rcv, 1) * col(self.sew), row(self.rcv) * col(self.sew)            # calc.f90:148-151
        a_ew = row(self.rv) * row(self.snsv)
        vol = row(self.rv) * col(self.sew) * row(self.snsv)
        gn = 0.5 * (den[I(), J(1)] + den[I(), J()]) * v[I(), J(1)]                            # :153-159
        gp = 0.5 * (den[I(), J()] + den[I(), J(-1)]) * v[I(), J()]
        gs = 0.5 * (den[I(), J(-1)] + den[I(), J(-2)]) * v[I(), J(-1)]
        ge = 0.5 * (den[I(1), J()] + den[I(), J()]) * u[I(1), J()]
        gse = 0.5 * (den[I(), J(-1)] + den[I(1), J(-1)]) * u[I(1), J(-1)]
        gw = 0.5 * (den[I(), J()] + den[I(-1), J()]) * u[I(), J()]
        gsw = 0.5 * (den[I(), J(-1)] + den[I(-1), J(-1)]) * u[I(), J(-1)]
        cn, cs = 0.5 * (gn + gp) * a_n, 0.5 * (gp + gs) * a_s
        ce, cw = 0.5 * (ge + gse) * a_ew, 0.5 * (gw + gsw) * a_ew
        vise = 0.25 * (vis[I(), J()] + vis[I(1), J()] + vis[I(), J(-1)] + vis[I(1), J(-1)])   # :165-170
        visw = 0.25 * (vis[I(), J()] + vis[I(-1), J()] + vis[I(), J(-1)] + vis[I(-1), J(-1)])
        dn, ds = vis[I(), J()] * a_n / row(self.dynpv), vis[I(), J(-1)] * a_s / row(self.dypsv)
        de, dw = vise * a_ew / col(self.dxep), visw * a_ew / col(self.dxpw)
        cp = np.maximum(0.0, cn - cs + ce - cw)
        an, as_ = self._hybrid(cn, dn, -1), self._hybrid(cs, ds, +1)
        ae, aw = self._hybrid(ce, de, -1), self._hybrid(cw, dw, +1)
        dv = 0.5 * (a_n + a_s)
        su = cp * v[I(), J()] + dv * (p[I(), J(-1)] - p[I(), J()])                            # :181
        sp = -cp
        if self.indcos == 2:                                                                  # :183-184: subtracted TWICE as written
            sink = vis[I(), J()] * vol / row(self.rv) ** 2
            sp = sp - sink
            sp = sp - sink
        dudyp = (u[I(1), J()] - u[I(1), J(-1)]) / row(self.dyps)                              # :185-189
        dudym = (u[I(), J()] - u[I(), J(-1)]) / row(self.dyps)
        su = su + (vise * dudyp - visw * dudym) / col(self.sew) * vol
        dvdyp = (v[I(), J(1)] - v[I(), J()]) / row(self.sns)                                  # :190-194
        dvdym = (v[I(), J()] - v[I(), J(-1)]) / row(self.sns, -1)
        su = su + (vis[I(), J()] * row(self.r) * dvdyp - vis[I(), J(-1)] * row(self.r, -1) * dvdym) / (row(self.r) * row(self.sns)) * vol
        for name, val in (("ae", ae), ("aw", aw), ("an", an), ("as_", as_), ("su", su), ("sp", sp), ("dv", dv)):
            getattr(self, name)[I(), J()] = val
        # modv (promod.f90:99-133); laminar runs: xplusw = 11 (0 at jstep) <= 11.63  ->  tmult = viscos / xp
        Jw = W(self.jstep + 1, njm1)
        tmult = self._tmult_w()
        self.sp[self.istep, Jw] -= tmult * self.snsv[Jw] * self.rv[Jw]
        self.aw[self.istep, Jw] = 0.0
        self.an[W(self.istep, nim1), njm1] = 0.0
        self.as_[W(2, nim1), 3] = 0.0
        ae, aw, an, as_, su, sp = (getattr(self, n)[I(), J()] for n in ("ae", "aw", "an", "as_", "su", "sp"))
        ap = an + as_ + ae + aw - sp                                                          # :207-219
        dv = self.dv[I(), J()] / ap
        rs = an * v[I(), J(1)] + as_ * v[I(), J(-1)] + ae * v[I(1), J()] + aw * v[I(-1), J()] - ap * v[I(), J()] + su
        self.res[1] = float(np.abs(rs).sum())
        ap = ap / self.urf[1]
        self.su[I(), J()] = su + (1.0 - self.urf[1]) * ap * v[I(), J()]
        self.ap[I(), J()] = ap
        self.dv[I(), J()] = dv * self.urf[1]
        for _ in range(self.nswp[1]):
            line_sweep(2, 3, ni, nj, v, self.ap, self.an, self.as_, self.ae, self.aw, self.su)

    def calcp(self):
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        u, v, p, den, du, dv, pp = self.u, self.v, self.p, self.den, self.du, self.dv, self.pp
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        a_n, a_s = row(self.rv, 1) * col(self.sew), row(self.rv) * col(self.sew)              # calc.f90:245-247
        a_ew = row(self.r) * row(self.sns) + 0.0 * a_n
        dnn, dns = 0.5 * (den[I(), J()] + den[I(), J(1)]), 0.5 * (den[I(), J()] + den[I(), J(-1)])
        dne, dnw = 0.5 * (den[I(), J()] + den[I(1), J()]), 0.5 * (den[I(), J()] + den[I(-1), J()])
        an, as_ = dnn * a_n * dv[I(), J(1)], dns * a_s * dv[I(), J()]                         # :253-256
        ae, aw = dne * a_ew * du[I(1), J()], dnw * a_ew * du[I(), J()]
        smp = dnn * v[I(), J(1)] * a_n - dns * v[I(), J()] * a_s + dne * u[I(1), J()] * a_ew - dnw * u[I(), J()] * a_ew
        self.res[2] = float(np.abs(smp).sum())
        for name, val in (("ae", ae), ("aw", aw), ("an", an), ("as_", as_), ("ap", an + as_ + ae + aw), ("su", -smp)):
            getattr(self, name)[I(), J()] = val
        for _ in range(self.nswp[2]):
            line_sweep(2, 2, ni, nj, pp, self.ap, self.an, self.as_, self.ae, self.aw, self.su)
        Iu, Jv = W(3, nim1), W(3, njm1)                                                       # :290-304: velocities first
        u[Iu, J()] += du[Iu, J()] * (pp[W(3, nim1, -1), J()] - pp[Iu, J()])
        v[I(), Jv] += dv[I(), Jv] * (pp[I(), W(3, njm1, -1)] - pp[I(), Jv])
        ppref = pp[2, 2]
        p[I(), J()] += self.urf[2] * (pp[I(), J()] - ppref)
        pp[I(), J()] = 0.0


class TurbTeachRef(TeachRef):
    """The program as shipped: k-epsilon + energy + props on top of the laminar core.  control.f90:95-179 (constants,
    initial fields), calc.f90:309-577 (calct, calcte, calced), promod.f90:36-60, 105-121 (wall law in modu / modv),
    :146-296 (modt, modte, moded), general.f90:165-190 (props); loop order control.f90:210-217."""

    def __init__(self, ni, nj, jstep, tin=0.0, twalln=100.0, turbin=0.03, alamda=0.005, prandl=7.0,
                 urf_turb=(0.7, 0.7, 1.0, 0.7), nswp_turb=(3, 3, 3), uin=0.5, rlarge=0.0124, **kw):
        super().__init__(ni, nj, jstep, uin=uin, rlarge=rlarge, **kw)
        nim1, njm1, jstp1 = ni - 1, nj - 1, jstep + 1
        self.cmu, self.cd, self.c1, self.c2, self.cappa, self.elog = 0.09, 1.00, 1.44, 1.92, .4187, 9.793
        self.pred = self.cappa * self.cappa / (self.c2 - self.c1) / (self.cmu ** .5)
        self.prte, self.prandt, self.prandl, self.twalln = 1.0, 0.9, prandl, twalln
        pf = prandl / self.prandt
        self.pfun = 9.24 * (pf ** 0.75 - 1.0) * (1.0 + 0.28 * np.exp(-0.007 * pf))
        self.urfk, self.urfe, self.urft, self.urfvis = urf_turb
        self.nswpk, self.nswpd, self.nswpt = nswp_turb
        z2 = lambda: np.zeros((ni + 2, nj + 2))
        for n in ("te", "ed", "t", "gen", "gamh", "sukd", "spkd"):
            setattr(self, n, z2())
        self.taun, self.yplusn = np.zeros(ni + 2), np.zeros(ni + 2)
        self.tauw, self.xplusw = np.zeros(nj + 2), np.zeros(nj + 2)
        tein = turbin * uin ** 2
        edin = tein ** 1.5 / (alamda * rlarge)
        self.t[1, W(2, jstep)] = tin                                       # control.f90:148-179
        self.te[1, W(2, jstep)] = tein
        self.ed[1, W(2, jstep)] = edin
        self.te[2, W(2, jstep)] = tein                                     # i = 2 < istp1: jfin = jstp1
        self.ed[2, W(2, jstep)] = edin
        self.te[W(3, ni), W(2, njm1)] = tein
        self.ed[W(3, ni), W(2, njm1)] = edin
        self.t[W(2, nim1), nj] = twalln
        self.yplusn[W(2, nim1)] = 11.0
        self.t[1, W(jstep, nj)] = tin
        self.xplusw[W(jstep, nj)] = 11.0
        self.xplusw[jstep] = 0.0
        self.res = [0.0] * 6
        self.props()

    # ---- wall law of modu / modv (promod.f90:36-60, 105-121)
    def _tmult_n(self):
        ni, nj = self.ni, self.nj
        I, j = W(3, ni - 1), nj - 1
        Im = W(3, ni - 1, -1)
        yp = self.yv[nj] - self.y[j]
        sqrtk = np.sqrt(0.5 * (self.te[I, j] + self.te[Im, j]))
        denu = 0.5 * (self.den[I, j] + self.den[Im, j])
        ypa = 0.5 * (self.yplusn[I] + self.yplusn[Im])
        with np.errstate(divide="ignore", invalid="ignore"):
            tm = np.where(ypa <= 11.63, self.viscos / yp, denu * self.cmu ** 0.25 * sqrtk * self.cappa / np.log(self.elog * ypa))
        self.taun[I] = -tm * self.u[I, j]
        self.taun[2], self.taun[ni] = self.taun[3], self.taun[ni - 1]
        return tm

    def _tmult_w(self):
        nj, i = self.nj, self.istep
        J, Jm = W(self.jstep + 1, nj - 1), W(self.jstep + 1, nj - 1, -1)
        xp = self.x[i] - self.xu[i]
        sqrtk = np.sqrt(0.5 * (self.te[i, J] + self.te[i, Jm]))
        denv = 0.5 * (self.den[i, J] + self.den[i, Jm])
        xpa = 0.5 * (self.xplusw[J] + self.xplusw[Jm])
        with np.errstate(divide="ignore", invalid="ignore"):
            tm = np.where(xpa <= 11.63, self.viscos / xp, denv * self.cmu ** 0.25 * sqrtk * self.cappa / np.log(self.elog * xpa))
        self.tauw[J] = -tm * self.v[i, J]
        self.tauw[self.jstep], self.tauw[nj] = self.tauw[self.jstep + 1], self.tauw[nj - 1]
        return tm

    # ---- chapter 1 shared by the three scalar equations: convection + diffusion with exchange coefficient `gam`
    def _scalar_coefficients(self, gam):
        nim1, njm1 = self.ni - 1, self.nj - 1
        u, v, den = self.u, self.v, self.den
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
        col, row = (lambda a, d=0: a[I(d)][:, None]), (lambda a, d=0: a[J(d)][None, :])
        a_n, a_s = row(self.rv, 1) * col(self.sew), row(self.rv) * col(self.sew)
        a_ew = row(self.r) * row(self.sns) + 0.0 * a_n
        vol = row(self.r) * row(self.sns) * col(self.sew)
        cn = 0.5 * (den[I(), J()] + den[I(), J(1)]) * v[I(), J(1)] * a_n
        cs = 0.5 * (den[I(), J()] + den[I(), J(-1)]) * v[I(), J()] * a_s
        ce = 0.5 * (den[I(), J()] + den[I(1), J()]) * u[I(1), J()] * a_ew
        cw = 0.5 * (den[I(), J()] + den[I(-1), J()]) * u[I(), J()] * a_ew
        dn = 0.5 * (gam[I(), J()] + gam[I(), J(1)]) * a_n / row(self.dynp)
        ds = 0.5 * (gam[I(), J()] + gam[I(), J(-1)]) * a_s / row(self.dyps)
        de = 0.5 * (gam[I(), J()] + gam[I(1), J()]) * a_ew / col(self.dxep)
        dw = 0.5 * (gam[I(), J()] + gam[I(-1), J()]) * a_ew / col(self.dxpw)
        cp = np.maximum(0.0, cn - cs + ce - cw)
        self.an[I(), J()], self.as_[I(), J()] = self._hybrid(cn, dn, -1), self._hybrid(cs, ds, +1)
        self.ae[I(), J()], self.aw[I(), J()] = self._hybrid(ce, de, -1), self._hybrid(cw, dw, +1)
        return I, J, col, row, vol, cp

    def _finish(self, phi, urf, nswp):
        """chapter 3 + 4 of calct / calcte / calced: ap, residual (normalised where sp holds a `great` source), relaxation."""
        ni, nj, nim1, njm1 = self.ni, self.nj, self.ni - 1, self.nj - 1
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
        an, as_, ae, aw, su, sp = (getattr(self, n)[I(), J()] for n in ("an", "as_", "ae", "aw", "su", "sp"))
        ap = an + as_ + ae + aw - sp
        rs = an * phi[I(), J(1)] + as_ * phi[I(), J(-1)] + ae * phi[I(1), J()] + aw * phi[I(-1), J()] - ap * phi[I(), J()] + su
        sorvol = self.great * (self.r[J()][None, :] * self.sew[I()][:, None] * self.sns[J()][None, :])
        rs = np.where(-sp > 0.5 * sorvol, rs / sorvol, rs)
        ap = ap / urf
        self.su[I(), J()] = su + (1.0 - urf) * ap * phi[I(), J()]
        self.ap[I(), J()] = ap
        for _ in range(nswp):
            line_sweep(2, 2, ni, nj, phi, self.ap, self.an, self.as_, self.ae, self.aw, self.su)
        return float(np.abs(rs).sum())

    def _dudy(self, i, j):           # cell-centre du/dy of promod.f90 / calc.f90 (i, j may be windows)
        u = self.u
        ip = slice(i.start + 1, i.stop + 1) if isinstance(i, slice) else i + 1
        jp = slice(j.start + 1, j.stop + 1) if isinstance(j, slice) else j + 1
        jm = slice(j.start - 1, j.stop - 1) if isinstance(j, slice) else j - 1
        return (u[i, j] + u[ip, j] + u[i, jp] + u[ip, jp]) / 4. - (u[i, j] + u[ip, j] + u[i, jm] + u[ip, jm]) / 4.

    def _dvdx(self, i, j):
        v = self.v
        ip = slice(i.start + 1, i.stop + 1) if isinstance(i, slice) else i + 1
        im = slice(i.start - 1, i.stop - 1) if isinstance(i, slice) else i - 1
        jp = slice(j.start + 1, j.stop + 1) if isinstance(j, slice) else j + 1
        return (v[i, j] + v[i, jp] + v[ip, j] + v[ip, jp]) / 4. - (v[i, j] + v[i, jp] + v[im, j] + v[im, jp]) / 4.

    def calcte(self):
        ni, nj, nim1, njm1, jstp1, istep = self.ni, self.nj, self.ni - 1, self.nj - 1, self.jstep + 1, self.istep
        u, v, den, vis, te, gen = self.u, self.v, self.den, self.vis, self.te, self.gen
        I, J, col, row, vol, cp = self._scalar_coefficients(vis / self.prte)
        dudx = (u[I(1), J()] - u[I(), J()]) / col(self.sew)                                    # calc.f90:430-436
        dvdy = (v[I(), J(1)] - v[I(), J()]) / row(self.sns)
        dudy = self._dudy(I(), J()) / row(self.sns)
        dvdx = self._dvdx(I(), J()) / col(self.sew)
        g = (2. * (dudx ** 2 + dvdy ** 2) + (dudy + dvdx) ** 2) * vis[I(), J()]
        if self.indcos == 2:                                                                   # :438-441
            with np.errstate(divide="ignore", invalid="ignore"):
                vdr = v[I(), J()] / row(self.rv)                                               # rv(2) = 0 on the axis ...
            vdr[:, 0] = 0.0                                                                    # ... j == 2: vdr = 0
            g = g + vis[I(), J()] * 0.5 * (vdr + v[I(), J(1)] / row(self.rv, 1)) ** 2
        gen[I(), J()] = g
        self.sukd[I(), J()] = cp * te[I(), J()]
        self.su[I(), J()] = self.sukd[I(), J()] + g * vol
        self.spkd[I(), J()] = -cp
        self.sp[I(), J()] = -cp - self.cd * self.cmu * den[I(), J()] ** 2 * te[I(), J()] * vol / vis[I(), J()]
        # ---- modte (promod.f90:182-256): top wall, side wall, symmetry axis -- in this order (they share corner cells)
        cdterm, cmu75 = self.cmu ** 0.25, self.cmu ** .75
        j, Iw = njm1, W(2, nim1)
        yp = self.yv[nj] - self.y[njm1]
        sqrtk = np.sqrt(te[Iw, j])
        vol_w = self.r[j] * self.sns[j] * self.sew[Iw]
        gencou = 0.5 * (np.abs(self.taun[W(2, nim1, 1)] * u[W(2, nim1, 1), j]) + np.abs(self.taun[Iw] * u[Iw, j])) / yp
        self.yplusn[Iw] = den[Iw, j] * sqrtk * cdterm * yp / self.viscos
        dudy_w = self._dudy(Iw, j) / self.sns[j]
        gen[Iw, j] = (gen[Iw, j] - vis[Iw, j] * dudy_w ** 2) + gencou
        ypl = self.yplusn[Iw]
        with np.errstate(divide="ignore", invalid="ignore"):
            diterm = np.where(ypl <= 11.63, den[Iw, j] * cmu75 * sqrtk * ypl / yp,
                              den[Iw, j] * cmu75 * sqrtk * np.log(self.elog * ypl) / (self.cappa * yp))
        self.su[Iw, j] = gen[Iw, j] * vol_w + self.sukd[Iw, j]
        self.sp[Iw, j] = -diterm * vol_w + self.spkd[Iw, j]
        self.an[Iw, j] = 0.0
        i, Jw = istep, W(jstp1, njm1)
        xp = self.x[istep] - self.xu[istep]
        sqrtk = np.sqrt(te[i, Jw])
        vol_s = self.r[Jw] * self.sns[Jw] * self.sew[i]
        self.xplusw[Jw] = den[i, Jw] * sqrtk * cdterm * xp / self.viscos
        gencou = 0.5 * (np.abs(self.tauw[W(jstp1, njm1, 1)] * v[i, W(jstp1, njm1, 1)]) + np.abs(self.tauw[Jw] * v[i, Jw])) / xp
        dvdx_s = self._dvdx(i, Jw) / self.sew[i]
        gen[i, Jw] = (gen[i, Jw] - vis[i, Jw] * dvdx_s ** 2) + gencou
        xpl = self.xplusw[Jw]
        with np.errstate(divide="ignore", invalid="ignore"):
            diterm = np.where(xpl <= 11.63, den[i, Jw] * cmu75 * sqrtk * xpl / xp,
                              den[i, Jw] * cmu75 * sqrtk * np.log(self.elog * xpl) / (self.cappa * xp))
        self.su[i, Jw] = self.sukd[i, Jw] + gen[i, Jw] * vol_s
        self.sp[i, Jw] = self.spkd[i, Jw] - diterm * vol_s
        self.aw[i, Jw] = 0.0
        j = 2
        te[Iw, 1] = te[Iw, 2]
        dudy_a = self._dudy(Iw, j) / self.sns[j]
        gen[Iw, j] = gen[Iw, j] - vis[Iw, j] * dudy_a ** 2
        self.su[Iw, j] = self.sukd[Iw, j] + gen[Iw, j] * (self.r[j] * self.sns[j] * self.sew[Iw])
        self.as_[Iw, 2] = 0.0
        self.res[4] = self._finish(te, self.urfk, self.nswpk)

    def calced(self):
        nj, nim1, njm1, jstp1, istep = self.nj, self.ni - 1, self.nj - 1, self.jstep + 1, self.istep
        den, vis, te, ed = self.den, self.vis, self.te, self.ed
        I, J, col, row, vol, cp = self._scalar_coefficients(vis / self.pred)
        self.su[I(), J()] = cp * ed[I(), J()] + self.c1 * self.cmu * self.gen[I(), J()] * vol * den[I(), J()] * te[I(), J()] / vis[I(), J()]
        with np.errstate(divide="ignore", invalid="ignore"):      # 0/0 behind the step (te = ed = 0): moded overwrites those cells
            self.sp[I(), J()] = -cp - self.c2 * den[I(), J()] * ed[I(), J()] * vol / te[I(), J()]
        cmu75, Iw = self.cmu ** .75, W(2, nim1)                                               # moded (promod.f90:260-296)
        term = cmu75 / (self.cappa * (self.yv[nj] - self.y[njm1]))
        self.su[Iw, njm1] = self.great * term * te[Iw, njm1] ** 1.5
        self.sp[Iw, njm1] = -self.great
        term = cmu75 / (self.cappa * (self.x[istep] - self.xu[istep]))
        Jw = W(jstp1, nj - 2)
        self.su[istep, Jw] = self.great * term * te[istep, Jw] ** 1.5
        self.sp[istep, Jw] = -self.great
        self.as_[Iw, 2] = 0.0
        self.res[5] = self._finish(ed, self.urfe, self.nswpd)

    def calct(self):
        nj, nim1, njm1, jstp1 = self.nj, self.ni - 1, self.nj - 1, self.jstep + 1
        I, J, col, row, vol, cp = self._scalar_coefficients(self.gamh)
        self.su[I(), J()] = cp * self.t[I(), J()]
        self.sp[I(), J()] = -cp
        j, Iw = njm1, W(2, nim1)                                                               # modt (promod.f90:146-178)
        dyn = self.yv[nj] - self.y[njm1]
        self.an[Iw, j] = 0.0
        area = self.rv[j + 1] * self.sew[Iw]
        ypl = self.yplusn[Iw]
        with np.errstate(divide="ignore", invalid="ignore"):
            uplus = np.log(self.elog * ypl) / self.cappa
            gt = np.where(ypl <= 11.63, self.viscos / (self.prandl * dyn),
                          self.den[Iw, j] * self.cmu ** 0.25 * np.sqrt(self.te[Iw, j]) / (self.prandt * (uplus + self.pfun)))
        term = gt * area
        self.su[Iw, j] += term * self.twalln
        self.sp[Iw, j] -= term
        self.aw[2, W(jstp1, njm1)] = 0.0
        self.as_[Iw, 2] = 0.0
        self.res[3] = self._finish(self.t, self.urft, self.nswpt)

    def props(self):                                                                           # general.f90:165-190
        nim1, njm1 = self.ni - 1, self.nj - 1
        I, J = W(2, nim1), W(2, njm1)
        vis, te, ed = self.vis, self.te, self.ed
        vis[I, 1] = vis[I, 2]
        old = vis[I, J].copy()
        with np.errstate(divide="ignore", invalid="ignore"):
            new = np.where(ed[I, J] == 0., self.viscos, self.den[I, J] * te[I, J] ** 2 * self.cmu / ed[I, J] + self.viscos)
        vis[I, J] = self.urfvis * new + (1. - self.urfvis) * old
        self.gamh[I, J] = self.viscos / self.prandl + (vis[I, J] - self.viscos) / self.prandt

    def iterate(self):                                  # control.f90:210-217
        self.calcu()
        self.calcv()
        self.calcp()
        self.calcte()
        self.calced()
        self.calct()
        self.props()


def gen_teach_turb():
    out = {"_source": "tools/gen_kat.py (numpy restatement of backward-step-flow-2d as shipped: k-epsilon + energy + props + wall "
                      "functions; FP64; independent of oracle/teach_turb.inc)"}
    for key, kw in (("ni40_nj30", dict(ni=40, nj=30, jstep=8, epsx=1.05, indcos=2)),
                    ("ni34_nj26_plane", dict(ni=34, nj=26, jstep=6, epsx=1.0, indcos=1))):
        t = TurbTeachRef(**kw)
        imon, jmon = 10, 10
        hist = {}
        for it in range(1, 61):
            t.iterate()
            if it in (1, 2, 10, 30, 60):
                ni, nj = t.ni, t.nj
                hist[str(it)] = {"res": list(t.res),
                                 "mon": {n: float(getattr(t, n)[imon, jmon]) for n in ("u", "v", "p", "te", "ed", "t", "vis")},
                                 "sum_abs": {n: float(np.abs(getattr(t, n)[1:ni + 1, 1:nj + 1]).sum()) for n in ("u", "v", "p", "te", "ed", "t", "vis", "gamh")},
                                 "taun_sum": float(np.abs(t.taun[1:ni + 1]).sum()), "yplusn_sum": float(t.yplusn[1:ni + 1].sum()),
                                 "tauw_sum": float(np.abs(t.tauw[1:nj + 1]).sum()), "xplusw_sum": float(t.xplusw[1:nj + 1].sum()),
                                 "yplusn_max": float(t.yplusn.max())}
        out[key] = dict(kw, imon=imon, jmon=jmon, history=hist)
    return out


def gen_teach():
    out = {"_source": "tools/gen_kat.py (numpy restatement of backward-step-flow-2d: control.f90, general.f90, calc.f90 "
                      "calcu/calcv/calcp, promod.f90 modu/modv; laminar, FP64; independent of oracle/teach_body.inc)"}
    for key, kw in (("ni40_nj30", dict(ni=40, nj=30, jstep=8, epsx=1.05, indcos=2)),
                    ("ni34_nj26_plane", dict(ni=34, nj=26, jstep=6, epsx=1.0, indcos=1))):
        t = TeachRef(**kw)
        imon, jmon = 10, 10
        hist = {}
        for it in range(1, 101):
            t.calcu()
            t.calcv()
            t.calcp()
            if it in (1, 2, 10, 100):
                ni, nj = t.ni, t.nj
                hist[str(it)] = {"res": list(t.res), "u_mon": float(t.u[imon, jmon]), "v_mon": float(t.v[imon, jmon]),
                                 "p_mon": float(t.p[imon, jmon]), "u_out": float(t.u[ni, jmon]),
                                 "sum_abs": [float(np.abs(a[1:ni + 1, 1:nj + 1]).sum()) for a in (t.u, t.v, t.p)]}
        out[key] = dict(kw, imon=imon, jmon=jmon, flowin=t.flowin, xmonin=t.xmonin, history=hist)
    return out



# =====================================================================================================================
# 4. lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90: explicit-Euler predictor, p' by Stone's SIP (sipsol
#    :24-87), correction, mirror-ghost boundary conditions, continuity norm (time loop :168-255).
#    SIP works on the system  ap*phi + aw*phiW + ae*phiE + as*phiS + an*phiN = su  (the program passes -aw .. -an).
# =====================================================================================================================
@numba.njit(cache=True, error_model="numpy")      # 0/0 -> NaN as in the Fortran program (a zero first residual)
def sip_solve(n1, n2, aw, ae, as_, an, ap, su, phi, alpha, tol, maxit):
    """sipsol on the index box 1..n1 x 1..n2 (interior 2..n1-1 x 2..n2-1).  Returns the iterations used."""
    lw = np.zeros((n1 + 2, n2 + 2)); ls = np.zeros((n1 + 2, n2 + 2)); lp = np.zeros((n1 + 2, n2 + 2))
    un = np.zeros((n1 + 2, n2 + 2)); ue = np.zeros((n1 + 2, n2 + 2)); rs = np.zeros((n1 + 2, n2 + 2))
    for j in range(2, n2):
        for i in range(2, n1):
            lw[i, j] = aw[i, j] / (1.0 + alpha * un[i - 1, j])
            ls[i, j] = as_[i, j] / (1.0 + alpha * ue[i, j - 1])
            p1 = alpha * lw[i, j] * un[i - 1, j]
            p2 = alpha * ls[i, j] * ue[i, j - 1]
            lp[i, j] = 1.0 / (ap[i, j] + p1 + p2 - lw[i, j] * ue[i - 1, j] - ls[i, j] * un[i, j - 1])
            un[i, j] = (an[i, j] - p1) * lp[i, j]
            ue[i, j] = (ae[i, j] - p2) * lp[i, j]
    res1 = 0.0
    used = 0
    for it in range(1, maxit + 1):
        used = it
        resl = 0.0
        for j in range(2, n2):
            for i in range(2, n1):
                r = su[i, j] - aw[i, j] * phi[i - 1, j] - ae[i, j] * phi[i + 1, j] - an[i, j] * phi[i, j + 1] \
                    - as_[i, j] * phi[i, j - 1] - ap[i, j] * phi[i, j]
                resl += abs(r)
                rs[i, j] = (r - ls[i, j] * rs[i, j - 1] - lw[i, j] * rs[i - 1, j]) * lp[i, j]
        if it == 1:
            res1 = resl
        rsm = resl / res1
        for j in range(n2 - 1, 1, -1):
            for i in range(n1 - 1, 1, -1):
                rs[i, j] = rs[i, j] - un[i, j] * rs[i, j + 1] - ue[i, j] * rs[i + 1, j]
                phi[i, j] += rs[i, j]
        if rsm < tol:
            break
        if it == 1 and res1 < 1.0e-10:
            break
    return used


def gen_fdcavity():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90; "
                      "independent of oracle/fdcavity_oracle.c)"}
    for key, ni, nj, re, dt, nsteps in (("n33", 33, 33, 10.0, 1.0e-4, 5), ("n21x13", 21, 13, 10.0, 1.0e-4, 5),
                                        ("n11_default", 11, 11, 10.0, 1.0e-4, 20)):
        utop, urfp = 1.0, 0.8
        dx, dy = 1.0 / (ni - 1), 1.0 / (nj - 1)
        Z = lambda a, b: np.zeros((a + 2, b + 2))
        u, v, p, pp = Z(ni, nj + 1), Z(ni + 1, nj), Z(ni + 1, nj + 1), Z(ni + 1, nj + 1)
        uo, vo, po = Z(ni, nj + 1), Z(ni + 1, nj), Z(ni + 1, nj + 1)
        uo[1:ni + 1, nj:nj + 2] = 1.0                                     # :162-165
        po[1:ni + 1 + 1, 1:nj + 1 + 1] = 1.0
        ae, aw, an, as_, ap, su = (Z(ni + 1, nj + 1) for _ in range(6))
        steps = []
        for itsp in range(1, nsteps + 1):
            I, J = (lambda d=0: W(2, ni - 1, d)), (lambda d=0: W(2, nj, d))            # u momentum :170-180
            t = -0.5 * (uo[I(1), J()] * uo[I(1), J()] - uo[I(-1), J()] * uo[I(-1), J()]) / dx \
                - 0.25 * ((uo[I(), J()] + uo[I(), J(1)]) * (vo[I(), J()] + vo[I(1), J()])
                          - (uo[I(), J()] + uo[I(), J(-1)]) * (vo[I(), J(-1)] + vo[I(1), J(-1)])) / dy \
                - (po[I(1), J()] - po[I(), J()]) / dx \
                + ((uo[I(-1), J()] - 2.0 * uo[I(), J()] + uo[I(1), J()]) / dx ** 2
                   + (uo[I(), J(-1)] - 2.0 * uo[I(), J()] + uo[I(), J(1)]) / dy ** 2) / re
            u[I(), J()] = uo[I(), J()] + dt * t
            I, J = (lambda d=0: W(2, ni, d)), (lambda d=0: W(2, nj - 1, d))            # v momentum :182-192
            t = -0.25 * ((uo[I(), J()] + uo[I(), J(1)]) * (vo[I(), J()] + vo[I(1), J()])
                         - (uo[I(-1), J()] + uo[I(-1), J(1)]) * (vo[I(), J()] + vo[I(-1), J()])) / dx \
                - 0.5 * (vo[I(), J(1)] * vo[I(), J(1)] - vo[I(), J(-1)] * vo[I(), J(-1)]) / dy \
                - (po[I(), J(1)] - po[I(), J()]) / dy \
                + ((vo[I(-1), J()] - 2.0 * vo[I(), J()] + vo[I(1), J()]) / dx ** 2
                   + (vo[I(), J(-1)] - 2.0 * vo[I(), J()] + vo[I(), J(1)]) / dy ** 2) / re
            v[I(), J()] = vo[I(), J()] + dt * t
            pp[:, :] = 0.0                                                               # :194-205
            I, J = (lambda d=0: W(2, ni, d)), (lambda d=0: W(2, nj, d))
            ae[I(), J()] = aw[I(), J()] = dt / dx ** 2
            an[I(), J()] = as_[I(), J()] = dt / dy ** 2
            su[I(), J()] = -(u[I(), J()] - u[I(-1), J()]) / dx - (v[I(), J()] - v[I(), J(-1)]) / dy
            ap[I(), J()] = 2 * (dt / dx ** 2 + dt / dy ** 2)
            nit = sip_solve(ni + 1, nj + 1, -aw, -ae, -as_, -an, ap, su, pp, 0.88, 1.0e-5, 600)   # :206
            p[I(), J()] += urfp * pp[I(), J()]                                           # :208-222
            Iu = W(2, ni - 1)
            u[Iu, J()] -= dt / dx * (pp[W(2, ni - 1, 1), J()] - pp[Iu, J()])
            Jv = W(2, nj - 1)
            v[I(), Jv] -= dt / dy * (pp[I(), W(2, nj - 1, 1)] - pp[I(), Jv])
            u[1:ni + 1, 1] = -u[1:ni + 1, 2]                                             # :224-237
            u[1:ni + 1, nj + 1] = 2.0 * utop - u[1:ni + 1, nj]
            u[1, 2:nj + 1] = 0.0
            u[ni, 2:nj + 1] = 0.0
            v[1, 1:nj + 1] = -v[2, 1:nj + 1]
            v[ni + 1, 1:nj + 1] = -v[ni, 1:nj + 1]
            v[2:ni + 1, 1] = 0.0
            v[2:ni + 1, nj] = 0.0
            p[1, 1:ni + 2] = p[2, 1:ni + 2]                                              # extent ni+1 as written (:234)
            p[ni + 1, 1:nj + 2] = p[ni, 1:nj + 2]
            p[2:ni + 1, nj + 1] = p[2:ni + 1, nj]
            p[2:ni + 1, 1] = p[2:ni + 1, 2]
            I, J = (lambda d=0: W(2, ni, d)), (lambda d=0: W(2, nj, d))                  # :239-247
            err = float(np.sqrt((((u[I(), J()] - u[I(-1), J()]) / dx + (v[I(), J()] - v[I(), J(-1)]) / dy) ** 2).sum()))
            steps.append({"sip_iters": int(nit), "error": err, "u_mon": float(u[ni // 2, nj // 2]),
                          "sum_abs": [float(np.abs(a[1:, 1:]).sum()) for a in (u[:ni + 1, :nj + 2], v[:ni + 2, :nj + 1], p[:ni + 2, :nj + 2])]})
            uo[:, :], vo[:, :], po[:, :] = u, v, p                                       # :252-254
        out[key] = {"ni": ni, "nj": nj, "re": re, "dt": dt, "steps": steps}
    return out


# =====================================================================================================================
# 4b. lid-driven-cavity-2d/fdm-implicit/fdm_implicit_cavity.f90: implicit-Euler (or steady) momentum equations linearised
#     about the previous velocities, each solved by SIP (sipsol :29-92: alpha 0.80, tol 1e-6, maxit 3000), p' from
#     continuity, correction, mirror-ghost boundary values, continuity norm (time loop :174-288).
# =====================================================================================================================
def gen_fdcavity_implicit():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of lid-driven-cavity-2d/fdm-implicit/fdm_implicit_cavity.f90; "
                      "independent of oracle/fdcavity_oracle.c)"}
    for key, ni, nj, re, dt, urf, steady, nsteps in (("n11_default", 11, 11, 10.0, 1.0e-4, (1.0, 1.0, 0.5), False, 12),
                                                     ("n25x17_relaxed", 25, 17, 100.0, 2.0e-3, (0.8, 0.7, 0.5), False, 8),
                                                     ("n17_steady", 17, 17, 10.0, 1.0e-4, (0.7, 0.7, 0.3), True, 8)):
        utop = 1.0
        urfu, urfv, urfp = urf
        dx, dy = 1.0 / (ni - 1), 1.0 / (nj - 1)
        Z = lambda: np.zeros((ni + 3, nj + 3))                     # every array (1:ni+1, 1:nj+1) as in the program (:128-133)
        u, v, p, pp, ae, aw, an, as_, ap, su = (Z() for _ in range(10))
        u[1:ni + 1, nj:nj + 2] = 1.0                                # :166-169
        rex, rey = 1.0 / re / dx ** 2, 1.0 / re / dy ** 2
        pterm = 1.0 / dt + 2.0 * (rex + rey)
        steps = []
        for itsp in range(1, nsteps + 1):
            I, J = (lambda d=0: W(2, ni - 1, d)), (lambda d=0: W(2, nj, d))             # u momentum :177-197
            vel1, vel2 = 0.5 * (v[I(), J()] + v[I(1), J()]), 0.5 * (v[I(), J(-1)] + v[I(1), J(-1)])
            ae[I(), J()] = -0.5 * u[I(1), J()] / dx + rex
            aw[I(), J()] = 0.5 * u[I(-1), J()] / dx + rex
            an[I(), J()] = -0.5 * vel1 / dy + rey
            as_[I(), J()] = 0.5 * vel2 / dy + rey
            gradp = (p[I(), J()] - p[I(1), J()]) / dx
            s0, a0 = (gradp, 2.0 * (rex + rey) + 0.0 * gradp) if steady else (u[I(), J()] / dt + gradp, pterm + 0.0 * gradp)
            su[I(), J()] = s0 + (1 - urfu) * a0 / urfu * u[I(), J()]
            ap[I(), J()] = a0 / urfu
            it_u = sip_solve(ni, nj + 1, -aw, -ae, -as_, -an, ap, su, u, 0.80, 1.0e-6, 3000)   # :198
            I, J = (lambda d=0: W(2, ni, d)), (lambda d=0: W(2, nj - 1, d))             # v momentum :200-219
            vel1, vel2 = 0.5 * (u[I(), J()] + u[I(), J(1)]), 0.5 * (u[I(-1), J()] + u[I(-1), J(1)])
            ae[I(), J()] = -0.5 * vel1 / dx + rex
            aw[I(), J()] = 0.5 * vel2 / dx + rex
            an[I(), J()] = -0.5 * v[I(), J(1)] / dy + rey
            as_[I(), J()] = -0.5 * v[I(), J(-1)] / dy + rey                              # the sign the source has (:208)
            gradp = (p[I(), J()] - p[I(), J(1)]) / dy
            s0, a0 = (gradp, 2.0 * (rex + rey) + 0.0 * gradp) if steady else (v[I(), J()] / dt + gradp, pterm + 0.0 * gradp)
            su[I(), J()] = s0 + (1 - urfv) * a0 / urfv * v[I(), J()]
            ap[I(), J()] = a0 / urfv
            it_v = sip_solve(ni + 1, nj, -aw, -ae, -as_, -an, ap, su, v, 0.80, 1.0e-6, 3000)   # :220
            pp[:, :] = 0.0                                                                # p' :222-240
            I, J = (lambda d=0: W(2, ni, d)), (lambda d=0: W(2, nj, d))
            den = 2.0 * (rex + rey) if steady else pterm
            ae[I(), J()] = aw[I(), J()] = 1.0 / den / dx ** 2
            an[I(), J()] = as_[I(), J()] = 1.0 / den / dy ** 2
            ap[I(), J()] = ae[I(), J()] + aw[I(), J()] + an[I(), J()] + as_[I(), J()]
            su[I(), J()] = (u[I(-1), J()] - u[I(), J()]) / dx + (v[I(), J(-1)] - v[I(), J()]) / dy
            it_p = sip_solve(ni + 1, nj + 1, -aw, -ae, -as_, -an, ap, su, pp, 0.80, 1.0e-6, 3000)
            p[I(), J()] += urfp * pp[I(), J()]                                            # :242-258 (pterm in both modes, as written)
            Iu, Jv = W(2, ni - 1), W(2, nj - 1)
            u[Iu, J()] -= (pp[W(2, ni - 1, 1), J()] - pp[Iu, J()]) / pterm / dx
            v[I(), Jv] -= (pp[I(), W(2, nj - 1, 1)] - pp[I(), Jv]) / pterm / dy
            u[1:ni + 1, 1] = -u[1:ni + 1, 2]                                              # :260-273
            u[1:ni + 1, nj + 1] = 2.0 * utop - u[1:ni + 1, nj]
            u[1, 2:nj + 1] = 0.0
            u[ni, 2:nj + 1] = 0.0
            v[1, 1:nj + 1] = -v[2, 1:nj + 1]
            v[ni + 1, 1:nj + 1] = -v[ni, 1:nj + 1]
            v[2:ni + 1, 1] = 0.0
            v[2:ni + 1, nj] = 0.0
            m = min(ni + 1, nj + 1)                                                       # `p(1,1:ni+1)` (:270): ni where nj is meant
            p[1, 1:m + 1] = p[2, 1:m + 1]
            p[ni + 1, 1:nj + 2] = p[ni, 1:nj + 2]
            p[2:ni + 1, nj + 1] = p[2:ni + 1, nj]
            p[2:ni + 1, 1] = p[2:ni + 1, 2]
            err = float(np.sqrt((((u[I(), J()] - u[I(-1), J()]) / dx + (v[I(), J()] - v[I(), J(-1)]) / dy) ** 2).sum()))
            steps.append({"sip_iters": [int(it_u), int(it_v), int(it_p)], "error": err, "u_mon": float(u[ni // 2, nj // 2]),
                          "sum_abs": [float(np.abs(a[1:ni + 2, 1:nj + 2]).sum()) for a in (u, v, p)]})
        out[key] = {"ni": ni, "nj": nj, "re": re, "dt": dt, "urf": list(urf), "steady": steady, "steps": steps}
    return out


# =====================================================================================================================
# 5. stagnation-point-flow-2d/steady/fvm2d_stag_point.f90: coefficients (:462-498), boundary edits (:500-515), the
#    iterative line solvers tdma_we / tdma_sn (:82-160) with wesweep / snsweep / tdma (:162-260), wall flux (:528-533).
# =====================================================================================================================
@numba.njit(cache=True)
def stag_sweep(direction, ni, nj, aw, ae, as_, an, ap, su, phi):
    """direction 0: wesweep (lines along j, W->E); 1: snsweep (lines along i, S->N).  tdma :232-260: two divisions."""
    n = nj if direction == 0 else ni
    P = np.zeros(n + 2); Q = np.zeros(n + 2)
    outer = ni if direction == 0 else nj
    for o in range(2, outer):
        P[1] = 0.0
        Q[1] = phi[o, 1] if direction == 0 else phi[1, o]
        for k in range(2, n):
            if direction == 0:
                a, b, c = ap[o, k], an[o, k], as_[o, k]
                d = su[o, k] + aw[o, k] * phi[o - 1, k] + ae[o, k] * phi[o + 1, k]
            else:
                a, b, c = ap[k, o], ae[k, o], aw[k, o]
                d = su[k, o] + as_[k, o] * phi[k, o - 1] + an[k, o] * phi[k, o + 1]
            P[k] = b / (a - c * P[k - 1])
            Q[k] = (d + c * Q[k - 1]) / (a - c * P[k - 1])
        for k in range(n - 1, 1, -1):
            if direction == 0:
                phi[o, k] = P[k] * phi[o, k + 1] + Q[k]
            else:
                phi[k, o] = P[k] * phi[k + 1, o] + Q[k]


def gen_stagnation():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of stagnation-point-flow-2d/steady/fvm2d_stag_point.f90; "
                      "independent of oracle/lin_oracle.c); den 1.2, gam 0.1, uniform grid, stop rule rsm < 1e-4"}
    for ncv in (20, 32):
        for isch in (1, 2):
            for isol in (1, 2):
                ni = nj = ncv + 2
                nim1, njm1 = ni - 1, nj - 1
                den, gam = 1.2, 0.1
                x, y, xc, yc, fx, fy = (np.zeros(ni + 2) for _ in range(6))
                dx = 1.0 / ncv
                x[1] = 0.0
                for i in range(2, nim1 + 1):
                    x[i] = x[i - 1] + dx
                x[ni] = x[nim1]
                xc[1] = x[1]
                xc[W(2, nim1)] = 0.5 * (x[W(2, nim1)] + x[W(2, nim1, -1)])
                xc[ni] = x[nim1]
                y[:], yc[:] = x, xc                                          # same construction in y (ymin = 0, ymax = 1)
                fx[W(2, nim1)] = (x[W(2, nim1)] - xc[W(2, nim1)]) / (xc[W(2, nim1, 1)] - xc[W(2, nim1)])
                fy[:] = fx
                u = np.repeat(x[:, None], nj + 2, axis=1)                   # u(i,j) = x(i), v(i,j) = -y(j)   (:443-448)
                v = -np.repeat(y[None, :], ni + 2, axis=0)
                phi = np.zeros((ni + 2, nj + 2))
                phi[1, 1:njm1 + 1] = 1.0 - yc[1:njm1 + 1]                   # :452
                I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
                hy = (y[J()] - y[J(-1)])[None, :]
                hx = (x[I()] - x[I(-1)])[:, None]
                de = -gam * hy / (xc[I(1)] - xc[I()])[:, None]              # :456-459
                dw = -gam * hy / (xc[I()] - xc[I(-1)])[:, None]
                dn = -gam * hx / (yc[J(1)] - yc[J()])[None, :]
                ds = -gam * hx / (yc[J()] - yc[J(-1)])[None, :]
                ge, gw = den * u[I(), J()] * hy, -den * u[I(-1), J()] * hy   # :466-469
                gn, gs = den * v[I(), J()] * hx, -den * v[I(), J(-1)] * hx
                if isch == 1:
                    ce, cw, cn, cs = (np.minimum(g, 0.0) for g in (ge, gw, gn, gs))
                else:
                    ce, cw = ge * fx[I()][:, None], gw * (1.0 - fx[I(-1)])[:, None]
                    cn, cs = gn * fy[J()][None, :], gs * (1.0 - fy[J(-1)])[None, :]
                A = {k: np.zeros((ni + 2, nj + 2)) for k in ("ae", "aw", "an", "as", "ap", "su")}
                A["ae"][I(), J()], A["aw"][I(), J()] = ce + de, cw + dw
                A["an"][I(), J()], A["as"][I(), J()] = cn + dn, cs + ds
                A["ap"][I(), J()] = A["ae"][I(), J()] + A["aw"][I(), J()] + A["an"][I(), J()] + A["as"][I(), J()]
                A["ap"][nim1, J()] -= A["ae"][nim1, J()]                   # :505-507
                A["ae"][nim1, J()] = 0.0
                A["ap"][I(), 2] -= A["as"][I(), 2]                         # :512-514
                A["as"][I(), 2] = 0.0
                iters, res1, rsm = 0, 0.0, 0.0
                for it in range(1, 1001):
                    iters = it
                    stag_sweep(isol - 1, ni, nj, A["aw"], A["ae"], A["as"], A["an"], A["ap"], A["su"], phi)
                    r = A["ae"][I(), J()] * phi[I(1), J()] + A["aw"][I(), J()] * phi[I(-1), J()] + A["an"][I(), J()] * phi[I(), J(1)] \
                        + A["as"][I(), J()] * phi[I(), J(-1)] + A["su"][I(), J()] - A["ap"][I(), J()] * phi[I(), J()]
                    res = float(np.abs(r).sum())
                    if it == 1:
                        res1 = res
                    rsm = res / res1
                    if rsm < 1.0e-4 or (it == 1 and res1 < 1.0e-10):
                        break
                phi[W(2, nim1), 1] = phi[W(2, nim1), 2]                    # :524-525
                phi[ni, 1:nj + 1] = phi[nim1, 1:nj + 1]
                fwall = float((gam * (y[J()] - y[J(-1)]) * (phi[2, J()] - phi[1, J()]) / (xc[2] - xc[1])).sum())
                out[f"cv{ncv}_isch{isch}_isol{isol}"] = {"ncv": ncv, "isch": isch, "isol": isol, "iters": iters, "rsm": rsm,
                                                         "fwall": fwall, "phi_mid": float(phi[ni // 2, nj // 2]),
                                                         "sum_abs_phi": float(np.abs(phi[1:ni + 1, 1:nj + 1]).sum())}
    return out


# =====================================================================================================================
# 5b. stagnation-point-flow-2d/unsteady/stag_point_unsteady.f90: the same transport problem marched in time (explicit
#     Euler, implicit Euler, Crank-Nicolson :634-688) with its own solver copies, which keep every term on the LEFT:
#     a x(k-1) + b x(k) + c x(k+1) = d  (tdma :232-261: x(k) = q(k) - p(k) x(k+1)),  residual ae*phiE + ... + ap*phi - su.
# 5c. diagonal-convection-2d/fvm2d_oblique_conv.f90: pure upwind convection of a step profile oblique to the grid,
#     the same left-hand-side convention, a tdma that starts from p(2) = c(2)/b(2) (boundary terms already in su) :346-387.
# =====================================================================================================================
@numba.njit(cache=True, error_model="numpy")
def lhs_sweep(direction, first_from_boundary, ni, nj, aw, ae, as_, an, ap, su, phi):
    """One wesweep (direction 0) / snsweep (1) with all neighbour terms on the left.  first_from_boundary: the recurrence starts
    from the boundary value x(1) (stag_point_unsteady.f90:245-249) or from the first interior equation alone (oblique :371-372)."""
    n = nj if direction == 0 else ni
    outer = ni if direction == 0 else nj
    p = np.zeros(n + 2); q = np.zeros(n + 2)
    for o in range(2, outer):
        p[1] = 0.0
        q[1] = (phi[o, 1] if direction == 0 else phi[1, o]) if first_from_boundary else 0.0
        for k in range(2, n):
            if direction == 0:
                a, b, c = as_[o, k], ap[o, k], an[o, k]
                d = su[o, k] - aw[o, k] * phi[o - 1, k] - ae[o, k] * phi[o + 1, k]
            else:
                a, b, c = aw[k, o], ap[k, o], ae[k, o]
                d = su[k, o] - as_[k, o] * phi[k, o - 1] - an[k, o] * phi[k, o + 1]
            if first_from_boundary or k > 2:
                den = b - a * p[k - 1]
                p[k] = c / den
                q[k] = (d - a * q[k - 1]) / den
            else:
                p[k] = c / b
                q[k] = d / b
        last = (phi[o, n] if direction == 0 else phi[n, o]) if first_from_boundary else 0.0
        x_next = last
        for k in range(n - 1, 1, -1):
            x = q[k] - p[k] * x_next if (first_from_boundary or k < n - 1) else q[k]
            if direction == 0:
                phi[o, k] = x
            else:
                phi[k, o] = x
            x_next = x


def lhs_residual(ni, nj, aw, ae, as_, an, ap, su, phi):
    I, J = (lambda d=0: W(2, ni - 1, d)), (lambda d=0: W(2, nj - 1, d))
    return float(np.abs(ae[I(), J()] * phi[I(1), J()] + aw[I(), J()] * phi[I(-1), J()] + an[I(), J()] * phi[I(), J(1)]
                        + as_[I(), J()] * phi[I(), J(-1)] + ap[I(), J()] * phi[I(), J()] - su[I(), J()]).sum())


def uniform_grid(ncv):
    """Faces x(1..ni) and centres xc of the demos' uniform unit grid, interpolation factors fx (:522-566)."""
    ni = ncv + 2
    nim1 = ni - 1
    x, xc, fx = (np.zeros(ni + 2) for _ in range(3))
    dx = 1.0 / ncv
    for i in range(2, nim1 + 1):
        x[i] = x[i - 1] + dx
    x[ni] = x[nim1]
    xc[W(2, nim1)] = 0.5 * (x[W(2, nim1)] + x[W(2, nim1, -1)])
    xc[1], xc[ni] = x[1], x[nim1]
    fx[W(2, nim1)] = (x[W(2, nim1)] - xc[W(2, nim1)]) / (xc[W(2, nim1, 1)] - xc[W(2, nim1)])
    return x, xc, fx


def gen_stag_unsteady():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of stagnation-point-flow-2d/unsteady/stag_point_unsteady.f90 "
                      ":583-700; independent of oracle/lin_oracle.c); den 1.2, gam 0.1, dt 1e-3, uniform 20 x 20 and 28 x 28 grids"}
    for ncv in (20, 28):
        for isch in (1, 2):
            for itsch in (1, 2, 3):
                for isol in (1, 2):
                    ni = nj = ncv + 2
                    nim1, njm1 = ni - 1, nj - 1
                    den, gam, dt = 1.2, 0.1, 1.0e-3
                    x, xc, fx = uniform_grid(ncv)
                    y, yc, fy = x, xc, fx
                    u = np.repeat(x[:, None], nj + 2, axis=1)
                    v = -np.repeat(y[None, :], ni + 2, axis=0)
                    phi = np.zeros((ni + 2, nj + 2))
                    phi[1, 1:njm1 + 1] = 1.0 - yc[1:njm1 + 1]
                    I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
                    hy, hx = (y[J()] - y[J(-1)])[None, :], (x[I()] - x[I(-1)])[:, None]
                    Z = lambda: np.zeros((ni + 2, nj + 2))
                    steps = []
                    for it in range(1, 9):
                        phio = phi.copy()
                        phi[1, 1:njm1 + 1] = 1.0 - yc[1:njm1 + 1]
                        ge, gw = den * u[I(), J()] * hy, -den * u[I(-1), J()] * hy
                        gn, gs = den * v[I(), J()] * hx, -den * v[I(), J(-1)] * hx
                        de = -gam * hy / (xc[I(1)] - xc[I()])[:, None]
                        dw = -gam * hy / (xc[I()] - xc[I(-1)])[:, None]
                        dn = -gam * hx / (yc[J(1)] - yc[J()])[None, :]
                        ds = -gam * hx / (yc[J()] - yc[J(-1)])[None, :]
                        if isch == 1:
                            ce, cw, cn, cs = (np.minimum(g, 0.0) for g in (ge, gw, gn, gs))
                        else:
                            ce, cw = ge * fx[I()][:, None], gw * (1.0 - fx[I(-1)])[:, None]
                            cn, cs = gn * fy[J()][None, :], gs * (1.0 - fy[J(-1)])[None, :]
                        ae, aw, an, as_, ap, su = (Z() for _ in range(6))
                        ae[I(), J()], aw[I(), J()], an[I(), J()], as_[I(), J()] = ce + de, cw + dw, cn + dn, cs + ds
                        ae[nim1, J()] = 0.0                                  # outflow (:624)
                        as_[I(), 2] = 0.0                                    # symmetry (:626)
                        ct = den / dt * hx * hy
                        nb = lambda: (ae[I(), J()] * phio[I(1), J()] + aw[I(), J()] * phio[I(-1), J()]
                                      + an[I(), J()] * phio[I(), J(1)] + as_[I(), J()] * phio[I(), J(-1)])
                        tot = lambda: ae[I(), J()] + aw[I(), J()] + an[I(), J()] + as_[I(), J()]
                        if itsch == 1:                                       # :634-651
                            su[I(), J()] = ct * phio[I(), J()] - nb() + tot() * phio[I(), J()]
                            ap[I(), J()] = ct + 0.0 * ge
                            for a in (ae, aw, an, as_):
                                a[I(), J()] = 0.0
                        elif itsch == 2:                                     # :653-662
                            su[I(), J()] = ct * phio[I(), J()]
                            ap[I(), J()] = ct - tot()
                        else:                                                # :664-682
                            for a in (ae, aw, an, as_):
                                a[I(), J()] = 0.5 * a[I(), J()]
                            su[I(), J()] = ct * phio[I(), J()] - nb() + tot() * phio[I(), J()]
                            ap[I(), J()] = ct - tot()
                        iters, res1, rsm = 0, 0.0, 0.0
                        for k in range(1, 1001):                             # tdma_we / tdma_sn :79-160
                            iters = k
                            lhs_sweep(isol - 1, True, ni, nj, aw, ae, as_, an, ap, su, phi)
                            res = lhs_residual(ni, nj, aw, ae, as_, an, ap, su, phi)
                            if k == 1:
                                res1 = res
                            with np.errstate(divide="ignore", invalid="ignore"):
                                rsm = float(np.float64(res) / np.float64(res1))
                            if rsm < 1.0e-4 or (k == 1 and res1 < 1.0e-10):
                                break
                        steps.append({"iters": iters, "rsm": rsm, "phi_mid": float(phi[ni // 2, nj // 2]),
                                      "sum_abs_phi": float(np.abs(phi[1:ni + 1, 1:nj + 1]).sum())})
                    fwall = float((gam * (y[J()] - y[J(-1)]) * (phi[2, J()] - phi[1, J()]) / (xc[2] - xc[1])).sum())
                    out[f"cv{ncv}_isch{isch}_itsch{itsch}_isol{isol}"] = {"ncv": ncv, "isch": isch, "itsch": itsch, "isol": isol,
                                                                          "steps": steps, "fwall": fwall}
    return out


def gen_oblique():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of diagonal-convection-2d/fvm2d_oblique_conv.f90; independent of "
                      "oracle/lin_oracle.c); UDS, uniform grids, tol 1e-4, maxit 100"}
    for nicv, njcv, vel, bc in ((20, 20, 2.0, 1), (20, 20, 2.0, 2), (32, 24, 1.5, 1), (24, 32, 0.7, 2)):
        ni, nj = nicv + 2, njcv + 2
        nim1, njm1 = ni - 1, nj - 1
        x, _, _ = uniform_grid(nicv)
        y, _, _ = uniform_grid(njcv)
        phi = np.zeros((ni + 2, nj + 2))
        phi[1, W(2, njm1)] = 1.0                                            # :168-177
        phi[1, 1] = 0.5
        if bc == 2:
            phi[W(1, nim1), nj] = 1.0
            phi[ni, W(1, njm1)] = 0.0
            phi[ni, nj] = 0.5
        I, J = (lambda d=0: W(2, nim1, d)), (lambda d=0: W(2, njm1, d))
        hy, hx = (y[J()] - y[J(-1)])[None, :], (x[I()] - x[I(-1)])[:, None]
        Z = lambda: np.zeros((ni + 2, nj + 2))
        ae, aw, an, as_, ap, su = (Z() for _ in range(6))
        ae[I(), J()] = np.minimum(vel * hy + 0.0 * hx, 0.0)                  # :180-201
        aw[I(), J()] = np.minimum(-vel * hy + 0.0 * hx, 0.0)
        an[I(), J()] = np.minimum(vel * hx + 0.0 * hy, 0.0)
        as_[I(), J()] = np.minimum(-vel * hx + 0.0 * hy, 0.0)
        ap[I(), J()] = -(ae[I(), J()] + aw[I(), J()] + an[I(), J()] + as_[I(), J()])
        su[2, J()] -= aw[2, J()] * phi[1, J()]                              # west, Dirichlet :203-208
        aw[2, J()] = 0.0
        if bc == 1:                                                         # east :210-219
            ap[nim1, J()] = ae[nim1, J()] + ap[nim1, J()]
        else:
            su[nim1, J()] -= ae[nim1, J()] * phi[ni, J()]
        ae[nim1, J()] = 0.0
        if bc == 1:                                                         # north :221-230
            ap[I(), njm1] = an[I(), njm1] + ap[I(), njm1]
        else:
            su[I(), njm1] -= an[I(), njm1] * phi[I(), nj]
        an[I(), njm1] = 0.0
        su[I(), 2] -= as_[I(), 2] * phi[I(), 1]                             # south, Dirichlet :232-237
        as_[I(), 2] = 0.0
        iters, res1, rsm = 0, 0.0, 0.0
        for k in range(1, 101):                                             # tdma_solver :284-315
            iters = k
            lhs_sweep(0, False, ni, nj, aw, ae, as_, an, ap, su, phi)
            res = lhs_residual(ni, nj, aw, ae, as_, an, ap, su, phi)
            if k == 1:
                res1 = res
            rsm = res / res1
            if res < 1.0e-10 or rsm < 1.0e-4:
                break
        if bc == 1:
            phi[ni, 1:nj + 1] = phi[nim1, 1:nj + 1]
            phi[1:ni + 1, nj] = phi[1:ni + 1, njm1]
        out[f"cv{nicv}x{njcv}_bc{bc}"] = {"nicv": nicv, "njcv": njcv, "vel": vel, "bc": bc, "iters": iters, "rsm": rsm,
                                         "phi_mid": float(phi[ni // 2, nj // 2]), "phi_row": [float(z) for z in phi[1:ni + 1, nj // 2]],
                                         "sum_abs_phi": float(np.abs(phi[1:ni + 1, 1:nj + 1]).sum())}
    return out


# =====================================================================================================================
# 6. finite-difference-method/heat-conduction/transient-conduction-1d: the explicit (explicit-time/
#    fdm1d_explicit_heat_cond.f90:88-141) and implicit (implicit-time/fdm1d_explicit_heat_cond.f90:76-184) programs as
#    shipped -- nx = 21, k = 5, rho = cp = 1, QDOT = 1e3 where x < 0.1, endtime = 10 -- and two shorter variants.  The time
#    loop `do while (time <= endtime)` counts with a floating-point sum, so the step count is part of the answer.
# =====================================================================================================================
@numba.njit(cache=True)
def thomas_pq(a, b, c, d, x):
    """thomas (:147-184) for a(i) x(i-1) + b(i) x(i) + c(i) x(i+1) = d(i), i = 1..n (0-based here), in P-Q form:
    x(i) = P(i) x(i+1) + Q(i).  The reference eliminates with e = c/temp, f = (d - a f_prev)/temp: the same numbers."""
    n = len(b)
    P = np.zeros(n); Q = np.zeros(n)
    P[0] = -c[0] / b[0]
    Q[0] = d[0] / b[0]
    for i in range(1, n):
        den = b[i] + a[i] * P[i - 1]
        P[i] = -c[i] / den
        Q[i] = (d[i] - a[i] * Q[i - 1]) / den
    x[n - 1] = Q[n - 1]
    for i in range(n - 2, -1, -1):
        x[i] = P[i] * x[i + 1] + Q[i]


def gen_heat1d():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of the two transient-conduction-1d programs; independent of "
                      "oracle/heat_oracle.c)"}
    for key, nx, endtime, cfl, dt_imp in (("shipped", 21, 10.0, 0.5, 2.0e-4), ("nx41_short", 41, 0.05, 0.4, 1.0e-4),
                                          ("nx33_t1", 33, 1.0, 0.25, 5.0e-4)):
        k, rho, cp, xa, xb, ta, tb = 5.0, 1.0, 1.0, 0.0, 1.0, 0.0, 0.0
        dx = (xb - xa) / float(nx - 1)
        x = xa + np.arange(nx) * dx
        q = np.where(x < 0.1, 1.0e3, 0.0)
        # ---- explicit: t(i) = tcopy(i) + (k/rho/cp*(tcopy(i+1) - 2 tcopy(i) + tcopy(i-1))/dx**2 + QDOT/rho/cp)*dt  (:130-139)
        dt = cfl * dx ** 2 / (k / rho / cp)
        tc = np.zeros(nx)
        t = tc.copy()
        time_, n_exp = 0.0, 0
        while time_ <= endtime:
            t[0], t[-1] = ta, tb
            t[1:-1] = tc[1:-1] + (k / rho / cp * (tc[2:] - 2 * tc[1:-1] + tc[:-2]) / dx ** 2 + q[1:-1] / rho / cp) * dt
            tc[:] = t
            time_ += dt
            n_exp += 1
        t_exp = t.copy()
        # ---- implicit: (1 + 2 fo) t(i) - fo t(i-1) - fo t(i+1) = tcopy(i) + dt QDOT/rho/cp, Dirichlet rows  (:118-141)
        fo = k * dt_imp / rho / cp / dx ** 2
        a = np.full(nx, -fo); b = np.full(nx, 1.0 + 2.0 * fo); c = np.full(nx, -fo)
        a[0] = c[0] = a[-1] = c[-1] = 0.0
        b[0] = b[-1] = 1.0
        tc = np.zeros(nx)
        t = tc.copy()
        time_, n_imp = 0.0, 0
        while time_ <= endtime:
            rhs = tc + dt_imp * q / rho / cp
            rhs[0], rhs[-1] = ta, tb
            thomas_pq(a, b, c, rhs, t)
            tc[:] = t
            time_ += dt_imp
            n_imp += 1
        out[key] = {"nx": nx, "endtime": endtime, "dt_explicit": dt, "dt_implicit": dt_imp, "steps_explicit": n_exp,
                    "steps_implicit": n_imp, "t_explicit": [float(z) for z in t_exp], "t_implicit": [float(z) for z in t]}
    return out


# =====================================================================================================================
# 7. finite-difference-method/heat-conduction/steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90 as written: the DENSE
#    (nx*ny)^2 matrix of interior() (:131-191) and boundary() (:193-246), then gauss_seidel (:248-287) over full rows,
#    zeros included.  The oracle and the device never build the matrix; this restatement does, exactly as the program.
# =====================================================================================================================
@numba.njit(cache=True)
def dense_gauss_seidel(a, b, x0, imax, tol, rf):
    n = len(b)
    x1 = x0.copy()
    x2 = x1.copy()
    passes = 0
    dx2 = 0.0
    for k in range(1, imax + 1):
        passes = k
        for i in range(n):
            s = 0.0
            for j in range(n):
                if j < i:
                    s = s + a[i, j] * x2[j]
                elif j > i:
                    s = s + a[i, j] * x1[j]
            x2[i] = (b[i] - s) * rf / a[i, i] + (1 - rf) * x1[i]
        dx2 = 0.0
        for i in range(n):
            dx2 = dx2 + (x1[i] - x2[i]) ** 2
        dx2 = np.sqrt(dx2)
        if dx2 < tol:
            break
        x1[:] = x2
    return x2, passes, dx2


def gen_heat2d_steady():
    out = {"_source": "tools/gen_kat.py (numpy/numba restatement of steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90 with "
                      "its dense matrix; independent of oracle/heat_oracle.c)"}
    cases = (("shipped_7x5", 7, 5, "const"), ("variable_13x9", 13, 9, "var"), ("variable_21x17", 21, 17, "var"))
    for key, nx, ny, kind in cases:
        xmin, xmax, ymin, ymax = 0.0, 0.03, 0.0, 0.02
        tleft, tright, tbottom, ttop = (300.0, 300.0, 300.0, 300.0) if kind == "const" else (300.0, 320.0, 310.0, 290.0)
        lam = (lambda x, y: 20.0) if kind == "const" else (lambda x, y: 20.0 * (1.0 + 10.0 * x + 5.0 * y))
        qd = (lambda x, y: 5.0e7) if kind == "const" else (lambda x, y: 5.0e7 * (1.0 + 20.0 * x))
        dx, dy = (xmax - xmin) / (nx - 1), (ymax - ymin) / (ny - 1)            # xymesh :107-129
        x = np.zeros(nx + 1); y = np.zeros(ny + 1)
        x[1] = xmin
        for i in range(2, nx + 1):
            x[i] = x[1] + (i - 1) * dx
        y[1] = ymin
        for j in range(2, ny + 1):
            y[j] = y[1] + (j - 1) * dy
        n = nx * ny
        a = np.zeros((n + 1, n + 1)); rhs = np.zeros(n + 1)
        ddx, ddy = x[2] - x[1], y[2] - y[1]
        for i in range(2, nx):
            for j in range(2, ny):
                k = (j - 1) * nx + i
                aw = lam(0.5 * (x[i] + x[i - 1]), y[j]) / ddx ** 2
                ae = lam(0.5 * (x[i] + x[i + 1]), y[j]) / ddx ** 2
                as_ = lam(x[i], 0.5 * (y[j] + y[j - 1])) / ddy ** 2
                an = lam(x[i], 0.5 * (y[j] + y[j + 1])) / ddy ** 2
                a[k, k - 1], a[k, k + 1], a[k, k - nx], a[k, k + nx] = aw, ae, as_, an
                a[k, k] = -(aw + ae + as_ + an)
                rhs[k] = -qd(x[i], y[j])
        for j in range(2, ny):                                                  # boundary :209-222
            for i, tv in ((1, tleft), (nx, tright)):
                k = (j - 1) * nx + i
                a[k, k], rhs[k] = 1.0, tv
        for j, tv in ((1, tbottom), (ny, ttop)):                                # :224-237 (rows include the corners)
            for i in range(1, nx + 1):
                k = (j - 1) * nx + i
                a[k, k], rhs[k] = 1.0, tv
        t, passes, dx2 = dense_gauss_seidel(a[1:, 1:].copy(), rhs[1:].copy(), np.zeros(n), 500, 1.0e-7, 0.5)
        out[key] = {"nx": nx, "ny": ny, "kind": kind, "tbc": [tleft, tright, tbottom, ttop], "passes": int(passes), "dx2": float(dx2),
                    "t": [float(z) for z in t], "t_centre": float(t[(ny // 2) * nx + nx // 2])}
    return out


GENERATORS = {"cavity_regen": gen_cavity, "cylinder_regen": gen_cylinder, "buoyancy": gen_buoyancy, "teach": gen_teach, "teach_turb": gen_teach_turb, "fdcavity": gen_fdcavity, "fdcavity_implicit": gen_fdcavity_implicit, "stagnation": gen_stagnation, "stag_unsteady": gen_stag_unsteady, "oblique": gen_oblique, "heat1d": gen_heat1d, "heat2d_steady": gen_heat2d_steady}



def main():
    which = sys.argv[1:] or list(GENERATORS)
    for name in which:
        t0 = time.time()
        data = GENERATORS[name]()
        path = os.path.join(GOLD, f"{name}_kat.json")
        with open(path, "w") as f:
            json.dump(data, f, indent=1)
        print(f"{path}: {time.time() - t0:.1f} s")


if __name__ == "__main__":
    main()
