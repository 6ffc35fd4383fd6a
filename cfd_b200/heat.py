"""Host-side mirror of the heat-conduction programs (finite-difference-method/heat-conduction/) over the C ABI.

`Heat2D.step(n)` is n passes of the update loop of fd1d_explicit_heat_cond
(transient-conduction-1d/explicit-time/fdm1d_explicit_heat_cond.f90:130-139) with the 5-point operator of
steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90:171-188; the argument names follow that subroutine
(k, rho, cp, xa, xb, ta/tb -> tleft/tright/tbottom/ttop, dt).  All arithmetic runs in libcfd_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import CfdHeatConfig, check

QTHR_REF = float(np.float32(0.1))   # QDOT compares x with the default-real literal 0.1 (…explicit…f90:83)


class Heat2D:
    def __init__(self, nx, ny, *, k=5.0, rho=1.0, cp=1.0, xa=0.0, xb=1.0, ya=0.0, yb=1.0,
                 tleft=0.0, tright=0.0, tbottom=0.0, ttop=0.0, dt=None, fo=0.2,
                 qthr=QTHR_REF, qval=1.0e3, device=-1, j_goff=0, ny_global=0, halo_s=0, halo_n=0):
        self.lib = _lib.load()
        self.nx, self.ny = nx, ny
        nyg = ny_global if ny_global else ny
        dx = (xb - xa) / (nx - 1)
        if dt is None:
            dt = fo * dx * dx / (k / rho / cp)      # grid Fourier number; 2-D stability needs Fo <= 0.25
        self.dt = dt
        cfg = CfdHeatConfig(nx=nx, ny=ny, k=k, rho=rho, cp=cp, xa=xa, xb=xb, ya=ya, yb=yb,
                            tleft=tleft, tright=tright, tbottom=tbottom, ttop=ttop, dt=dt,
                            qthr=qthr, qval=qval, device=device, j_goff=j_goff, ny_global=nyg if ny_global else 0,
                            halo_s=halo_s, halo_n=halo_n)
        self._cfg = cfg
        self._h = C.c_void_p()
        check(self.lib.cfd_heat_create(C.byref(self._h), C.byref(cfg)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.cfd_heat_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int):
        check(self.lib.cfd_heat_set_stream(self._h, C.c_void_p(cuda_stream)))

    def upload(self, t: np.ndarray):
        t = np.asfortranarray(t, dtype=np.float64)
        assert t.shape == (self.nx, self.ny)
        check(self.lib.cfd_heat_upload(self._h, t.ctypes.data_as(C.c_void_p), self.nx))

    def upload_ptr(self, ptr: int, ld: int):
        check(self.lib.cfd_heat_upload(self._h, C.c_void_p(ptr), ld))

    def download_ptr(self, ptr: int, ld: int):
        check(self.lib.cfd_heat_download(self._h, C.c_void_p(ptr), ld))

    def download(self) -> np.ndarray:
        t = np.zeros((self.nx, self.ny), order="F")
        check(self.lib.cfd_heat_download(self._h, t.ctypes.data_as(C.c_void_p), self.nx))
        return t

    def step(self, nsteps: int = 1, exact: bool = False):
        fn = self.lib.cfd_heat2d_step_exact if exact else self.lib.cfd_heat2d_step
        check(fn(self._h, nsteps))

    def implicit_step(self, nsteps: int = 1, exact: bool = False):
        fn = self.lib.cfd_heat2d_implicit_step_exact if exact else self.lib.cfd_heat2d_implicit_step
        check(fn(self._h, nsteps))

    # -- variable conductivity: interior() / gauss_seidel of steady-heat-conduction-2d/fdm2d_steady_heat_cond.f90 ----
    def face_points(self):
        """The points interior() evaluates LAMBDA at (:171-174), as four pairs of (nx,ny) arrays: (x-dx/2, y),
        (x+dx/2, y), (x, y-dy/2), (x, y+dy/2), plus the node coordinates; xymesh (:112-129) for x, y."""
        g, nx, ny = self._cfg, self.nx, self.ny
        dx, dy = (g.xb - g.xa) / (nx - 1), (g.yb - g.ya) / (ny - 1)
        x, y = g.xa + np.arange(nx) * dx, g.ya + np.arange(ny) * dy
        x[0], y[0] = g.xa, g.ya
        X, Y = x[:, None] + 0.0 * y[None, :], 0.0 * x[:, None] + y[None, :]
        xm, xp, ym, yp = (np.zeros_like(X) for _ in range(4))
        xm[1:, :] = 0.5 * (x[1:] + x[:-1])[:, None]
        xp[:-1, :] = 0.5 * (x[:-1] + x[1:])[:, None]
        ym[:, 1:] = 0.5 * (y[1:] + y[:-1])[None, :]
        yp[:, :-1] = 0.5 * (y[:-1] + y[1:])[None, :]
        return (xm, Y), (xp, Y), (X, ym), (X, yp), (X, Y)

    def set_operator(self, lam, qdot=None):
        """`lam(x, y)`, `qdot(x, y)`: the program's functions LAMBDA and QDOT as vectorised callables (host logic,
        evaluated once); or pass four ready-made face arrays as a tuple for `lam` and an array for `qdot`."""
        f = lambda a: np.asfortranarray(np.broadcast_to(np.asarray(a, dtype=np.float64), (self.nx, self.ny)).copy())
        if callable(lam):
            w, e, s, n, c = self.face_points()
            faces = [f(lam(*w)), f(lam(*e)), f(lam(*s)), f(lam(*n))]
            q = f(qdot(*c)) if qdot is not None else None
        else:
            faces = [f(a) for a in lam]
            q = f(qdot) if qdot is not None else None
        ptr = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
        check(self.lib.cfd_heat_set_operator(self._h, *(ptr(a) for a in faces), ptr(q), self.nx))

    def apply(self) -> np.ndarray:
        r = np.zeros((self.nx, self.ny), order="F")
        check(self.lib.cfd_heat2d_apply(self._h, r.ctypes.data_as(C.c_void_p), self.nx))
        return r

    def step_var(self, nsteps: int = 1):
        check(self.lib.cfd_heat2d_step_var(self._h, nsteps))

    def steady_gs(self, maxit=500, tol=1.0e-7, rf=0.5):
        """gauss_seidel (:248-287), matrix-free, two colours; returns (passes, last |x1-x2|_2)."""
        it, d = C.c_int(), C.c_double()
        check(self.lib.cfd_heat2d_steady_gs(self._h, maxit, tol, rf, C.byref(it), C.byref(d)))
        return it.value, d.value

    def synchronize(self):
        check(self.lib.cfd_heat_synchronize(self._h))

    def device_view(self):
        base, stride, n = C.c_void_p(), C.c_longlong(), C.c_int()
        check(self.lib.cfd_heat_device_view(self._h, C.byref(base), C.byref(stride), C.byref(n)))
        return base.value, stride.value, n.value
