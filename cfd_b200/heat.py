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

    def synchronize(self):
        check(self.lib.cfd_heat_synchronize(self._h))

    def device_view(self):
        base, stride, n = C.c_void_p(), C.c_longlong(), C.c_int()
        check(self.lib.cfd_heat_device_view(self._h, C.byref(base), C.byref(stride), C.byref(n)))
        return base.value, stride.value, n.value
