"""Host-side mirror of the explicit finite-difference cavity program
(lid-driven-cavity-2d/fdm-explicit/fdm_explicit_cavity.f90:90-255) over the C ABI.  The stencil passes of a time
step run on the GPU; the p' Poisson solve between them (Stone's SIP, `sipsol` :24-87) is the caller's -- pass a
`solve_pp(su) -> pp` callable to `step` (SURVEY 8f-1 marks a device SIP as "next")."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import CfdFdcavConfig, FDC_FIELDS, check


class FDCavity:
    def __init__(self, ni=11, nj=11, *, re=10.0, dt=1.0e-4, utop=1.0, urfp=0.8, device=-1, implicit=False,
                 urfu=1.0, urfv=1.0, steady=False):
        self.lib = _lib.load()
        self.ni, self.nj, self.implicit = ni, nj, implicit
        self._h = C.c_void_p()
        cfg = CfdFdcavConfig(ni=ni, nj=nj, re=re, dt=dt, utop=utop, urfp=urfp, device=device, steady=int(steady),
                             urfu=urfu, urfv=urfv)
        check(self.lib.cfd_fdcav_create(C.byref(self._h), C.byref(cfg)))
        uo = np.zeros((ni + 1, nj + 1), order="F")
        uo[:ni, nj - 1:nj + 1] = 1.0
        if implicit:      # initial condition fdm_implicit_cavity.f90:160-164 (u itself; p = 0)
            self.upload("u", uo)
        else:             # fdm_explicit_cavity.f90:160-166
            self.upload("uo", uo)
            self.upload("po", np.ones((ni + 1, nj + 1), order="F"))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.cfd_fdcav_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, name, a):
        a = np.asfortranarray(a, dtype=np.float64)
        assert a.shape == (self.ni + 1, self.nj + 1)
        check(self.lib.cfd_fdcav_upload(self._h, FDC_FIELDS.index(name), a.ctypes.data_as(C.c_void_p), self.ni + 1))

    def download(self, name):
        a = np.zeros((self.ni + 1, self.nj + 1), order="F")
        check(self.lib.cfd_fdcav_download(self._h, FDC_FIELDS.index(name), a.ctypes.data_as(C.c_void_p), self.ni + 1))
        return a

    def predict(self):
        check(self.lib.cfd_fdcav_predict(self._h))

    def correct(self) -> float:
        e = C.c_double()
        check(self.lib.cfd_fdcav_correct(self._h, C.byref(e)))
        return e.value

    def solve_pp(self) -> int:
        """sipsol on the device (fdm_explicit_cavity.f90:24-87); returns the SIP iterations used."""
        it = C.c_int()
        check(self.lib.cfd_fdcav_solve_pp(self._h, C.byref(it)))
        return it.value

    def step(self, solve_pp=None) -> float:
        """One time step.  solve_pp=None: everything on the GPU; else a host callable su -> pp."""
        if solve_pp is None:
            e, it = C.c_double(), C.c_int()
            check(self.lib.cfd_fdcav_step(self._h, C.byref(e), C.byref(it)))
            self.sip_iters = it.value
            return e.value
        self.predict()
        self.upload("pp", solve_pp(self.download("su")))
        return self.correct()

    # ---- implicit-Euler variant (fdm_implicit_cavity.f90:177-288) ----------------------------------------------
    def implicit_assemble(self, which: int):
        check(self.lib.cfd_fdcav_implicit_assemble(self._h, which))

    def implicit_solve(self, which: int) -> int:
        it = C.c_int()
        check(self.lib.cfd_fdcav_implicit_solve(self._h, which, C.byref(it)))
        return it.value

    def implicit_correct(self) -> float:
        e = C.c_double()
        check(self.lib.cfd_fdcav_implicit_correct(self._h, C.byref(e)))
        return e.value

    def implicit_step(self):
        e, it = C.c_double(), (C.c_int * 3)()
        check(self.lib.cfd_fdcav_implicit_step(self._h, C.byref(e), it))
        self.sip_iters = list(it)
        return e.value


def sipsol(aw, ae, as_, an, ap, su, phi, *, i2=None, j2=None, negate=True, alpha=0.80, tol=1.0e-6, maxit=3000, device=-1):
    """`call sipsol(1, i2, 1, j2, -aw, -ae, -as, -an, ap, su, phi)` (fdm_implicit_cavity.f90:29-92; negate=False: the arrays as
    given) on host arrays of one common shape (ld, ncols), Fortran order; phi is updated in place.  Returns (iterations, rsm)."""
    lib = _lib.load()
    arrs = [np.asfortranarray(a, dtype=np.float64) for a in (aw, ae, as_, an, ap, su)]
    if not (phi.flags.f_contiguous and phi.dtype == np.float64):
        raise ValueError("phi must be a Fortran-ordered float64 array (it is updated in place)")
    ld, ncols = phi.shape
    i2 = ld if i2 is None else i2
    j2 = ncols if j2 is None else j2
    it, rsm = C.c_int(), C.c_double()
    check(lib.cfd_sipsol(device, i2, j2, ld, *(a.ctypes.data_as(C.c_void_p) for a in arrs), phi.ctypes.data_as(C.c_void_p),
                         1 if negate else 0, alpha, tol, maxit, C.byref(it), C.byref(rsm)))
    return it.value, rsm.value
