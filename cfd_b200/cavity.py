"""Host-side mirror of `program cavity_fvm` (lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90) over the C ABI.

The names are the reference's: module variables of vars_mod (cavity_fvm.f90:22-41) become attributes, and
calcu / calcv / calcp / lisolv keep their meaning (cavity_fvm.f90:92-355, 49-82) -- but the arrays live on the
GPU, and the host only sees them through `upload` / `download` (the reference driver touches them between
calls at cavity_fvm.f90:534-538 and 559-607).  Everything numeric runs in libcfd_b200.so; nothing here
computes a field on the CPU, and a missing library or GPU raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import CfdConfig, FIELD_ID, GEOM_ID, check


def setgrid(nicv: int, njcv: int, xstart: float, xend: float, ystart: float, yend: float):
    """setgrid (cavity_fvm.f90:366-414): uniform, cell-centred, backward-staggered grid.  Host logic (O(ni+nj))."""
    ni, nj = nicv + 2, njcv + 2
    x, xu, sew, sewu = (np.zeros(ni) for _ in range(4))
    y, yv, sns, snsv = (np.zeros(nj) for _ in range(4))
    xu[0] = xstart
    xu[1] = xstart
    dx = (xend - xstart) / nicv
    for i in range(2, ni):
        xu[i] = xu[i - 1] + dx
    x[0] = xu[1]
    for i in range(1, ni - 1):
        x[i] = 0.5 * (xu[i] + xu[i + 1])
        sew[i] = xu[i + 1] - xu[i]
    x[ni - 1] = xu[ni - 1]
    for i in range(2, ni - 1):
        sewu[i] = x[i] - x[i - 1]
    yv[0] = ystart
    yv[1] = ystart
    dy = (yend - ystart) / njcv
    for j in range(2, nj):
        yv[j] = yv[j - 1] + dy
    y[0] = yv[1]
    for j in range(1, nj - 1):
        y[j] = 0.5 * (yv[j] + yv[j + 1])
        sns[j] = yv[j + 1] - yv[j]
    y[nj - 1] = yv[nj - 1]
    for j in range(2, nj - 1):
        snsv[j] = y[j] - y[j - 1]
    return dict(x=x, xu=xu, sew=sew, sewu=sewu, y=y, yv=yv, sns=sns, snsv=snsv)


class CavityFVM:
    """One GPU context for the cavity family.  Indices are 1-based like the Fortran wherever they are arguments."""

    def __init__(self, nicv=10, njcv=10, *, iadv=3, order="redblack", urf=(0.5, 0.5, 0.8), nswp=(3, 3, 3),
                 ipref=None, jpref=None, device=-1, tdma_m=0,
                 ni_local=None, goff=0, halo_w=0, halo_e=0, sew_nim1=0.0,
                 nj_local=None, joff=0, halo_s=0, halo_n=0, sns_njm1=0.0, raw_coefficients=False):
        self.lib = _lib.load()
        self.nicv, self.njcv = nicv, njcv
        self.ni_global, self.nj = nicv + 2, njcv + 2
        self.nj_global, self.joff = self.nj, joff
        if nj_local is not None:
            self.nj = nj_local
        self.ni = ni_local if ni_local is not None else self.ni_global
        self.nim1, self.njm1 = self.ni_global - 1, self.nj_global - 1
        self.iadv = iadv
        self.urfu, self.urfv, self.urfp = urf
        self.nswpu, self.nswpv, self.nswpp = nswp
        self.ipref = ipref if ipref is not None else self.ni_global // 2      # cavity_fvm.f90:549-550
        self.jpref = jpref if jpref is not None else self.nj_global // 2
        self.imon, self.jmon = self.ipref, self.jpref
        self.goff = goff
        cfg = CfdConfig()
        cfg.family = _lib.FAMILY_CAVITY
        cfg.ni, cfg.nj = self.ni, self.nj
        cfg.iadv = iadv
        cfg.order = _lib.ORDER_SEQUENTIAL if order in ("sequential", "seq", 1) else _lib.ORDER_REDBLACK
        for k in range(3):
            cfg.nswp[k] = nswp[k]
            cfg.urf[k] = urf[k]
        cfg.ipref, cfg.jpref = self.ipref, self.jpref
        cfg.device = device
        cfg.tdma_m = tdma_m
        cfg.goff = goff
        cfg.ni_global = self.ni_global if (ni_local is not None) else 0
        cfg.halo_w, cfg.halo_e = halo_w, halo_e
        cfg.sew_nim1 = sew_nim1
        cfg.joff = joff
        cfg.nj_global = self.nj_global if (nj_local is not None) else 0
        cfg.halo_s, cfg.halo_n = halo_s, halo_n
        cfg.sns_njm1 = sns_njm1
        cfg.raw_coefficients = 1 if raw_coefficients else 0
        self._cfg = cfg
        self._h = C.c_void_p()
        check(self.lib.cfd_create(C.byref(self._h), C.byref(cfg)))
        self.resoru = self.resorv = self.resorm = 0.0

    # -- life cycle ---------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.cfd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def layout(self):
        m, t, ld = C.c_int(), C.c_int(), C.c_int()
        check(self.lib.cfd_layout_info(self._h, C.byref(m), C.byref(t), C.byref(ld)))
        return dict(M=m.value, T=t.value, LD=ld.value)

    @property
    def launches(self) -> int:
        return int(self.lib.cfd_launch_count(self._h))

    def set_stream(self, cuda_stream: int):
        check(self.lib.cfd_set_stream(self._h, C.c_void_p(cuda_stream)))

    def synchronize(self):
        check(self.lib.cfd_synchronize(self._h))

    # -- data movement ------------------------------------------------------------------------------
    def set_geometry(self, **arrays):
        for name, a in arrays.items():
            a = np.ascontiguousarray(a, dtype=np.float64)
            check(self.lib.cfd_set_geometry(self._h, GEOM_ID[name], a.ctypes.data_as(C.c_void_p), a.size))

    def setgrid(self, xstart=0.0, xend=1.0, ystart=0.0, yend=1.0):
        """call setgrid(...) of the reference driver (cavity_fvm.f90:532); single-GPU contexts only."""
        g = setgrid(self.nicv, self.njcv, xstart, xend, ystart, yend)
        for k, v in g.items():
            setattr(self, k, v)
        self.set_geometry(**g)
        return g

    def upload(self, name: str, a: np.ndarray):
        a = np.asfortranarray(a, dtype=np.float64)
        assert a.shape == (self.ni, self.nj), (a.shape, self.ni, self.nj)
        check(self.lib.cfd_upload(self._h, FIELD_ID[name], a.ctypes.data_as(C.c_void_p), a.shape[0]))

    def upload_noalloc(self, name: str, a: np.ndarray):
        """cfd_upload_noalloc: same effect as `upload`, staged in the library's built-in 32 MiB device buffer
        (no cudaMalloc at run time); grids up to ni*nj = 4 194 304 nodes.  `a` may be the leading (ni, nj) part of a
        larger column-major array (leading dimension > ni), like the reference's (it, jt) arrays."""
        assert a.dtype == np.float64 and a.shape == (self.ni, self.nj), (a.dtype, a.shape, self.ni, self.nj)
        if not (a.strides[0] == 8 and a.strides[1] % 8 == 0 and a.strides[1] >= 8 * self.ni):
            a = np.asfortranarray(a)
        check(self.lib.cfd_upload_noalloc(self._h, FIELD_ID[name], a.ctypes.data_as(C.c_void_p), a.strides[1] // 8))

    def upload_ptr(self, name: str, ptr: int, ld: int):
        check(self.lib.cfd_upload(self._h, FIELD_ID[name], C.c_void_p(ptr), ld))

    def download_ptr(self, name: str, ptr: int, ld: int):
        check(self.lib.cfd_download(self._h, FIELD_ID[name], C.c_void_p(ptr), ld))

    def download(self, name: str) -> np.ndarray:
        a = np.zeros((self.ni, self.nj), order="F")
        check(self.lib.cfd_download(self._h, FIELD_ID[name], a.ctypes.data_as(C.c_void_p), a.shape[0]))
        return a

    def probe(self, name: str, i: int, j: int) -> float:
        v = C.c_double()
        check(self.lib.cfd_probe(self._h, FIELD_ID[name], i, j, C.byref(v)))
        return v.value

    # -- post-processing on the device fields (no whole-field download) -------------------------------------
    def extract_line(self, name: str, i: int) -> np.ndarray:
        """A(i, 1:nj), local 1-based i."""
        out = np.zeros(self.nj)
        check(self.lib.cfd_extract_line(self._h, FIELD_ID[name], i, out.ctypes.data_as(C.POINTER(C.c_double)), self.nj))
        return out

    def extract_row(self, name: str, j: int) -> np.ndarray:
        """A(1:ni, j)."""
        out = np.zeros(self.ni)
        check(self.lib.cfd_extract_row(self._h, FIELD_ID[name], j, out.ctypes.data_as(C.POINTER(C.c_double)), self.ni))
        return out

    def centreline_uc(self, imon: int) -> np.ndarray:
        """uc(imon, 1:nj) of centred_vel (cavity_fvm.f90:416-430): the third column of uvel.txt (:600-603)."""
        out = np.zeros(self.nj)
        check(self.lib.cfd_centreline(self._h, 0, imon, out.ctypes.data_as(C.POINTER(C.c_double)), self.nj))
        return out

    def centreline_vc(self, jmon: int) -> np.ndarray:
        """vc(1:ni, jmon): the third column of vvel.txt (cavity_fvm.f90:604-607)."""
        out = np.zeros(self.ni)
        check(self.lib.cfd_centreline(self._h, 1, jmon, out.ctypes.data_as(C.POINTER(C.c_double)), self.ni))
        return out

    def reference_pressure(self):
        """p = p - p(ipref,jpref) and the corner values (cavity_fvm.f90:568-573), on the device."""
        check(self.lib.cfd_reference_pressure(self._h))

    def device_view(self, name: str):
        base, stride, n = C.c_void_p(), C.c_longlong(), C.c_int()
        check(self.lib.cfd_field_device_view(self._h, FIELD_ID[name], C.byref(base), C.byref(stride), C.byref(n)))
        return base.value, stride.value, n.value

    # -- the reference's subroutines ----------------------------------------------------------------
    def calcu(self) -> float:
        r = C.c_double()
        check(self.lib.cfd_calcu(self._h, C.byref(r)))
        self.resoru = r.value
        return r.value

    def calcv(self) -> float:
        r = C.c_double()
        check(self.lib.cfd_calcv(self._h, C.byref(r)))
        self.resorv = r.value
        return r.value

    def calcp(self) -> float:
        r = C.c_double()
        check(self.lib.cfd_calcp(self._h, C.byref(r)))
        self.resorm = r.value
        return r.value

    def lisolv(self, istart: int, jstart: int, phi: str):
        check(self.lib.cfd_lisolv(self._h, istart, jstart, FIELD_ID[phi]))

    def iterate(self, n: int = 1):
        res = (C.c_double * 8)()
        check(self.lib.cfd_iterate(self._h, n, res))
        self.resoru, self.resorv, self.resorm = res[0], res[1], res[2]
        return self.resoru, self.resorv, self.resorm

    def iterate_async(self, n: int = 1):
        check(self.lib.cfd_iterate(self._h, n, None))

    def residuals(self):
        res = (C.c_double * 8)()
        check(self.lib.cfd_residuals(self._h, res))
        return tuple(res)

    # -- stage-level calls (multi-GPU driver) -------------------------------------------------------
    def assemble_u(self):
        check(self.lib.cfd_assemble_u(self._h))

    def assemble_v(self):
        check(self.lib.cfd_assemble_v(self._h))

    def assemble_p(self):
        check(self.lib.cfd_assemble_p(self._h))

    def sweep(self, istart, jstart, phi, colour, first=False):
        check(self.lib.cfd_sweep(self._h, istart, jstart, FIELD_ID[phi], colour, 1 if first else 0))

    def sweep_part(self, istart, jstart, phi, colour, first=False, part=0):
        check(self.lib.cfd_sweep_part(self._h, istart, jstart, FIELD_ID[phi], colour, 1 if first else 0, part))

    def sweeps(self, istart, jstart, phi, nsweeps, first=False):
        """cfd_sweeps: the `do n = 1, nswp; call lisolv` loop, all colour passes in one flag-ordered launch."""
        check(self.lib.cfd_sweeps(self._h, istart, jstart, FIELD_ID[phi], nsweeps, 1 if first else 0))

    # -- i-slab neighbours over peer memory -------------------------------------------------------------------
    IPC_FLAGS = -1

    def ipc_export(self, what) -> bytes:
        """64-byte CUDA IPC handle of a plain field's allocation (`what` = field name) or of the flag block ("flags")."""
        buf = C.create_string_buffer(64)
        check(self.lib.cfd_ipc_export(self._h, self.IPC_FLAGS if what == "flags" else FIELD_ID[what], buf))
        return buf.raw

    def ipc_open(self, handle: bytes) -> int:
        p = C.c_void_p()
        check(self.lib.cfd_ipc_open(self._h, C.create_string_buffer(handle, 64), C.byref(p)))
        return p.value

    def peer_attach(self, side: int, what, devptr: int, peer_ni: int, peer_hw: int, peer_he: int):
        check(self.lib.cfd_peer_attach(self._h, side, self.IPC_FLAGS if what == "flags" else FIELD_ID[what],
                                       C.c_void_p(devptr), peer_ni, peer_hw, peer_he))

    def halo_push(self, spec):
        """spec: sequence of (field name, w, e): my last w owned lines go to the east neighbour's west halo, my first e
        owned lines to the west neighbour's east halo; returns when (on the stream) the neighbours' lines are here."""
        n = len(spec)
        arr = lambda vals: (C.c_int * n)(*vals)
        check(self.lib.cfd_halo_push(self._h, n, arr([FIELD_ID[f] for f, _, _ in spec]), arr([w for _, w, _ in spec]),
                                     arr([e for _, _, e in spec])))

    def sweeps_single_launch(self) -> bool:
        y = C.c_int()
        check(self.lib.cfd_sweeps_single_launch(self._h, C.byref(y)))
        return bool(y.value)

    def peer_detach(self):
        check(self.lib.cfd_peer_detach(self._h))

    def bcast_attach(self, rank: int, flag_block_ptr, rank_ni: int):
        check(self.lib.cfd_bcast_attach(self._h, rank, C.c_void_p(flag_block_ptr) if flag_block_ptr else None, rank_ni))

    def scalar_bcast(self, owner: bool, name: str, i: int, j: int) -> int:
        """All ranks call it; returns the device address where the owner's A(i,j) arrives (peer-memory broadcast)."""
        p = C.c_void_p()
        check(self.lib.cfd_scalar_bcast(self._h, 1 if owner else 0, FIELD_ID[name], i, j, C.byref(p)))
        return p.value

    def peer_status(self) -> bool:
        """True if a bounded wait on a neighbour ever timed out."""
        a = C.c_int()
        check(self.lib.cfd_peer_status(self._h, C.byref(a)))
        return bool(a.value)

    def correct(self, ppref_device_ptr: int | None = None):
        check(self.lib.cfd_correct(self._h, C.c_void_p(ppref_device_ptr) if ppref_device_ptr else None))

    # -- j-slabs: partition method for lines cut across contexts (device pointers as ints) ------------------
    def spike_basis(self, istart, jstart):
        check(self.lib.cfd_spike_basis(self._h, istart, jstart))

    def spike_pack(self, istart, jstart, phi, colour, which, buf_ptr: int):
        check(self.lib.cfd_spike_pack(self._h, istart, jstart, FIELD_ID[phi], colour, which, C.c_void_p(buf_ptr)))

    def spike_finish(self, istart, jstart, phi, colour, wz_all_ptr: int, y_all_ptr: int, rank: int, world: int):
        check(self.lib.cfd_spike_finish(self._h, istart, jstart, FIELD_ID[phi], colour, C.c_void_p(wz_all_ptr),
                                        C.c_void_p(y_all_ptr), rank, world))

    def rows_pack(self, name, j_first, nrows, buf_ptr: int):
        check(self.lib.cfd_rows_pack(self._h, FIELD_ID[name], j_first, nrows, C.c_void_p(buf_ptr)))

    def rows_unpack(self, name, j_first, nrows, buf_ptr: int):
        check(self.lib.cfd_rows_unpack(self._h, FIELD_ID[name], j_first, nrows, C.c_void_p(buf_ptr)))

    def residual_device_ptr(self) -> int:
        p = C.c_void_p()
        check(self.lib.cfd_residual_device_ptr(self._h, C.byref(p)))
        return p.value

    # -- the literals of the reference driver ----------------------------------------------------------
    def init_lid_driven(self, reynolds=100.0, utop=1.0, rho=1000.0, xlim=(0.0, 1.0), ylim=(0.0, 1.0)):
        """cavity_fvm.f90:527-538: grid, fluid properties, lid velocity (single-GPU contexts)."""
        self.setgrid(xlim[0], xlim[1], ylim[0], ylim[1])
        den = np.full((self.ni, self.nj), rho, order="F")
        vis = np.asfortranarray(den * utop * (xlim[1] - xlim[0]) / reynolds)
        u = np.zeros((self.ni, self.nj), order="F")
        u[1:, self.nj - 1] = utop
        self.upload("den", den)
        self.upload("vis", vis)
        self.upload("u", u)

    def run(self, maxit=90000, tol=None, check_every=1, rule="literal", flux_scales=None, stationary_checks=3):
        """The outer loop cavity_fvm.f90:553-566.  check_every>1 tests the stop rule every that many passes.

        rule="literal"     the reference's own test, max(resoru, resorv, resorm) < 1e-10 (cavity_fvm.f90:565).  The sums
                           run over all cells and are not normalised, so their rounding floor grows with the grid:
                           the reference's own 52 x 52 cylinder case already floors at 1.26e-10 and never stops
                           (SURVEY 8c), and at 4096^2 the rule cannot fire.  Kept as the default (it is what reproduces
                           the reference's 232 / 2093 / 8258 iterations).
        rule="scaled"      the same three sums divided by what one cell carries: (number of cells) x (reference momentum
                           flux rho*U^2*dx for u, v; mass flux rho*U*dx for the continuity sum) -- the mean relative
                           imbalance per cell, which does not grow with the grid; default tol 1e-13.  `flux_scales` =
                           (momentum, mass) overrides rho*U^2*dx, rho*U*dx (defaults: the lid-driven case's rho = 1000,
                           U = 1, dx = 1/nicv, i.e. 1e-10 of the literal rule at the default 10 x 10 grid <-> 1e-14).
        rule="stationary"  stop when u, v, p at the monitor node are bit-for-bit unchanged over `stationary_checks`
                           consecutive checks: the iteration has reached its floating-point fixed point.
        """
        if rule not in ("literal", "scaled", "stationary"):
            raise ValueError("rule must be 'literal', 'scaled' or 'stationary'")
        if tol is None:
            tol = 1.0e-10 if rule == "literal" else 1.0e-13
        ncell = float(self.nicv * self.njcv)
        dx = 1.0 / self.nicv
        fm, fc = flux_scales if flux_scales is not None else (1000.0 * dx, 1000.0 * dx)
        it, same, last = 0, 0, None
        while it < maxit:
            n = min(check_every, maxit - it)
            ru, rv, rm = self.iterate(n)
            it += n
            if rule == "literal":
                if max(ru, rv, rm) < tol:
                    break
            elif rule == "scaled":
                if max(ru / (ncell * fm), rv / (ncell * fm), rm / (ncell * fc)) < tol:
                    break
            else:
                mon = tuple(self.probe(f, self.imon, self.jmon) for f in ("u", "v", "p"))
                same = same + 1 if mon == last else 0
                last = mon
                if same >= stationary_checks:
                    break
        return it


def centred_vel(u: np.ndarray, v: np.ndarray):
    """centred_vel (cavity_fvm.f90:416-430) on downloaded fields -- post-processing, host side."""
    ni, nj = u.shape
    uc, vc = np.zeros_like(u), np.zeros_like(v)
    uc[0, :] = u[1, :]
    uc[ni - 1, :] = u[ni - 1, :]
    vc[:, 0] = v[:, 1]
    vc[:, nj - 1] = v[:, nj - 1]
    uc[1:ni - 1, :] = 0.5 * (u[1:ni - 1, :] + u[2:ni, :])
    vc[:, 1:nj - 1] = 0.5 * (v[:, 1:nj - 1] + v[:, 2:nj])
    return uc, vc
