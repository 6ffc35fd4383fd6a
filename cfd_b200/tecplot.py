"""Result writers of the reference drivers (host side, I/O only): Tecplot ASCII POINT format and the centre-line
text profiles of `program cavity_fvm` (lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90:567-607), plus the corner
pressure extrapolation (:568-573).  The on-disk format is the contract for anyone diffing result files:
list-directed header lines, then `(*(1x,es14.7))` records with i fastest."""
from __future__ import annotations

import numpy as np


def es14_7(v: float) -> str:
    """Fortran edit descriptor es14.7 (one leading digit, 7 decimals, width 14).  Exponents of one or two digits
    print as E+dd; a three-digit exponent drops the letter (`-1.2345678-100`), as the Fortran standard prescribes
    for Ew.d output, so the field keeps its width."""
    s = f"{v:.7E}"                      # e.g. -1.5782975E-01
    mant, exp = s.split("E")
    e = int(exp)
    sign = "+" if e >= 0 else "-"
    out = f"{mant}E{sign}{abs(e):02d}" if abs(e) < 100 else f"{mant}{sign}{abs(e):03d}"
    return out.rjust(14)


def i3(n: int) -> str:
    """Fortran edit descriptor i3: a value that does not fit three characters prints as `***` (the reference's zone
    header does exactly that beyond 999 nodes, cavity_fvm.f90:593; teach_tecprint below follows the same rule)."""
    s = f"{n:3d}"
    return s if len(s) <= 3 else "***"


def corner_pressure(p: np.ndarray, ipref: int, jpref: int) -> np.ndarray:
    """cavity_fvm.f90:568-573: shift by p(ipref,jpref) and extrapolate the four corner values."""
    p = np.array(p, dtype=np.float64, order="F", copy=True)
    ni, nj = p.shape
    p -= p[ipref - 1, jpref - 1]
    p[0, 0] = p[1, 0] + p[0, 1] - p[1, 1]
    p[0, nj - 1] = p[0, nj - 2] + p[1, nj - 1] - p[1, nj - 2]
    p[ni - 1, 0] = p[ni - 2, 0] + p[ni - 1, 1] - p[ni - 2, 1]
    p[ni - 1, nj - 1] = p[ni - 1, nj - 2] + p[ni - 2, nj - 1] - p[ni - 2, nj - 2]
    return p


def write_result_dat(path, x, y, uc, vc, p, title="Lid Driven Cavity Flow"):
    """cavity_fvm.f90:590-598: `result.dat`."""
    ni, nj = uc.shape
    with open(path, "w") as f:
        f.write(f' title="{title}"\n')
        f.write(' variables="x", "y", "u", "v", "p"\n')
        f.write(f"zone  i={i3(ni)}  j={i3(nj)}  f=point\n")
        for j in range(nj):
            for i in range(ni):
                f.write("".join(" " + es14_7(v) for v in (x[i], y[j], uc[i, j], vc[i, j], p[i, j])) + "\n")


def write_centrelines(path_u, path_v, x, y, uc, vc, imon, jmon):
    """cavity_fvm.f90:600-607: `uvel.txt` (u along the vertical centre line) and `vvel.txt`."""
    ni, nj = uc.shape
    with open(path_u, "w") as f:
        for j in range(nj):
            f.write("".join(" " + es14_7(v) for v in (x[imon - 1], y[j], uc[imon - 1, j])) + "\n")
    with open(path_v, "w") as f:
        for i in range(ni):
            f.write("".join(" " + es14_7(v) for v in (x[i], y[jmon - 1], vc[i, jmon - 1])) + "\n")


# ---- general family: postproc (buoyancy-driven-cavity-2d/buoyancy_cav.f90:435-469) -------------------------------
def general_postproc(u, v, p, t, xdif, ydif, ipref, jpref, lprint=(True, True, True, True)):
    """Cell-centred velocities, pressure extrapolated to the boundaries / corners and referenced to (ipref,jpref),
    corner temperatures -- on downloaded fields, same statement order as the reference.  Returns (uc, vc, p, t)."""
    ni, nj = u.shape
    uc, vc = np.zeros_like(u), np.zeros_like(v)
    p, t = np.array(p, order="F", copy=True), np.array(t, order="F", copy=True)
    if lprint[0]:
        uc[0, :] = u[1, :]
        uc[ni - 1, :] = u[ni - 1, :]
        uc[1:ni - 1, :] = 0.5 * (u[1:ni - 1, :] + u[2:ni, :])
    if lprint[1]:
        vc[:, 0] = v[:, 1]
        vc[:, nj - 1] = v[:, nj - 1]
        vc[:, 1:nj - 1] = 0.5 * (v[:, 1:nj - 1] + v[:, 2:nj])
    if lprint[2]:
        p[0, :] = p[1, :] + (p[1, :] - p[2, :]) * xdif[1] / xdif[2]
        p[ni - 1, :] = p[ni - 2, :] + (p[ni - 2, :] - p[ni - 3, :]) * xdif[ni - 1] / xdif[ni - 2]
        p[:, 0] = p[:, 1] + (p[:, 1] - p[:, 2]) * ydif[1] / ydif[2]
        p[:, nj - 1] = p[:, nj - 2] + (p[:, nj - 2] - p[:, nj - 3]) * ydif[nj - 1] / ydif[nj - 2]
        p[0, 0] = p[1, 0] + p[0, 1] - p[1, 1]
        p[0, nj - 1] = p[0, nj - 2] + p[1, nj - 1] - p[1, nj - 2]
        p[ni - 1, 0] = p[ni - 2, 0] + p[ni - 1, 1] - p[ni - 2, 1]
        p[ni - 1, nj - 1] = p[ni - 1, nj - 2] + p[ni - 2, nj - 1] - p[ni - 2, nj - 2]
        p[:, :] = p - p[ipref - 1, jpref - 1]
    if lprint[3]:
        t[0, 0] = t[1, 0] + t[0, 1] - t[1, 1]
        t[0, nj - 1] = t[0, nj - 2] + t[1, nj - 1] - t[1, nj - 2]
        t[ni - 1, 0] = t[ni - 2, 0] + t[ni - 1, 1] - t[ni - 2, 1]
        t[ni - 1, nj - 1] = t[ni - 1, nj - 2] + t[ni - 2, nj - 1] - t[ni - 2, nj - 2]
    return uc, vc, p, t


# ---- TEACH family: centred_vels, tecprint, velprint (backward-step-flow-2d/general.f90:238-289, 331-350) ---------
def teach_centred_vels(u, v):
    ni, nj = u.shape
    uc, vc = np.zeros_like(u), np.zeros_like(v)
    uc[2:ni - 1, 1:nj] = u[2:ni - 1, 1:nj] + 0.5 * (u[3:ni, 1:nj] - u[2:ni - 1, 1:nj])
    vc[1:ni, 2:nj - 1] = v[1:ni, 2:nj - 1] + 0.5 * (v[1:ni, 2:nj - 1] - v[1:ni, 3:nj])
    return uc, vc


def _es10_3(v: float) -> str:
    return "%10.3E" % v


def teach_tecprint(path, istart, jstart, x, y, phi, mode="w"):
    """tecprint: format 120 `/'ZONE',3x,'I=',i3,3x,'j=',i3,3x,'F=POINT'` (i3 prints *** beyond 999, as Fortran does),
    then one `1x,es10.3,3x,es10.3,3x,es10.3` record per node, i fastest."""
    ni, nj = phi.shape

    def i3(n):
        return "%3d" % n if -99 <= n <= 999 else "***"
    with open(path, mode) as f:
        f.write("\n" + "ZONE" + "   " + "I=" + i3(ni - 1) + "   " + "j=" + i3(nj - 1) + "   " + "F=POINT" + "\n")
        for j in range(jstart, nj + 1):
            for i in range(istart, ni + 1):
                f.write(" " + _es10_3(x[i - 1]) + "   " + _es10_3(y[j - 1]) + "   " + _es10_3(phi[i - 1, j - 1]) + "\n")


def teach_velprint(path, istart, jstart, x, y, uvel, vvel, mode="w"):
    ni, nj = uvel.shape

    def i3(n):
        return "%3d" % n if -99 <= n <= 999 else "***"
    with open(path, mode) as f:
        f.write("\n" + "ZONE" + "   " + "I=" + i3(ni - 1) + "   " + "j=" + i3(nj - 1) + "   " + "F=POINT" + "\n")
        for j in range(jstart, nj + 1):
            for i in range(istart, ni + 1):
                f.write(" " + _es10_3(x[i - 1]) + "   " + _es10_3(y[j - 1]) + "   " + _es10_3(uvel[i - 1, j - 1]) + "   "
                        + _es10_3(vvel[i - 1, j - 1]) + "\n")
