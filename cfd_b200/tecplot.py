"""Result writers of the reference drivers (host side, I/O only): Tecplot ASCII POINT format and the centre-line
text profiles of `program cavity_fvm` (lid-driven-cavity-2d/fvm-simple/cavity_fvm.f90:567-607), plus the corner
pressure extrapolation (:568-573).  The on-disk format is the contract for anyone diffing result files:
list-directed header lines, then `(*(1x,es14.7))` records with i fastest."""
from __future__ import annotations

import numpy as np


def es14_7(v: float) -> str:
    """Fortran edit descriptor es14.7 (one leading digit, 7 decimals, two-digit exponent, width 14)."""
    s = f"{v:.7E}"                      # e.g. -1.5782975E-01
    mant, exp = s.split("E")
    e = int(exp)
    out = f"{mant}E{'+' if e >= 0 else '-'}{abs(e):02d}"
    return out.rjust(14)


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
        f.write(f"zone  i={ni:3d}  j={nj:3d}  f=point\n")
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
