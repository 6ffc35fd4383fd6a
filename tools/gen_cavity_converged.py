#!/usr/bin/env python
"""Converged-field fixtures of the lid-driven cavity at grids the CPU oracle cannot converge inside a test run.

Runs oracle/cavity_oracle.c (the bit-reference build, the reference's sequential line order and literal stop rule
max(resoru, resorv, resorm) < 1e-10, cavity_fvm.f90:553-566) to convergence and stores what a GPU parity test needs,
small enough to commit: iteration count, monitor values, |.|-sums of u, v, p and the two centre-line profiles the
reference itself writes (uvel.txt / vvel.txt, cavity_fvm.f90:600-607) plus a strided sample of u, v, p.

    python tools/gen_cavity_converged.py 256 100      ->  tests/golden/cavity_cv256_re100.npz   (about an hour of one core)

The generating run happens in the build container (the GPU box has no use for it); the .npz is the fixture.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402


def main():
    cv, re = int(sys.argv[1]), float(sys.argv[2])
    maxit = int(sys.argv[3]) if len(sys.argv) > 3 else 2000000
    orc.build()
    o = orc.Cavity(cv, cv, re)
    t0 = time.time()
    n = o.run(maxit=maxit)
    uc, vc = o.centred_vel()
    st = max(1, cv // 32)
    out = os.path.join(ROOT, "tests", "golden", f"cavity_cv{cv}_re{int(re)}.npz")
    np.savez_compressed(
        out, cv=cv, reynolds=re, iters=n, residuals=np.array(o.residuals), imon=o.imon, jmon=o.jmon,
        u_mon=o.u[o.imon - 1, o.jmon - 1], v_mon=o.v[o.imon - 1, o.jmon - 1], uc_mon=uc[o.imon - 1, o.jmon - 1],
        sum_abs=np.array([np.abs(o.u).sum(), np.abs(o.v).sum(), np.abs(o.p).sum()]),
        u_centre=o.u[o.imon - 1, :].copy(), v_centre=o.v[:, o.jmon - 1].copy(), p_centre=o.p[o.imon - 1, :].copy(),
        stride=st, u_s=o.u[::st, ::st].copy(), v_s=o.v[::st, ::st].copy(), p_s=o.p[::st, ::st].copy())
    print(f"{out}: {n} iterations in {time.time() - t0:.0f} s, residuals {o.residuals}")


if __name__ == "__main__":
    main()
