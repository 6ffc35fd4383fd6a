#!/usr/bin/env python
"""The converged 256^2 cavity fixtures once more, from the SECOND restatement (tools/gen_kat.py:CavityRef) instead of the C
oracle: iteration count of the literal stop rule, monitor values, field sums and the centre-line profile.  About an hour
and a half of one core per Reynolds number.

    python tools/gen_cavity_cv256_regen.py 100     ->  tests/golden/cavity_cv256_re100_regen.json
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gen_kat as g  # noqa: E402


def main():
    re = float(sys.argv[1])
    cv = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    c = g.CavityRef(cv, cv, re)
    t0 = time.time()
    n, _ = c.run(maxit=2000000)
    ni, nj = c.ni, c.nj
    out = {"cv": cv, "reynolds": re, "iters": int(n), "residuals": [float(r) for r in c.res], "imon": int(c.imon), "jmon": int(c.jmon),
           "u_mon": float(c.u[c.imon, c.jmon]), "v_mon": float(c.v[c.imon, c.jmon]), "sum_abs": c.sums(),
           "u_centre": [float(z) for z in c.u[c.imon, 1:nj + 1]], "v_centre": [float(z) for z in c.v[1:ni + 1, c.jmon]],
           "seconds": time.time() - t0,
           "_source": "tools/gen_cavity_cv256_regen.py (second restatement, independent of oracle/cavity_oracle.c)"}
    path = os.path.join(g.GOLD, f"cavity_cv{cv}_re{int(re)}_regen.json")
    json.dump(out, open(path, "w"))
    print(path, n, "iterations in", round(time.time() - t0), "s")


if __name__ == "__main__":
    main()
