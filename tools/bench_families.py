"""Timing of the other BASELINE.json configs on one GPU (not the headline bench line): buoyancy cavity 8192^2
(config 4), backward step 16384 x 4096 (config 5), implicit heat 16384^2.  Prints one JSON line per case with the
SURVEY 8(d) algorithmic bytes (1224 B and 1056 B per cell and outer iteration)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import cfd_b200

PEAK = 6551.0
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(fn, n, warm=3):
    fn(warm)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fn(n)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    which = sys.argv[1:] or ["buoyancy", "step", "implicit"]
    if "buoyancy" in which:
        n = 8194
        g = cfd_b200.BuoyantCavity(n, n, order="redblack")
        ms = timed(lambda k: g.iterate(k), 10)
        cells = (n - 2) ** 2
        r = g.iterate(1)
        print(json.dumps({"case": "buoyancy cavity Ra=1e6 8192^2 (BASELINE configs[3])", "ms_per_iteration": ms,
                          "GLUPS": cells / ms / 1e6, "algorithmic_GBs": 1224 * cells / ms / 1e6,
                          "frac_of_measured_hbm": 1224 * cells / ms / 1e6 / PEAK, "layout": g.layout,
                          "finite": bool(np.isfinite(r).all())}), flush=True)
        g.close()
    if "step" in which:
        ni, nj = 16384, 4096
        g = cfd_b200.TeachFVM(ni, nj, epsx=1.0, order="redblack")
        ms = timed(lambda k: g.iterate(k), 10)
        cells = (ni - 2) * (nj - 2)
        r = g.iterate(1)
        print(json.dumps({"case": "backward step 16384x4096 laminar, nswp 3/3/5 (BASELINE configs[4])", "ms_per_iteration": ms,
                          "GLUPS": cells / ms / 1e6, "algorithmic_GBs": 1056 * cells / ms / 1e6,
                          "frac_of_measured_hbm": 1056 * cells / ms / 1e6 / PEAK, "layout": g.layout,
                          "finite": bool(np.isfinite(r).all())}), flush=True)
        g.close()
    if "step_turb" in which:
        ni, nj = 16384, 4096
        g = cfd_b200.TeachFVM(ni, nj, epsx=1.0, order="redblack", turbulent=True)
        ms = timed(lambda k: g.iterate_async(k), 10)
        cells = (ni - 2) * (nj - 2)
        r = g.iterate_turb(1)
        print(json.dumps({"case": "backward step 16384x4096 as shipped: k-epsilon + energy + props (SURVEY 8d: ~1992 B/cell)",
                          "ms_per_iteration": ms, "GLUPS": cells / ms / 1e6, "algorithmic_GBs": 1992 * cells / ms / 1e6,
                          "frac_of_measured_hbm": 1992 * cells / ms / 1e6 / PEAK, "layout": g.layout,
                          "finite": bool(np.isfinite(r).all()), "residuals": list(r)}), flush=True)
        g.close()
    if "implicit" in which:
        n = 16384
        h = cfd_b200.Heat2D(n, n, fo=5.0)
        ms = timed(lambda k: (h.implicit_step(k), h.synchronize()), 5, warm=1)
        print(json.dumps({"case": "implicit (LOD) heat step 16384^2", "ms_per_step": ms, "GLUPS": n * n / ms / 1e6,
                          "note": "4 passes x (read+write) = 64 B/cell: %.0f GB/s" % (64 * n * n / ms / 1e6)}), flush=True)
        h.close()


if __name__ == "__main__":
    main()
