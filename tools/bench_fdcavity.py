"""Timing of the explicit FD cavity (fdm_explicit_cavity.f90) on one GPU: the stencil passes of one time step
(predict + correct, SURVEY 8d: ~120 B/cell) and the device SIP solve of the same step (17 doubles/cell per inner
iteration), at a few grid sizes.  Prints one JSON line per size."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import cfd_b200


def main():
    for n in [int(a) for a in sys.argv[1:]] or [1025, 4097]:
        g = cfd_b200.FDCavity(n, n, re=100.0, dt=1.0e-4)
        for _ in range(2):
            g.predict(); it = g.solve_pp(); g.correct()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        K = 2
        for _ in range(K):
            g.predict()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        its = 0
        for _ in range(K):
            its += g.solve_pp()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        for _ in range(K):
            g.correct()
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        cells = n * n
        sten_ms = ((t1 - t0) + (t3 - t2)) / K * 1e3
        sip_ms = (t2 - t1) / K * 1e3
        print(json.dumps({"case": f"explicit FD cavity {n}x{n}", "stencil_ms_per_step": sten_ms,
                          "stencil_GBs": 120 * cells / sten_ms / 1e6, "sip_ms_per_step": sip_ms,
                          "sip_inner_iterations_per_step": its / K,
                          "sip_GBs": 17 * 8 * cells * (its / K) / sip_ms / 1e6}), flush=True)
        g.close()


if __name__ == "__main__":
    main()
