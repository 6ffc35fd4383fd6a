"""Assembly launches of the general family (buoyancy cavity, n x n nodes) timed alone, and reproducibility checks.
  python tools/bench_gen_assemble.py 8194                 time assemble_u/v/p/t (10 launches each)
  python tools/bench_gen_assemble.py 8194 check           cfd_iterate against the subroutine path, bit for bit, 4 iterations
  python tools/bench_gen_assemble.py 8194 race 30 [uvpt]  the same launch 30 times on rough fields: identical bits every time?
  python tools/bench_gen_assemble.py cavity 4096 30       the same for the cavity family's TMA-staged u and v kernels
  python tools/bench_gen_assemble.py cavtime 8192         cavity family: assembly launches and one iteration
CFD_B200_GEN_TMA=0 / CFD_B200_ASM_TMA=0 select the register-window kernels for comparison."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import cfd_b200


def race_check(g, names, reps):
    """Run-to-run determinism of each assembly launch on rough fields: a stale read shows as a large difference."""
    for name in names:
        fn = getattr(g, "assemble_" + name)
        fn()
        ref = {c: g.download(c) for c in ("su", "ae", "an")}
        bad = 0
        for r in range(reps):
            fn()
            for c, a in ref.items():
                b = g.download(c)
                d = np.argwhere(a != b)
                if len(d):
                    bad += 1
                    print("RACE", name, c, "rep", r, "nodes", len(d), "first", d[:3].tolist(), "i", int(d[:, 0].min()),
                          int(d[:, 0].max()), "j", int(d[:, 1].min()), int(d[:, 1].max()), "max abs", float(np.abs(a - b).max()))
        print("race check", name, "reps", reps, "bad", bad, flush=True)


def main():
    if sys.argv[1] == "cavtime":       # cavity family: assembly launches and the whole iteration at n x n CV
        n = int(sys.argv[2])
        g = cfd_b200.CavityFVM(n, n, order="redblack")
        g.init_lid_driven(reynolds=1000.0)
        g.iterate(3)
        out = {"cavity_n": n, "tma": os.environ.get("CFD_B200_ASM_TMA", "1")}
        for name, fn in (("u", g.assemble_u), ("v", g.assemble_v), ("iteration", lambda: g.iterate(1))):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10):
                fn()
            b.record()
            torch.cuda.synchronize()
            out[name + "_us"] = round(a.elapsed_time(b) * 100.0, 1)
        print(json.dumps(out))
        return
    if sys.argv[1] == "cavity":
        n = int(sys.argv[2])
        g = cfd_b200.CavityFVM(n, n, order="redblack")
        g.init_lid_driven(reynolds=1000.0)
        g.iterate(2)
        rng = np.random.default_rng(5)
        for name, scale in (("u", 2e-1), ("v", 2e-1), ("p", 1e-2)):
            g.upload(name, np.asfortranarray(scale * (rng.random((n + 2, n + 2)) - 0.5)))
        race_check(g, ("u", "v"), int(sys.argv[3]))
        return
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8194
    g = cfd_b200.BuoyantCavity(n, n, order="redblack")
    g.iterate(3)
    out = {"n": n, "tma": os.environ.get("CFD_B200_GEN_TMA", "1"), "stages": os.environ.get("CFD_B200_ASM_STAGES", "")}
    cells = (n - 2) ** 2
    for name in ("u", "v", "p", "t"):
        fn = getattr(g, "assemble_" + name)
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            fn()
        b.record()
        torch.cuda.synchronize()
        us = a.elapsed_time(b) * 100.0
        out[name + "_us"] = round(us, 1)
        out[name + "_B_per_cell_at_7TBs"] = round(us * 1e-6 * 7.0e12 / cells, 1)
    print(json.dumps(out))
    if len(sys.argv) > 2 and sys.argv[2] == "race":
        # run-to-run determinism of each assembly launch on rough (random) fields: a stale read shows as a large difference
        rng = np.random.default_rng(5)
        for name, scale, base in (("u", 2e-3, 0.0), ("v", 2e-3, 0.0), ("p", 1e-6, 0.0), ("t", 0.4, 0.5), ("den", 0.2, 1.0),
                                  ("cp", 0.2, 1.0)):
            g.upload(name, np.asfortranarray(base + scale * (rng.random((n, n)) - 0.5)))
        race_check(g, tuple(sys.argv[4]) if len(sys.argv) > 4 else ("u", "v", "p", "t"), int(sys.argv[3]) if len(sys.argv) > 3 else 6)
        return
    if len(sys.argv) > 2:
        g.close()
        g = cfd_b200.BuoyantCavity(n, n, order="redblack")
        h = cfd_b200.BuoyantCavity(n, n, order="redblack")
        for it in range(4):
            g.iterate(1)
            h.calcu(); h.calcv(); h.calcp(); h.calct()
            for name in ("u", "v", "p", "t"):
                a, b = g.download(name), h.download(name)
                d = np.argwhere(a != b)
                if len(d):
                    print("iteration", it + 1, name, "differs in", len(d), "nodes; first", d[:5].tolist(),
                          "max abs", float(np.abs(a - b).max()), "values", float(a[tuple(d[0])]), float(b[tuple(d[0])]), "i range", int(d[:, 0].min()), int(d[:, 0].max()),
                          "j range", int(d[:, 1].min()), int(d[:, 1].max()))
        print("check done")


if __name__ == "__main__":
    main()
