#!/usr/bin/env python
"""Time the line solver both ways on one GPU: one launch per colour pass (cfd_sweep) and all passes of an equation in one
flag-ordered launch (cfd_sweeps).  Usage: python tools/bench_sweeps.py [ncv=4096] [nsweeps=3]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cfd_b200

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nsw = int(sys.argv[2]) if len(sys.argv) > 2 else 3
g = cfd_b200.CavityFVM(n, n, order="redblack")
g.init_lid_driven(reynolds=1000.0)
st = torch.cuda.Stream()
g.set_stream(st.cuda_stream)
g.iterate(2)


def timed(fn, reps=5):
    out = []
    with torch.cuda.stream(st):
        for _ in range(reps):
            g.assemble_u()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st); fn(); b.record(st)
            torch.cuda.synchronize()
            out.append(a.elapsed_time(b))
    return out


def per_pass():
    for _ in range(nsw):
        g.sweep(3, 2, "u", 0)
        g.sweep(3, 2, "u", 1)


print("layout", g.layout, "single launch eligible", g.sweeps_single_launch())
print("per-pass  x%d sweeps [ms]:" % nsw, ["%.3f" % t for t in timed(per_pass)])
print("one launch x%d sweeps [ms]:" % nsw, ["%.3f" % t for t in timed(lambda: g.sweeps(3, 2, "u", nsw))])
print("per-pass  again       [ms]:", ["%.3f" % t for t in timed(per_pass)])
cells = (n - 1) * n * nsw
for name, fn in (("per-pass", per_pass), ("one launch", lambda: g.sweeps(3, 2, "u", nsw))):
    t = min(timed(fn, 8))
    print("%-10s best %.3f ms = %.0f GB/s on 64 B/cell" % (name, t, 64 * cells / t / 1e6))
print("timed out:", g.peer_status())
if os.environ.get("CFD_B200_FLOW_TIMING"):
    import ctypes as C
    out = (C.c_double * 8)()
    g.lib.cfd_debug_flow_timing(g._h, out)
    g.sweeps(3, 2, "u", nsw)
    torch.cuda.synchronize()
    g.lib.cfd_debug_flow_timing(g._h, out)
    nl = max(out[0], 1.0)
    names = ("slices", "backward", "reduction", "exchange", "stores", "end barrier")
    print("block 0: %d lines; cycles per line: " % nl + ", ".join("%s %.0f" % (k, out[1 + i] / nl) for i, k in enumerate(names)))
