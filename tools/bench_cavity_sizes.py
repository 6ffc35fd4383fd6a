#!/usr/bin/env python
"""Outer-iteration time of the cavity case at other grid sizes than the headline one.  Usage: python tools/bench_cavity_sizes.py 8192 [16384 ...]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cfd_b200

for n in [int(a) for a in sys.argv[1:]] or [8192]:
    g = cfd_b200.CavityFVM(n, n, order="redblack")
    g.init_lid_driven(reynolds=1000.0)
    g.iterate(3)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k = 10
    a.record(); g.iterate_async(k); b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / k
    print(json.dumps({"cavity": f"{n}^2", "ms_per_iteration": ms, "GLUPS": n * n / ms / 1e6, "algorithmic_GBs": 928 * n * n / ms / 1e6,
                      "layout": g.layout, "single_launch_sweeps": g.sweeps_single_launch()}), flush=True)
    g.close()
