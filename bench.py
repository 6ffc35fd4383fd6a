#!/usr/bin/env python
"""bench.py -- the headline benchmark of BASELINE.json on B200: SIMPLE outer iterations of the lid-driven cavity
(Re=1000, 4096^2 control volumes per GPU, red/black line TDMA), reported as cell-updates/s (GLUPS) and
iterations/s, with the HBM roofline of the dominant kernel and a CPU baseline beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cavity|heat|step|buoyancy|cavity16k]

A "step" is one pass of the hot path: calcu + calcv + calcp (cavity_fvm.f90:554-557) over the whole grid.
For N>1 launch with torch.distributed.run (one rank per GPU); the grid is slab-decomposed along i (weak scaling:
4096 columns per GPU).  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NICV = NJCV = 4096          # BASELINE.json configs[2]
REYNOLDS = 1000.0
BYTES_PER_CELL_ITER = 928   # 116 doubles / cell / outer iteration (stored-coefficient algorithm)
SWEEP_BYTES_PER_CELL = 64   # one TDMA sweep: an,as,ae,aw,ap,su,phi read + phi written
HEAT_N = 16384              # BASELINE.json configs[1]
HEAT_BYTES_PER_CELL = 16


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock, power and throttle reasons of one GPU sampled WHILE the timed region runs.  NVML is polled from a
    thread every ~5 ms (a 20-step run lasts ~50 ms: an nvidia-smi loop at 200 ms would see it once); if NVML is not
    importable the nvidia-smi loop is the fallback."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []
        self.nvml, self.handle, self.samples, self._stop = None, None, [], threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                h = None
            if h is None:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                phys = int(vis.split(",")[index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else index
                h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml, self.handle = pynvml, h
        except Exception:
            self.nvml = None

    def start(self):
        if self.nvml is not None:
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n, h = self.nvml, self.handle
        while not self._stop.is_set():
            try:
                sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
                pw = n.nvmlDeviceGetPowerUsage(h) / 1000.0
                rs = n.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else n.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((sm, mx, pw, rs))
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.nvml is not None:
            self._stop.set()
            self.th.join(timeout=2)
            n = self.nvml
            bits = {"hw_slowdown": getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8),
                    "hw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                    "sw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                    "sw_power_cap": getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4)}
            sm = sorted(x[0] for x in self.samples)
            reasons = sorted(k for k, b in bits.items() if any(x[3] & b for x in self.samples))
            return {"sm_mhz": float(sm[len(sm) // 2]) if sm else None,
                    "sm_max_mhz": float(max(x[1] for x in self.samples)) if sm else None,
                    "power_w_max": max(x[2] for x in self.samples) if sm else None, "samples": len(sm),
                    "reasons": reasons, "how": "NVML polled every ~5 ms during the timed region"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons),
                "how": "nvidia-smi -lms 200"}


# ----------------------------------------------------------------------------------------------------------
def cpu_baseline_cavity(budget_s=20.0, nicv=NICV, njcv=NJCV):
    """The oracle (C restatement, -O3 -march=native, 1 thread like the serial reference) on a bounded sample:
    whole SIMPLE iterations of the same 4096^2 case until ~budget_s seconds have passed (at least 2)."""
    from oracle import oracle as orc
    c = orc.Cavity(nicv, njcv, REYNOLDS, fast=True)
    c.run(maxit=1, tol=0.0)                       # warm-up iteration (page faults, caches)
    n, t0 = 0, time.perf_counter()
    while True:
        c.run(maxit=1, tol=0.0)
        n += 1
        dt = time.perf_counter() - t0
        if (dt > budget_s and n >= 2) or n >= 50:
            break
    glups = nicv * njcv * n / dt / 1e9
    return {"value": glups, "unit": "GLUPS", "iters_per_s": n / dt, "cores": 1, "kind": "port",
            "host_cores_available": os.cpu_count(),
            "sample": f"{n} SIMPLE iterations of the same {nicv}x{njcv} Re={REYNOLDS:g} case after 1 warm-up, "
                      f"{dt:.1f} s, oracle/cavity_oracle.c gcc -O3 -march=native, 1 thread (the reference is serial; "
                      "no Fortran compiler on the box, so the reference's own build is unavailable)"}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (sequential line order) timed on the host cores.  The
    reference is serial Fortran and cannot be compiled here (no Fortran compiler), so the C restatement in
    oracle/ stands in ("port").  One step = one SIMPLE iteration; the grid is the full 4096^2 case when the
    whole run fits a few minutes, else a 1024^2 sample of it (cell-updates/s is the comparable unit)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as orc
    K, W = args.steps, args.warmup
    full = (K + W) * 6.5 <= 200.0
    n = NICV if full else 1024
    c = orc.Cavity(n, n, REYNOLDS, fast=True)
    for _ in range(W):
        c.run(maxit=1, tol=0.0)
    t0 = time.perf_counter()
    for _ in range(K):
        c.run(maxit=1, tol=0.0)
    dt = time.perf_counter() - t0
    glups = n * n * K / dt / 1e9
    sample = (f"{K} SIMPLE iterations on {'the full' if full else 'a'} {n}x{n} grid (Re={REYNOLDS:g}, hybrid, nswp 3/3/3), "
              "sequential line order, oracle/cavity_oracle.c gcc -O3 -march=native, 1 thread: the reference is a serial "
              "program and its Fortran sources cannot be built on this box (no Fortran compiler)")
    out = {"impl": "reference", "metric": "simple_cell_updates_per_s", "value": glups, "unit": "GLUPS",
           "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": dt / K * 1e3, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "iters_per_s_4096": glups * 1e9 / (NICV * NJCV),
           "config": {"workload": f"lid-driven cavity FVM SIMPLE Re=1000 {NICV}x{NJCV} CV per GPU (BASELINE.json configs[2])",
                      "grid_timed": f"{n}x{n}", "order": "sequential (reference)"},
           "cpu_baseline": {"value": glups, "unit": "GLUPS", "cores": 1, "kind": "port", "sample": sample,
                            "host_cores_available": os.cpu_count()},
           "e2e": {"value": glups, "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(out)



# ----------------------------------------------------------------------------------------------------------
def parity_check(kind, slabs="i", iters=6, turbulent=False, cluster=False):
    """N > 1 only, before the timed region: a small problem of the same family runs once slab-decomposed over all
    ranks (the very driver, kernels and exchange the timed run uses) and once in a single context on rank 0; the
    gathered u, v, p are compared.  i-slabs must agree BIT FOR BIT (red/black iterates do not depend on the partition);
    j-slabs reorder the elimination (partition method) and must agree to 1e-10.  Returns the record for the JSON line;
    a mismatch makes every rank exit non-zero.  Lines are long enough (>= 1026 nodes) for the production TMA-staged
    kernels to be the ones that run."""
    import numpy as np
    import torch
    import torch.distributed as dist

    import cfd_b200
    from cfd_b200.dist import BuoyancySlab, JSlabCavity, SlabCavity, TeachSlab
    world, rank = dist.get_world_size(), dist.get_rank()
    # lines of 2050 nodes: a layout (160 threads x 13 nodes) the flag-ordered sweep kernel takes, so that the check runs
    # over the same halo path (peer memory or NCCL) as the timed run
    # cluster=True: the timed run solves its lines with clusters of thread blocks (k_tdma_flowc); the check then uses lines
    # of 8194 nodes (640 layout threads, two blocks per line) and forces that solver in BOTH runs -- its three-column
    # reduction rounds differently from the one-block solver, the partition must not matter
    if cluster:
        os.environ["CFD_B200_CLUSTER"] = "1"
    if kind == "buoyancy":
        n = 2050
        shape = (n, n)
        s = SlabCavity(n - 2, n - 2, slab_cls=BuoyancySlab)
        single = (lambda: cfd_b200.BuoyantCavity(n, n, order="redblack")) if rank == 0 else None
    elif kind == "step":
        shape = (64 * world + 2, 2050)
        s = SlabCavity(shape[0] - 2, shape[1] - 2, slab_cls=TeachSlab, epsx=1.0, turbulent=turbulent)
        single = (lambda: cfd_b200.TeachFVM(shape[0], shape[1], epsx=1.0, order="redblack", turbulent=turbulent)) if rank == 0 else None
    else:
        shape = ((64 * world, 2048) if not cluster else (64 * world, 8192)) if slabs == "i" else (96, 256 * world)
        s = (SlabCavity if slabs == "i" else JSlabCavity)(*shape, reynolds=REYNOLDS)

        def single():
            g = cfd_b200.CavityFVM(*shape, order="redblack")
            g.init_lid_driven(reynolds=REYNOLDS)
            return g
    s.iterate(iters)
    halo_path = "peer memory" if getattr(s, "peer", False) else "NCCL"
    names = ("u", "v", "p") + (("t",) if kind == "buoyancy" else ()) + (("te", "ed", "t", "vis") if turbulent else ())
    got = {n: s.gather_field(n) for n in names}
    s.ctx.close()
    rec = None
    if rank == 0:
        g = single()
        g.iterate(iters)
        max_abs, bitwise, scale = 0.0, True, 0.0
        for n in names:
            want = g.download(n)
            bitwise = bitwise and bool(np.array_equal(got[n], want))
            max_abs = max(max_abs, float(np.abs(got[n] - want).max()))
            scale = max(scale, float(np.abs(want).max()))
        g.close()
        tol = 0.0 if slabs == "i" else 1e-10 * max(1.0, scale)
        rec = {"world": world, "case": f"{kind}{' (k-epsilon + energy)' if turbulent else ''} {shape[0]}x{shape[1]}, {iters} outer "
                                       f"iterations, {slabs}-slabs vs one context",
               "halo_path": halo_path, "fields": list(names), "bitwise": bitwise, "max_abs": max_abs, "tolerance": tol,
               "ok": bool(bitwise if slabs == "i" else max_abs <= tol)}
    if cluster:
        os.environ.pop("CFD_B200_CLUSTER", None)
        if rec is not None:
            rec["line_solver"] = "clusters of thread blocks (k_tdma_flowc), forced in both runs"
    flag = torch.tensor([1 if (rec is None or rec["ok"]) else 0], device="cuda")
    dist.broadcast(flag, 0)
    if int(flag.item()) != 1:
        if rank == 0:
            sys.stderr.write("bench.py: multi-GPU parity check FAILED: %s\n" % json.dumps(rec))
        dist.destroy_process_group()
        raise SystemExit(3)
    return rec

# ----------------------------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import cfd_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    K, W = args.steps, args.warmup
    hbm_peak, peak_src = peaks()
    stream = torch.cuda.Stream()
    parity = parity_check("cavity", args.slabs) if world > 1 else None

    if world == 1:
        solver = cfd_b200.CavityFVM(NICV, NJCV, order="redblack", device=local)
        solver.set_stream(stream.cuda_stream)
        solver.init_lid_driven(reynolds=REYNOLDS)
        step = solver.iterate_async
        sync_res = solver.residuals
        ctx0 = solver
        cells_total = NICV * NJCV
        par = "single GPU"
    else:
        from cfd_b200.dist import JSlabCavity, SlabCavity
        if args.slabs == "j":
            solver = JSlabCavity(NICV, NJCV * world, reynolds=REYNOLDS, stream=stream)
        else:
            solver = SlabCavity(NICV * world, NJCV, reynolds=REYNOLDS, stream=stream)
        if args.slabs == "j":
            graphed = False
        elif os.environ.get("CFD_B200_GRAPH", "0") != "0":     # off by default: no gain measured, and teardown is slow
            with torch.cuda.stream(stream):
                graphed = solver.enable_graph(stream)
        else:
            graphed = False
        step = solver.iterate_async
        sync_res = solver.residuals
        ctx0 = solver.ctx
        cells_total = NICV * world * NJCV
        if getattr(solver, "peer", False):
            par = (f"i-slabs x{world}; halo lines over peer memory (NVLink): edge lines stored into the neighbour's halo from "
                   "inside the flag-ordered sweep kernels, cfd_halo_push between stages, pp(ipref,jpref) by peer stores from "
                   "its owner; residual all-reduce over NCCL when the host asks")
        else:
            par = (f"i-slabs x{world}, 1-line halo exchange (NCCL send/recv), all-reduced residuals, "
                   + ("one CUDA graph per outer iteration" if graphed else "eager launches"))
        if args.slabs == "j":
            par = (f"j-slabs x{world}: lines cut across GPUs, partitioned Thomas (piece solves + spike vectors, "
                   "all-gathered reduced system per line), halo rows once per iteration, all-reduced residuals")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(W):
            step(1)
        barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        l0 = ctx0.launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(K):
            step(1)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        launches = ctx0.launches - l0
        clocks = sampler.stop() if rank == 0 else None
        res = sync_res()
    if world > 1:
        tms = torch.tensor([ms, 1.0 if ctx0.peer_status() else 0.0], device="cuda", dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms[0].item())
        if float(tms[1].item()) != 0.0:
            raise SystemExit("bench.py: a wait on a neighbouring GPU timed out (peer halo path); the run is invalid")
    glups = cells_total * K / (ms * 1e-3) / 1e9

    out = {"metric": "simple_cell_updates_per_s", "value": glups, "unit": "GLUPS", "n_gpus": world, "steps": K,
           "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "iters_per_s": K / (ms * 1e-3),
           "config": {"workload": f"lid-driven cavity FVM SIMPLE Re=1000, {NICV}x{NJCV} CV per GPU, hybrid scheme, "
                                  "urf .5/.5/.8, nswp 3/3/3, red/black line TDMA (BASELINE.json configs[2])",
                      "global_grid": (f"{NICV}x{NJCV * world}" if (world > 1 and args.slabs == "j") else f"{NICV * world}x{NJCV}"),
                      "parallelism": par,
                      "l2": "inputs larger than L2 (1.9 GB of live fields per GPU vs 126 MB L2); no explicit flush",
                      "line_layout": ctx0.layout},
           "residuals_last": list(res[:3]), "gpu_launches": int(launches)}
    if parity is not None:
        out["parity_check"] = parity

    # ---- roofline of the dominant kernel (the red/black line solver), timed live with CUDA events ---------
    # the launch the step itself makes: all 2*nswp colour passes of one equation (cfd_sweeps)
    ctxr = ctx0
    nsw = 3
    one_launch = ctxr.sweeps_single_launch()
    with torch.cuda.stream(stream):
        ctxr.assemble_u()
        ctxr.sweeps(3, 2, "u", nsw)                       # untimed: first use of a kernel loads its module
        if world > 1:
            barrier()
        evs = []
        for _ in range(5):
            ctxr.assemble_u()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            ctxr.sweeps(3, 2, "u", nsw)
            b.record(stream)
            evs.append((a, b))
        torch.cuda.synchronize()
    t_launch = sum(a.elapsed_time(b) for a, b in evs) / len(evs) * 1e-3
    cells_launch = ((NICV - 1) if world == 1 else NICV) * NJCV * nsw   # u-lines i=3..ni-1, both colours, nsw sweeps
    achieved = SWEEP_BYTES_PER_CELL * cells_launch / t_launch / 1e9
    kname = "k_tdma_flow" if one_launch else "k_tdma_tma"
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get(kname, {}).get("dram_bytes_per_launch")
            if traffic is not None and not one_launch:
                traffic *= 2 * nsw                         # the file holds one colour pass of k_tdma_tma
        except Exception:
            traffic = None
    out["roofline"] = {"bound": "hbm",
                       "kernel": (f"{kname} (the {2 * nsw} colour passes of one equation's line sweeps"
                                  + (", one flag-ordered launch)" if one_launch else ", one launch per pass)")),
                       "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                       "peak_source": peak_src, "traffic": traffic, "launch_ms": t_launch * 1e3,
                       "algorithmic_bytes_per_launch": SWEEP_BYTES_PER_CELL * cells_launch,
                       "note": "algorithmic = 8 doubles/cell and sweep (6 coefficients + phi read, phi written; SURVEY 8d); the "
                               "kernel itself moves 7 (equations are stored divided by ap), so frac can exceed the copy peak",
                       "whole_iteration": {"algorithmic_bytes_per_cell": BYTES_PER_CELL_ITER,
                                           "achieved": BYTES_PER_CELL_ITER * glups / world,
                                           "frac": BYTES_PER_CELL_ITER * glups / world / hbm_peak,
                                           "frac_of_nominal_8TBs": BYTES_PER_CELL_ITER * glups / world / 8000.0}}
    if world == 1:
        # ---- e2e: the drop-in call sequence with HOST arrays (pinned), copies inside the timed region --------
        ni, nj = solver.ni, solver.nj
        host = {n: torch.empty((nj, ni), dtype=torch.float64, pin_memory=True) for n in ("u", "v", "p")}
        for n in host:
            solver.download_ptr(n, host[n].data_ptr(), ni)
        Ke = max(3, min(K, 10))
        fb = ni * nj * 8
        imon = jmon = 6                                       # the reference's monitor node (cavity_fvm.f90:528-529)

        def e2e_leg(full_download):
            torch.cuda.synchronize()
            with torch.cuda.stream(stream):
                for it in range(1 + Ke):
                    if it == 1:
                        torch.cuda.synchronize()
                        t0 = time.perf_counter()
                    for n in host:
                        solver.upload_ptr(n, host[n].data_ptr(), ni)
                    ru, rv, rm = solver.calcu(), solver.calcv(), solver.calcp()
                    if full_download:
                        for n in host:
                            solver.download_ptr(n, host[n].data_ptr(), ni)
                    else:
                        mon = [solver.probe(n, imon, jmon) for n in ("u", "v", "p")]
                torch.cuda.synchronize()
                return time.perf_counter() - t0

        dt = e2e_leg(False)
        dt_full = e2e_leg(True)
        out["e2e"] = {"value": NICV * NJCV * Ke / dt / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": 3 * fb,
                      "d2h_bytes_per_step": 48, "steps": Ke, "ms_per_step": dt / Ke * 1e3,
                      "what": "per step: cfd_upload(u,v,p) from pinned host arrays, cfd_calcu, cfd_calcv, cfd_calcp (each "
                              "returns its residual to the host), cfd_probe of u,v,p at the monitor node -- what the "
                              "reference's loop shows the host per iteration (cavity_fvm.f90:555-563)",
                      "roundtrip": {"value": NICV * NJCV * Ke / dt_full / 1e9, "unit": "GLUPS",
                                    "h2d_bytes_per_step": 3 * fb, "d2h_bytes_per_step": 3 * fb + 24,
                                    "ms_per_step": dt_full / Ke * 1e3,
                                    "what": "same, but cfd_download(u,v,p) of the whole fields every step"}}
    else:
        out["e2e"] = solver.e2e(max(3, min(K, 10)))

    out["clocks"] = clocks
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline_cavity(args.cpu_budget)
    if rank == 0:
        emit(out)
    if world > 1:
        dist.destroy_process_group()


def run_heat(args):
    """--workload heat: BASELINE.json configs[1], 2-D FTCS 16384^2 FP64 (one step = one explicit time step)."""
    import numpy as np
    import torch

    import cfd_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    K, W = args.steps, args.warmup
    hbm_peak, peak_src = peaks()
    n = HEAT_N
    stream = torch.cuda.Stream()
    if world > 1:
        # weak scaling: 16384 x 16384 rows per GPU, row slabs with one halo row per side and step (NCCL send/recv)
        from cfd_b200.dist import SlabHeat, implicit_halo_rows
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        implicit = bool(getattr(args, "implicit", False))
        if implicit:      # Fo = 5 (25x the explicit limit): 128 halo rows per inner side, refreshed every step
            halo = implicit_halo_rows(n, n * world, fo=5.0, yb=float(world))
            sh = SlabHeat(n, n * world, stream=stream, halo=halo, fo=5.0, yb=float(world))
            step = sh.implicit_step
        else:
            sh = SlabHeat(n, n * world, stream=stream)
            step = sh.step
        with torch.cuda.stream(stream):
            step(W)
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            sampler = ClockSampler(local)
            if rank == 0:
                sampler.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            step(K)
            e1.record(stream)
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            clocks = sampler.stop() if rank == 0 else None
        tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
        glups = n * n * world * K / (ms * 1e-3) / 1e9
        bpc = 64 if implicit else HEAT_BYTES_PER_CELL        # LOD: four passes x (read + write)
        out = {"metric": "lod_cell_updates_per_s" if implicit else "ftcs_cell_updates_per_s", "value": glups, "unit": "GLUPS",
               "n_gpus": world, "steps": K, "warmup": W,
               "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "steps_per_s": K / (ms * 1e-3),
               "config": {"workload": (f"2-D implicit (LOD) heat conduction, {n}x{n} nodes per GPU, FP64, Fo=5" if implicit else
                                       f"2-D FTCS heat conduction, {n}x{n} nodes per GPU, FP64, Fo=0.2 (BASELINE.json configs[1])"),
                          "global_grid": f"{n}x{n * world}",
                          "parallelism": (f"row slabs x{world}, {sh.slab.hs or sh.slab.hn} halo rows per inner side refreshed every step "
                                          "(NCCL send/recv)" if implicit else
                                          f"row slabs x{world}, one halo row per side and step (NCCL send/recv)"),
                          "l2": "field (2.1 GB per GPU) larger than L2; ping-pong buffers"},
               "gpu_launches": K * (5 if implicit else 1),
               "roofline": {"bound": "hbm", "kernel": "k_lod_x/y_chunk (four sweeps)" if implicit else "k_ftcs<false>",
                            "achieved": bpc * glups / world,
                            "peak": hbm_peak, "unit": "GB/s", "frac": bpc * glups / world / hbm_peak,
                            "peak_source": peak_src, "traffic": None},
               "e2e": None, "clocks": clocks}
        if rank == 0:
            emit(out)
        dist.destroy_process_group()
        return
    h = cfd_b200.Heat2D(n, n, device=local)
    h.set_stream(stream.cuda_stream)
    host = torch.zeros((n, n), dtype=torch.float64, pin_memory=True)
    host.view(-1)[:: 7] = 1.0
    h.upload_ptr(host.data_ptr(), n)
    sampler = ClockSampler(0)
    with torch.cuda.stream(stream):
        h.step(W)
        torch.cuda.synchronize()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        h.step(K)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        clocks = sampler.stop()
        Ke = 3
        t0 = time.perf_counter()
        for _ in range(Ke):
            h.upload_ptr(host.data_ptr(), n)
            h.step(1)
            h.download_ptr(host.data_ptr(), n)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    glups = n * n * K / (ms * 1e-3) / 1e9
    achieved = HEAT_BYTES_PER_CELL * glups
    from oracle import oracle as orc
    m = 4096
    t_in = np.asfortranarray(np.random.default_rng(1).random((m, m)))
    tc = time.perf_counter()
    nst = 0
    while time.perf_counter() - tc < args.cpu_budget / 2 or nst < 2:
        t_in = orc.heat2d_ftcs(t_in, 2, fast=True)
        nst += 2
    tcpu = time.perf_counter() - tc
    out = {"metric": "ftcs_cell_updates_per_s", "value": glups, "unit": "GLUPS", "n_gpus": 1, "steps": K, "warmup": W,
           "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic", "steps_per_s": K / (ms * 1e-3),
           "config": {"workload": f"transient heat conduction 2D, explicit FTCS 5-point, {n}x{n} FP64, Fo=0.2, source q=1e3 "
                                  "where x<0.1 (BASELINE.json configs[1])",
                      "l2": "two 2.1 GB ping-pong buffers, far larger than L2"},
           "gpu_launches": K,
           "roofline": {"bound": "hbm", "kernel": "k_ftcs<false>", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                        "frac": achieved / hbm_peak, "peak_source": peak_src, "traffic": None,
                        "algorithmic_bytes_per_launch": HEAT_BYTES_PER_CELL * n * n, "launch_ms": ms / K},
           "e2e": {"value": n * n * Ke / dt / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": n * n * 8,
                   "d2h_bytes_per_step": n * n * 8, "steps": Ke,
                   "what": "per step: cfd_heat_upload from pinned host memory, one FTCS step, cfd_heat_download"},
           "clocks": clocks,
           "cpu_baseline": {"value": m * m * nst / tcpu / 1e9, "unit": "GLUPS", "cores": 1, "kind": "port",
                            "sample": f"{nst} FTCS steps on a {m}x{m} sample grid, oracle/heat_oracle.c gcc -O3 -march=native, 1 thread"}}
    emit(out)


def run_fdcavity(args):
    """--workload fdcavity: the explicit finite-difference cavity (fdm_explicit_cavity.f90:168-255) on 4097 x 4097 nodes,
    one GPU.  A step = one time step: velocity predictor, p' source, the SIP solve of the p' equation (sipsol :24-87, up
    to 600 inner iterations, all inside one persistent kernel), correction + boundary values + continuity norm.  The
    SIP solve is >99 % of a step, so the metric is quoted per inner iteration as well."""
    import numpy as np
    import torch

    import cfd_b200
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(0)
    hbm_peak, peak_src = peaks()
    n = 4097
    K, W = max(1, min(args.steps, 5)), max(3, min(args.warmup, 3))
    g = cfd_b200.FDCavity(n, n, re=100.0, dt=1.0e-4)
    for _ in range(W):
        g.step()
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    its = 0
    e0.record()
    for _ in range(K):
        g.step()
        its += g.sip_iters
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    cells = (n - 1) * (n - 1)
    # end to end: host arrays in, one step, the continuity norm out
    host = {f: np.asfortranarray(g.download(f)) for f in ("uo", "vo", "po")}
    Ke = 2
    t0 = time.perf_counter()
    for _ in range(Ke):
        for f, a in host.items():
            g.upload(f, a)
        g.step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    sip_bytes = 17 * 8                                          # 17 doubles per cell and inner iteration
    t_it = ms / its                                             # the stencil passes are ~1 ms of a step: negligible
    achieved = sip_bytes * cells / (t_it * 1e-3) / 1e9
    out = {"metric": "fdcavity_cell_updates_per_s", "value": cells * K / (ms * 1e-3) / 1e9, "unit": "GLUPS", "n_gpus": 1,
           "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "config": {"workload": f"explicit FD lid-driven cavity {n}x{n} nodes, Re=100, dt=1e-4, SIP alpha 0.88 tol 1e-5 maxit 600 "
                                  "(fdm_explicit_cavity.f90; SURVEY 8 rows a12 + f-1)",
                      "l2": "13 arrays of 134 MB, far larger than L2", "sip_inner_iterations_per_step": its / K,
                      "ms_per_sip_iteration": t_it},
           "gpu_launches": K * 8,
           "roofline": {"bound": "hbm", "kernel": "k_sip_solve (one inner iteration: residual pass, forward and backward "
                                                  "substitution wavefronts)", "achieved": achieved, "peak": hbm_peak,
                        "unit": "GB/s", "frac": achieved / hbm_peak, "peak_source": peak_src, "traffic": None,
                        "algorithmic_bytes_per_launch": sip_bytes * cells * its / K, "launch_ms": ms / K,
                        "note": "a wavefront recurrence in both grid directions: bounded by the dependent chain of "
                                "(imax + jmax) cells per substitution, not by bandwidth (DESIGN.md section 4)"},
           "e2e": {"value": cells * Ke / dt / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": 3 * (n + 1) * (n + 1) * 8,
                   "d2h_bytes_per_step": 8, "steps": Ke,
                   "what": "per step: uo, vo, po uploaded from host arrays, cfd_fdcav_step, the continuity norm back"},
           "clocks": clocks}
    if not args.no_cpu_baseline:
        from oracle import oracle as orc
        m = 513
        o = orc.FDCavity(m, m, re=100.0, dt=1.0e-4, fast=True)
        tc = time.perf_counter()
        nst, oits = 0, 0
        while time.perf_counter() - tc < args.cpu_budget or nst < 1:
            o.step()
            oits += o.s.sip_iters
            nst += 1
        tcpu = time.perf_counter() - tc
        out["cpu_baseline"] = {"value": (m - 1) * (m - 1) * nst / tcpu / 1e9, "unit": "GLUPS", "cores": 1, "kind": "port",
                               "sample": f"{nst} time steps ({oits} SIP iterations) on a {m}x{m} sample grid, "
                                         "oracle/fdcavity_oracle.c gcc -O3 -march=native, 1 thread",
                               "ms_per_sip_iteration_scaled_to_4097": tcpu / max(oits, 1) * 1e3 * cells / ((m - 1) * (m - 1))}
    emit(out)
    g.close()


def run_cavity16k(args):
    """--workload cavity16k: north_star's "8 GPUs on 16384^2" case, STRONG scaling (the global grid is fixed).  Default:
    i-slabs -- every GPU owns 16384/N whole lines of 16386 nodes, solved by clusters of four thread blocks
    (tdma_cluster_kernels.cu), halo lines over peer memory.  --slabs j: the lines themselves are cut across the GPUs and
    solved with the partition method (N = 1 then runs two slabs on the one GPU)."""
    import torch
    import torch.distributed as dist

    import cfd_b200
    from cfd_b200.dist import JSlabCavity, LocalJSlabCavity, SlabCavity
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    K, W = args.steps, args.warmup
    hbm_peak, peak_src = peaks()
    NG = 16384
    stream = torch.cuda.current_stream()
    parity, jcut = None, args.slabs == "j"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        parity = parity_check("cavity", "j" if jcut else "i", cluster=not jcut)
        solver = JSlabCavity(NG, NG, reynolds=REYNOLDS) if jcut else SlabCavity(NG, NG, reynolds=REYNOLDS)
        step, ctx0 = solver.iterate_async, solver.ctx
    elif jcut:
        solver = LocalJSlabCavity(NG, NG, 2, reynolds=REYNOLDS)
        step, ctx0 = (lambda n: [solver._iterate_once() for _ in range(n)]), solver.slabs[0].ctx
    else:
        solver = cfd_b200.CavityFVM(NG, NG, order="redblack", device=local)
        solver.init_lid_driven(reynolds=REYNOLDS)
        step, ctx0 = solver.iterate_async, solver

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(W):
        step(1)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx0.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(K):
        step(1)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = (ctx0.launches - l0) * (2 if (world == 1 and jcut) else 1)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        tms = torch.tensor([ms, 1.0 if ctx0.peer_status() else 0.0], device="cuda", dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms[0].item())
        if float(tms[1].item()) != 0.0:
            raise SystemExit("bench.py: a wait on a neighbouring GPU timed out (peer halo path); the run is invalid")
        res = solver.residuals()
    elif jcut:
        res = tuple(sum(s.res.clone() for s in solver.slabs).cpu().tolist())
    else:
        res = solver.residuals()
    glups = NG * NG * K / (ms * 1e-3) / 1e9
    if jcut:
        par = (f"j-slabs x{world}: piece solves + spike vectors, all-gathered reduced system per line (NCCL), halo rows once "
               "per iteration" if world > 1 else "single GPU, two j-slabs in one process (same algorithm, copies instead of messages)")
    elif world > 1:
        par = (f"i-slabs x{world}: {NG // world} whole lines of {NG + 2} nodes per GPU, each line solved by a cluster of thread "
               "blocks; halo lines " + ("over peer memory (stores from inside the sweep kernels, cfd_halo_push)"
                                        if getattr(solver, "peer", False) else "by NCCL send/recv"))
    else:
        par = "single GPU, whole lines solved by clusters of thread blocks"
    out = {"metric": "simple_cell_updates_per_s", "value": glups, "unit": "GLUPS", "n_gpus": world, "steps": K,
           "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic", "iters_per_s": K / (ms * 1e-3),
           "config": {"workload": f"lid-driven cavity FVM SIMPLE Re=1000, {NG}x{NG} CV in total, hybrid scheme, urf .5/.5/.8, "
                                  "nswp 3/3/3, red/black lines",
                      "parallelism": par, "l2": "inputs larger than L2; no explicit flush", "line_layout": ctx0.layout},
           "residuals_last": [float(x) for x in res[:3]], "gpu_launches": int(launches),
           "roofline": {"bound": "hbm", "kernel": "whole outer iteration", "achieved": BYTES_PER_CELL_ITER * glups / world,
                        "peak": hbm_peak, "unit": "GB/s", "frac": BYTES_PER_CELL_ITER * glups / world / hbm_peak,
                        "peak_source": peak_src, "traffic": None,
                        "note": "algorithmic bytes of the stored-coefficient algorithm (928 B/cell, SURVEY 8d)"},
           "e2e": None, "clocks": clocks}
    if parity is not None:
        out["parity_check"] = parity
    if rank == 0:
        emit(out)
    if world > 1:
        dist.destroy_process_group()


def run_step(args, kind="step"):
    """--workload step: BASELINE.json configs[4], TEACH backward-step SIMPLE (laminar, nswp 3/3/5) on 16384 x 4096,
    i-slabs over the N GPUs (STRONG scaling: the global grid is fixed).  One step = one outer iteration."""
    import torch
    import torch.distributed as dist

    import cfd_b200
    from cfd_b200.dist import BuoyancySlab, SlabCavity, TeachSlab
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    K, W = args.steps, args.warmup
    hbm_peak, peak_src = peaks()
    NI, NJ = (16384, 4096) if kind == "step" else (8194, 8194)
    stream = torch.cuda.Stream()
    turb = bool(getattr(args, "turbulent", False)) and kind == "step"
    # (buoyancy 8192^2 on >= 3 GPUs: slabs of <= 3072 lines take the cluster line solver, tdma_flow_kernels.cu:
    # cfd_flow_eligible; the general family needs square grids, so its small parity case cannot have 8194-node lines --
    # the cluster solver over peer memory is checked by --workload cavity16k's parity case and tests/test_gpu_dist.py)
    parity = parity_check(kind, turbulent=turb) if world > 1 else None
    if world == 1:
        if kind == "step":
            solver = cfd_b200.TeachFVM(NI, NJ, epsx=1.0, order="redblack", device=local, turbulent=turb)
        else:
            solver = cfd_b200.BuoyantCavity(NI, NJ, order="redblack", device=local)
        solver.set_stream(stream.cuda_stream)
        step, ctx0 = solver.iterate_async, solver
    elif kind == "step":
        solver = SlabCavity(NI - 2, NJ - 2, stream=stream, slab_cls=TeachSlab, epsx=1.0, turbulent=turb)
        step, ctx0 = solver.iterate_async, solver.ctx
    else:
        solver = SlabCavity(NI - 2, NJ - 2, stream=stream, slab_cls=BuoyancySlab)
        step, ctx0 = solver.iterate_async, solver.ctx

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        for _ in range(W):
            step(1)
        barrier()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        l0 = ctx0.launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(K):
            step(1)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        launches = ctx0.launches - l0
        clocks = sampler.stop() if rank == 0 else None
        res = solver.residuals()
    if world > 1:
        tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    cells = (NI - 2) * (NJ - 2)
    glups = cells * K / (ms * 1e-3) / 1e9
    B = (1992 if turb else 1056) if kind == "step" else 1224   # 132 (249 as shipped) / 153 doubles per cell and outer iteration
    wl = (f"backward-facing step FVM SIMPLE (TEACH family, "
          f"{'as shipped: k-epsilon + energy + props, control.f90:210-217' if turb else 'laminar'}, axisymmetric, hybrid scheme, "
          f"urf .5/.5/1, nswp 3/3/5{' + 3/3/3' if turb else ''}) "
          f"on {NI}x{NJ}, red/black line TDMA (BASELINE.json configs[4])") if kind == "step" else (
          f"buoyancy-driven cavity Ra=1e6 (Boussinesq, coupled T equation) SIMPLE on {NI - 2}^2 CV, hybrid scheme, urf "
          ".3/.3/.9/1, nsweep 3/3/3/3, red/black line TDMA (BASELINE.json configs[3])")
    out = {"metric": "simple_cell_updates_per_s", "value": glups, "unit": "GLUPS", "n_gpus": world, "steps": K,
           "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic", "iters_per_s": K / (ms * 1e-3),
           "config": {"workload": wl,
                      "parallelism": "single GPU" if world == 1 else (
                          f"i-slabs x{world}; halo lines and pp(ipref,jpref) over peer memory (NVLink), all-reduced residuals"
                          if getattr(solver, "peer", False) else
                          f"i-slabs x{world}, 1-line halo exchange (NCCL send/recv), all-reduced residuals"),
                      "l2": "inputs larger than L2 (several GB of live fields); no explicit flush",
                      "line_layout": ctx0.layout},
           "residuals_last": [float(x) for x in res[:6 if turb else 3]], "gpu_launches": int(launches),
           "roofline": {"bound": "hbm", "kernel": "whole outer iteration", "achieved": B * glups / world, "peak": hbm_peak,
                        "unit": "GB/s", "frac": B * glups / world / hbm_peak, "peak_source": peak_src, "traffic": None,
                        "algorithmic_bytes_per_cell": B},
           "e2e": None, "clocks": clocks}
    if parity is not None:
        out["parity_check"] = parity
    if rank == 0:
        emit(out)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def _capture_stdout():
    """Rank 0 must print exactly ONE JSON line on stdout.  Libraries (NCCL prints its version banner) write to
    fd 1 directly, so fd 1 is pointed at stderr for the whole run and the JSON line goes to the saved fd."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, line)
    else:
        sys.stdout.write(line.decode())
        sys.stdout.flush()


def main():
    _capture_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cavity", choices=["cavity", "heat", "step", "cavity16k", "buoyancy", "fdcavity"])
    ap.add_argument("--slabs", default="i", choices=["i", "j"], help="N>1 cavity: cut along i (whole lines per GPU) or j")
    ap.add_argument("--turbulent", action="store_true", help="--workload step: the case as shipped (k-epsilon + energy)")
    ap.add_argument("--implicit", action="store_true", help="--workload heat at N > 1: implicit (LOD) steps on row slabs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "heat":
        run_heat(args)
    elif args.workload == "step":
        run_step(args)
    elif args.workload == "buoyancy":
        run_step(args, "buoyancy")
    elif args.workload == "fdcavity":
        run_fdcavity(args)
    elif args.workload == "cavity16k":
        run_cavity16k(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
