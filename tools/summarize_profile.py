"""Summarise gpurun_out/launches.csv and gpurun_out/prof_*.ncu-rep into profiles/ (tracked).
Usage: python tools/summarize_profile.py r01 [gpurun_out/prof_r01.ncu-rep]"""
import collections, csv, json, os, re, subprocess, sys

tag = sys.argv[1]
rep = sys.argv[2] if len(sys.argv) > 2 else f"gpurun_out/prof_{tag}.ncu-rep"
os.makedirs("profiles", exist_ok=True)

# ---- launch list ---------------------------------------------------------------------------------------
lines = [l for l in open("gpurun_out/launches.csv") if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    k = re.sub(r"\(.*", "", row["Kernel Name"])
    agg.setdefault(k, []).append(float(row["Metric Value"].replace(",", "")) / 1e3)
tot = sum(sum(v) for v in agg.values())
with open(f"profiles/{tag}_launches.md", "w") as f:
    f.write(f"# {tag}: ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`) of\n"
            "`python bench.py --steps 2 --warmup 3 --no-cpu-baseline` (cold-cache, serialised: compare SHARES)\n\n")
    f.write("| kernel | launches | mean us | min us | max us | share of GPU time |\n|---|---|---|---|---|---|\n")
    for k, v in agg.items():
        f.write(f"| `{k}` | {len(v)} | {sum(v)/len(v):.1f} | {min(v):.1f} | {max(v):.1f} | {sum(v)/tot*100:.1f} % |\n")
    f.write(f"\ntotal kernel time in the list: {tot/1e3:.2f} ms over {sum(len(v) for v in agg.values())} launches\n")
subprocess.run(["cp", "gpurun_out/launches.csv", f"profiles/{tag}_launches.csv"], check=True)

# ---- full capture -------------------------------------------------------------------------------------------
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "launch__grid_size", "launch__block_size", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
            "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_short_scoreboard",
            "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_wait",
            "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle", "smsp__pcsamp_warps_issue_stalled_lg_throttle",
            "smsp__pcsamp_warps_issue_stalled_selected", "smsp__pcsamp_warps_issue_stalled_membar",
            "smsp__pcsamp_warps_issue_stalled_sleeping", "smsp__pcsamp_warps_issue_stalled_branch_resolving"]
    seen, traffic = {}, {}
    with open(f"profiles/{tag}_ncu_summary.md", "w") as f:
        f.write(f"# {tag}: `ncu --set full --clock-control none --import-source on` (one capture per kernel kept)\n\n")
        for r in rows[2:]:
            name = re.sub(r"\(.*", "", r[idx["Kernel Name"]])
            if name in seen:
                continue
            seen[name] = 1
            f.write(f"## `{name}`\n\n| metric | value | unit |\n|---|---|---|\n")
            for w in want:
                if w in idx:
                    f.write(f"| {w} | {r[idx[w]]} | {units[idx[w]]} |\n")
            f.write("\n")

            def val(m):
                v, u = float(r[idx[m]].replace(",", "")), units[idx[m]]
                return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
            key = "k_tdma_rb" if "tdma_tma" in name or "tdma_rb" in name else name.replace("cfd::", "")
            traffic[key] = {"dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
                            "kernel": name, "gpu_time_us": float(r[idx["gpu__time_duration.sum"]])}
    json.dump(traffic, open("profiles/traffic.json", "w"), indent=1)
    print(json.dumps(traffic, indent=1))
