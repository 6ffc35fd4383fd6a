#!/bin/bash
# Round-2 scaling evidence on an 8-GPU box (under gpurun --gpus 8): the default cavity workload at N = 2, 4, 8 (weak),
# and the fixed-size workloads at N = 8 (strong).  One JSON line per run into gpurun_out/r02s_*.json.
set -u
mkdir -p gpurun_out
run() {  # name nproc args...
  local name=$1 n=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) \
    bench.py --gpus $n "$@" > gpurun_out/r02s_$name.json 2> gpurun_out/r02s_$name.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/r02s_$name.json"))
    pc = d.get("parity_check") or {}
    print("  ", d["n_gpus"], "GPUs", round(d["ms_per_step"], 3), "ms", round(d["value"], 2), d["unit"], "| parity", pc.get("halo_path"), pc.get("bitwise"), "|", d["config"].get("parallelism", "")[:60])
except Exception as e:
    print("   no JSON line:", e)
PY
}
run cav_n2 2 --steps 200 --warmup 20 --no-cpu-baseline
run cav_n4 4 --steps 200 --warmup 20 --no-cpu-baseline
run cav_n8 8 --steps 200 --warmup 20 --no-cpu-baseline
run c16k_n8 8 --workload cavity16k --steps 30 --warmup 5 --no-cpu-baseline
run step_n8 8 --workload step --steps 30 --warmup 5
run stepturb_n8 8 --workload step --turbulent --steps 30 --warmup 5
run buoy_n8 8 --workload buoyancy --steps 30 --warmup 5
run heatlod_n8 8 --workload heat --implicit --steps 20 --warmup 3
run heat_n8 8 --workload heat --steps 100 --warmup 5
ls -la gpurun_out/r02s_* | head -20
