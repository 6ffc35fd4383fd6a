#!/bin/bash
# multi-GPU bench lines: cavity i-slabs / j-slabs, backward step, heat; N = number of visible GPUs (or $1)
N=${1:-$(nvidia-smi -L | wc -l)}
mkdir -p gpurun_out
run() { # name, args...
  name=$1; shift
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) bench.py --gpus $N "$@" 2> gpurun_out/multi_${name}_$N.err > gpurun_out/multi_${name}_$N.json
  echo "$name N=$N rc=$?"; python - <<PY
import json
try:
    d = json.load(open("gpurun_out/multi_${name}_$N.json"))
    print({k: d.get(k) for k in ("value", "unit", "ms_per_step", "n_gpus", "scaling")}, d["config"].get("global_grid", ""), "e2e", (d.get("e2e") or {}).get("value"))
except Exception as e:
    print("no json:", e); print(open("gpurun_out/multi_${name}_$N.err").read()[-1500:])
PY
}
run cavity_i --steps 200 --warmup 10
run cavity_j --slabs j --steps 100 --warmup 10
run step --workload step --steps 50 --warmup 5
run heat --workload heat --steps 100 --warmup 10
run cavity16k --workload cavity16k --steps 20 --warmup 3
