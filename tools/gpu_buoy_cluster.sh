set -u
run() { local name=$1 n=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" > gpurun_out/r02s_$name.json 2> gpurun_out/r02s_$name.err
  echo "$name rc=$?"; python -c "
import json
d=json.load(open('gpurun_out/r02s_$name.json')); pc=d.get('parity_check') or {}
print('  ',d['n_gpus'],round(d['ms_per_step'],3),round(d['value'],2),pc.get('bitwise'),d['config'].get('parallelism','')[:70])"; }
CFD_B200_CLUSTER=1 run buoy_n8_cluster 8 --workload buoyancy --steps 30 --warmup 5
CFD_B200_CLUSTER=1 run buoy_n4_cluster 4 --workload buoyancy --steps 30 --warmup 5
run buoy_n4 4 --workload buoyancy --steps 30 --warmup 5
