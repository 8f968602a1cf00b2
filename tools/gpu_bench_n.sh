#!/bin/bash
# usage: gpu_run14.sh N
cd "$GRAFT_REPO_ROOT"
N=$1
mkdir -p gpurun_out
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 ) > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
tail -4 gpurun_out/r2_bench_n$N.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_n$N.json').read().strip().splitlines() if l.startswith('{')][-1])
print(json.dumps({k:d[k] for k in ('n_gpus','value','ms_per_step','e2e','step_breakdown_ms','parity')}, indent=1))
print(d['roofline']['kernel_ms_per_rank'], d['roofline']['hbm_fraction'])
PY
