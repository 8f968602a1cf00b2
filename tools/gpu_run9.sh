#!/bin/bash
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 ) > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
tail -5 gpurun_out/r2_bench_n2.err
python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2_bench_n2.json').read().strip().splitlines() if l.startswith('{')][-1])
    print(json.dumps({k:d[k] for k in ('value','ms_per_step','e2e','step_breakdown_ms','parity','n_gpus')}, indent=1))
    print(d['roofline']['kernel_ms_per_rank'], d['roofline']['hbm_fraction'])
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/r2_bench_n2.json').read()[-3000:])
PY
