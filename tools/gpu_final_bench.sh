#!/bin/bash
# the default bench line alone (the GPU tests are tools/gpu_final.sh)
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
( time timeout 1200 python bench.py ) > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err
tail -3 gpurun_out/final_bench.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/final_bench.json').read().strip().splitlines() if l.startswith('{')][-1])
print({k:d.get(k) for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['p50_batch_latency_ms'], d['e2e']['value'], d['e2e_api']['p50_batch_latency_ms'], d['step_breakdown_ms'])
r=d['roofline']; print({k:r.get(k) for k in ('frac','kernel_ms','traffic','lsu_fraction','issue_fraction','dram_fraction','traffic_source')})
a=d['strong_scaling_anchor']; print({k:a.get(k) for k in ('value','ms_per_step','kernel_ms','roofline_hbm_fraction','step_breakdown_ms')})
o=d['other_configs']; print(o['cfg1']['p50_ms_top50'], o['cfg2']['p50_ms_batched_match'], o['cfg5']['p50_ms_whole_pass'], [o[c]['parity'][:9] for c in o])
print(d['clocks'], d['cpu_baseline']['value'])
PY
