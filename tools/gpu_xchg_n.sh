#!/bin/bash
# usage: tools/gpu_xchg_n.sh N [bench]   -- exchange micro-benchmark on N GPUs, then (optionally) bench.py
cd "$GRAFT_REPO_ROOT"
N=$1
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/xchg_bench.py > gpurun_out/xchg_n$N.json 2> gpurun_out/xchg_n$N.err
echo "xchg rc=$?"; tail -3 gpurun_out/xchg_n$N.err; cat gpurun_out/xchg_n$N.json
if [ "$2" = "bench" ]; then bash tools/gpu_bench_n.sh $N; fi
