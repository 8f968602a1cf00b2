python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for rs in 0 1 2 3; do echo "rep_shift $rs"; DSM_REP_SHIFT=$rs python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value']/1e12, d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value']/1e12)"; done
DSM_LIB=dstack_b200/libdsmatch_dbg.so python tools/cta_times.py 2>&1 | head -7
