python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for u in 1 2 4 8; do echo "units/cta $u"; DSM_UNITS_PER_CTA=$u python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value']/1e12, d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value']/1e12)"; done
DSM_LIB=dstack_b200/libdsmatch_dbg.so python tools/cta_times.py 2>&1 | tail -5
