#!/bin/bash
# WIDE kernel variant: parity tests, then A/B on the sphere-grid workloads (BENCH_ENGINE_OPTS=wide_min_records=-1 = narrow)
python -m pytest tests -m gpu -q 2>&1 | tail -3
for wl in spheres1024_4k spheres255_8k; do
  for o in "wide_min_records=64" "wide_min_records=-1"; do
    echo "== $wl $o"
    BENCH_ENGINE_OPTS=$o python bench.py --steps 3 --warmup 3 --no-cpu --workload $wl 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step', round(d['ms_per_step'],3), 'kernel_ms', round(d['roofline']['kernel_ms'],3), 'frac', round(d['roofline']['frac'],4), d['config']['launch'])"
  done
done
