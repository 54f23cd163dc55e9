#!/bin/bash
# A/B: same bench, alternative builds of the library (RTB_LIB)
for lib in "$@"; do
  echo "== $lib"
  RTB_LIB=$PWD/$lib python bench.py --steps 20 --warmup 3 --no-cpu 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step', round(d['ms_per_step'],4), 'kernel_ms', round(d['roofline']['kernel_ms'],4), 'e2e', round(d['e2e']['ms_per_step'],4), d['config']['launch'])"
done
