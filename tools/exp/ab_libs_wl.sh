#!/bin/bash
# A/B of library builds on one workload: ab_libs_wl.sh workload steps name...
W=$1; S=$2; shift 2
for rep in 1 2; do
for n in "$@"; do
  if [ "$n" = base ]; then L=$PWD/example-gui-opencl-raytracer_b200/librtb200.so; else L=$PWD/example-gui-opencl-raytracer_b200/librtb200_$n.so; fi
  RTB_LIB=$L python bench.py --steps $S --warmup 3 --no-cpu --other-configs off --workload $W | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$n', round(d['ms_per_step'],4), d['config']['launch'])"
done; done
