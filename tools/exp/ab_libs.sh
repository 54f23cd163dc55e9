#!/bin/bash
# A/B of library builds: N=1 render.map 4K step time (20 steps) per build, twice, interleaved.  usage: ab_libs.sh name...
for rep in 1 2; do
for n in "$@"; do
  if [ "$n" = base ]; then L=$PWD/example-gui-opencl-raytracer_b200/librtb200.so; else L=$PWD/example-gui-opencl-raytracer_b200/librtb200_$n.so; fi
  RTB_LIB=$L python bench.py --steps 20 --warmup 3 --no-cpu --other-configs off | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$n', round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4))"
done; done
