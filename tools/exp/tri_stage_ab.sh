#!/bin/bash
# A/B of the staged triangle walk (tree kernel, every tile) against the cooperative walk on the 5k-triangle mesh at 4K
for v in base tb4; do
if [ "$v" = base ]; then L=$PWD/example-gui-opencl-raytracer_b200/librtb200.so; else L=$PWD/example-gui-opencl-raytracer_b200/librtb200_$v.so; fi
for o in "tri_tree_min=0" "tri_tree_min=1024,tree_ctas_per_sm=4"; do
RTB_LIB=$L BENCH_ENGINE_OPTS=$o python bench.py --steps 3 --warmup 3 --no-cpu --other-configs off --workload mesh5k_4k | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v $o', round(d['ms_per_step'],2), d['config']['launch']['tree_grid'], round(d['roofline']['achieved'],2))"
done; done
