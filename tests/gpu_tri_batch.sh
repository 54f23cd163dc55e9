#!/bin/bash
# Batched triangle walk (rt_device.cuh:for_triangle_hits): A/B of RT_TRI_BATCH on the 5,000-triangle mesh at 4K
# (variants: make variant NAME=tb8 DEFS=-DRT_TRI_BATCH=8 ...), parity of each variant first.
mkdir -p gpurun_out/tb
O=gpurun_out/tb
P=example-gui-opencl-raytracer_b200
for v in "" $@; do
  RTB_LIB=$PWD/$P/librtb200$v.so timeout 300 python -m pytest tests/test_triangles_gpu.py -m gpu -x -q 2>&1 | tail -1 | tee $O/pytest_tri$v.txt
  RTB_LIB=$PWD/$P/librtb200$v.so timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu --workload mesh5k_4k > $O/bench_mesh5k$v.json 2> $O/bench_mesh5k$v.err
  python -c "import json,sys; d=json.loads(open('$O/bench_mesh5k$v.json').read().strip().splitlines()[-1]); print('batch$v', d['ms_per_step'], d['roofline']['achieved'])"
done
