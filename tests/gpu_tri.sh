#!/bin/bash
mkdir -p gpurun_out/v10; O=gpurun_out/v10
python -m pytest tests -m gpu -x -q -k "triangle or mesh" 2>&1 | tail -2 | tee $O/pytest_tri.txt
for lib in librtb200.so librtb200_noprefilter.so; do
  RTB_LIB=$PWD/example-gui-opencl-raytracer_b200/$lib python bench.py --steps 3 --warmup 3 --no-cpu --workload mesh5k_4k 2>$O/err_$lib.txt | tail -1 > $O/bench_mesh5k_$lib.json
  python -c "import json,sys; d=json.loads(open('$O/bench_mesh5k_$lib.json').read()); print('$lib', round(d['ms_per_step'],2), round(d['roofline']['achieved'],2))"
done
timeout 300 ncu --set full --clock-control none --import-source on -k regex:render_frame -s 3 -c 1 -f -o $O/prof_tri python bench.py --steps 1 --warmup 3 --no-cpu --workload mesh5k_4k > $O/ncu_tri.log 2>&1
