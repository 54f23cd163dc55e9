#!/bin/bash
# r02 campaign D: render_tree as a kernel of its own beside render_frame
O=gpurun_out/r02d
mkdir -p $O
timeout 900 python -m pytest tests/test_tree_gpu.py tests/test_parity_gpu.py tests/test_fullsize_gpu.py -x -q 2>&1 | tail -8 | tee $O/pytest_tree.txt
python tools/exp/sweep_tree.py render_map_4k 0:0:0:0,1:0:0:0,1:32:44:64,1:36:44:64,1:36:48:64,1:40:48:64,1:32:40:64,1:36:64:64 2>&1 | tee $O/sweep_tree_render_map.txt
python tools/exp/sweep_tree.py spheres1024_4k 0:0:0:0,1:0:0:0 2>&1 | tee $O/sweep_tree_spheres1024.txt
python bench.py --steps 20 --warmup 3 --no-cpu --other-configs off > $O/bench_n1.json 2> $O/bench_n1.err; tail -3 $O/bench_n1.err; cut -c1-300 $O/bench_n1.json
