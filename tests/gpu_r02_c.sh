#!/bin/bash
# r02 campaign C: CTA-wide pixel-tree phase -- correctness, then threshold sweep, then A/B against a build without the phase
O=gpurun_out/r02c
mkdir -p $O
timeout 900 python -m pytest tests/test_tree_gpu.py tests/test_parity_gpu.py tests/test_fullsize_gpu.py -x -q 2>&1 | tail -8 | tee $O/pytest_tree.txt
python tools/exp/sweep_tree.py render_map_4k 2>&1 | tee $O/sweep_tree_render_map.txt
RTB_LIB=$PWD/example-gui-opencl-raytracer_b200/librtb200_notree.so python tools/exp/sweep_tree.py render_map_4k 0:0:0:0 2>&1 | tee $O/sweep_notree_render_map.txt
python tools/exp/sweep_tree.py spheres1024_4k 0:0:0:0,1:0:0:0,1:32:64:64,1:28:64:64,1:36:64:64,1:32:40:64 2>&1 | tee $O/sweep_tree_spheres1024.txt
