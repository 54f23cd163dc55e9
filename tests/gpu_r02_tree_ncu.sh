#!/bin/bash
# full ncu capture of render_tree on one N=8 part (the 27th launch of the sweep: past the whole-frame and N=4 frames)
O=gpurun_out/r02t
mkdir -p $O
ncu --set full --clock-control none --import-source on -k regex:render_tree -s 28 -c 1 -f -o $O/prof_tree python tools/exp/sweep_tree.py render_map_4k 1:0:0:0 > $O/ncu_full_tree.log 2>&1
tail -3 $O/ncu_full_tree.log
