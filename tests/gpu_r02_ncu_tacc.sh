#!/bin/bash
# one full ncu capture of render_frame<1,0,1> of the N4-over-triangles build (rt_kernels_tri_accel.cu) on the 49,928-triangle
# mesh at 3840x2160 (a history-scheduled frame)
O=gpurun_out/r02tacc
mkdir -p $O
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 5 -c 1 -f -o $O/prof python tools/exp/tacc_probe.py > $O/ncu_full.log 2>&1
tail -3 $O/ncu_full.log
ncu -i $O/prof.ncu-rep --page details > $O/details.txt 2>&1
ncu -i $O/prof.ncu-rep --page source --csv > $O/source.csv 2>&1
ls -la $O
