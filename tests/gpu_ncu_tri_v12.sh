#!/bin/bash
# r01 v12: one full ncu capture of the triangle unit's render_frame on the 5,000-triangle mesh at 4K (after the plain run exits 0).
mkdir -p gpurun_out/v12
O=gpurun_out/v12
CMD="python bench.py --steps 1 --warmup 3 --no-cpu --workload mesh5k_4k"
$CMD > $O/plain_mesh.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:render_frame -s 3 -c 1 -f -o $O/prof_tri $CMD > $O/ncu_full_tri.log 2>&1
tail -3 $O/ncu_full_tri.log; ls -la $O/prof_tri.ncu-rep
