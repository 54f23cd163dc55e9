#!/bin/bash
# one full ncu capture of render_frame<1,0,0> on render.map 4K (history frame)
O=gpurun_out/r02h
mkdir -p $O
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --other-configs off"
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 4 -c 1 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
tail -2 $O/ncu_full.log
