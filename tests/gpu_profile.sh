#!/bin/bash
# launch list + one full ncu capture of render_frame (bench.py must exit 0 without ncu first)
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 4 -c 1 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
tail -5 gpurun_out/ncu_full.log
ls -la gpurun_out
