#!/bin/bash
# r02 final evidence set on one B200: every GPU test, smoke, both bench arms at N=1 (with other_configs), launch list,
# full ncu captures of render_frame (render.map), the WIDE variant (1,024-sphere grid), render_tree (one N=8 part) and the
# N4-over-triangles build (49,928-triangle mesh).
O=gpurun_out/r02f
mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee $O/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee $O/smoke.txt
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cut -c1-300 $O/bench_ref.json
python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -3 $O/bench_n1.err; cut -c1-400 $O/bench_n1.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --other-configs off"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv $CMD > $O/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 4 -c 1 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
CMDW="python bench.py --steps 1 --warmup 3 --no-cpu --other-configs off --workload spheres1024_4k"
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 3 -c 1 -f -o $O/prof_wide $CMDW > $O/ncu_full_wide.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:render_tree -s 28 -c 1 -f -o $O/prof_tree python tools/exp/sweep_tree.py render_map_4k 1:0:0:0 > $O/ncu_full_tree.log 2>&1   # the 29th launch: an N=8 part
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 5 -c 1 -f -o $O/prof_tacc python tools/exp/tacc_probe.py > $O/ncu_full_tacc.log 2>&1   # the mesh through the box hierarchy, a history frame
tail -2 $O/ncu_full.log $O/ncu_full_wide.log $O/ncu_full_tree.log $O/ncu_full_tacc.log
ls -la $O
