#!/bin/bash
# r01 v9 evidence set (after the WIDE variant and the pow-inf fix): GPU tests, both bench arms, the other configs,
# launch list + full ncu capture of render_frame on render.map, and one full capture of the WIDE variant on the sphere grid.
mkdir -p gpurun_out/v9
O=gpurun_out/v9
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $O/pytest_gpu.txt
python bench.py --impl reference --steps 3 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err
python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err
python bench.py --steps 5 --warmup 3 --no-cpu --workload spheres1024_4k > $O/bench_spheres1024.json 2> $O/bench_spheres1024.err
python bench.py --steps 3 --warmup 3 --no-cpu --workload spheres255_8k > $O/bench_spheres255_8k.json 2> $O/bench_spheres255_8k.err
python bench.py --steps 2 --warmup 3 --no-cpu --workload mesh50k_4k > $O/bench_mesh50k_4k.json 2> $O/bench_mesh50k_4k.err
drivers/_bin/orbit_bench 2560 1440 240 > $O/orbit_1440p.json 2>&1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv $CMD > $O/ncu_list.log 2>&1
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 4 -c 1 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
CMDW="python bench.py --steps 1 --warmup 3 --no-cpu --workload spheres1024_4k"
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 3 -c 1 -f -o $O/prof_wide $CMDW > $O/ncu_full_wide.log 2>&1
tail -3 $O/ncu_full.log $O/ncu_full_wide.log
cat $O/bench_n1.json
ls -la $O
