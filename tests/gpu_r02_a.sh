#!/bin/bash
# r02 campaign A: every GPU test (incl. the new full-size config goldens, N2/N3, non-resident scene), smoke, the N=1 bench
# line with other_configs, the contraction A/B, launch list + full ncu captures (render.map and the WIDE variant).
O=gpurun_out/r02g
mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $O/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee $O/smoke.txt
python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -3 $O/bench_n1.err
cut -c1-600 $O/bench_n1.json

CMD="python bench.py --steps 3 --warmup 3 --no-cpu --other-configs off"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv $CMD > $O/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 4 -c 1 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
CMDW="python bench.py --steps 1 --warmup 3 --no-cpu --other-configs off --workload spheres1024_4k"
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 3 -c 1 -f -o $O/prof_wide $CMDW > $O/ncu_full_wide.log 2>&1
tail -2 $O/ncu_full.log $O/ncu_full_wide.log
ls -la $O
python tools/exp/sweep_tree.py spheres1024_4k 0:0:0:0 2>&1 | tee $O/sweep_spheres1024_auto.txt
for v in b4 b6; do RTB_LIB=$PWD/example-gui-opencl-raytracer_b200/librtb200_$v.so python bench.py --steps 20 --warmup 3 --no-cpu --other-configs off > $O/bench_$v.json 2>/dev/null; cut -c1-200 $O/bench_$v.json; done
