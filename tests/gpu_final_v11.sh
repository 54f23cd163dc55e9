#!/bin/bash
# r01 v11: end-of-round confirmation on a fresh box of the tree as committed (libraries rebuilt from source by build()):
# GPU tests, smoke(), both bench arms at N=1, launch list and one full ncu capture of render_frame on render.map.
mkdir -p gpurun_out/v11
O=gpurun_out/v11
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $O/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee $O/smoke.txt
python bench.py --impl reference --steps 3 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err
python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv $CMD > $O/ncu_list.log 2>&1
$CMD > $O/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render_frame -s 4 -c 1 -f -o $O/prof $CMD > $O/ncu_full.log 2>&1
tail -2 $O/ncu_full.log
cat $O/bench_n1.json
ls -la $O
