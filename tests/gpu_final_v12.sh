#!/bin/bash
# r01 v12: confirmation of the final tree (batched triangle walk in the triangle unit; main unit unchanged since v11).
mkdir -p gpurun_out/v12
O=gpurun_out/v12
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee $O/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee $O/smoke.txt
python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err
cut -c1-220 $O/bench_n1.json
python bench.py --steps 3 --warmup 3 --no-cpu --workload mesh5k_4k > $O/bench_mesh5k.json 2> $O/bench_mesh5k.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --workload mesh5k_4k"
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_mesh5k.csv $CMD > $O/ncu_list_mesh.log 2>&1
tail -3 $O/launches_mesh5k.csv | cut -c1-250
