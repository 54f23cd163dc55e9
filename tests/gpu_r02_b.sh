#!/bin/bash
# r02 campaign B: first light of the level-wise pixel-tree trace
O=gpurun_out/r02b
mkdir -p $O
timeout 600 python -m pytest tests/test_tree_gpu.py -x -q 2>&1 | tail -30 | tee $O/pytest_tree.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $O/pytest_gpu.txt
python bench.py --steps 20 --warmup 3 --no-cpu --other-configs off > $O/bench_n1.json 2> $O/bench_n1.err; tail -3 $O/bench_n1.err; cut -c1-300 $O/bench_n1.json
BENCH_ENGINE_OPTS=tree=0 python bench.py --steps 20 --warmup 3 --no-cpu --other-configs off > $O/bench_n1_notree.json 2> $O/bench_n1_notree.err; cut -c1-300 $O/bench_n1_notree.json
