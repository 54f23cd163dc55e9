#!/bin/bash
# rt_trace_solo.cuh on the GPU: parity of the new path, then A/B timing (kernel and bench lines).
mkdir -p gpurun_out/solo
O=gpurun_out/solo
timeout 300 python -m pytest tests/test_parity_gpu.py tests/test_fuzz_gpu.py -m gpu -x -q -k "solo" 2>&1 | tail -15 | tee $O/pytest_solo.txt
timeout 300 python tools/exp/sweep_solo.py 8 2>&1 | tail -12 | tee $O/sweep_solo.txt
BENCH_ENGINE_OPTS=solo=1 timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu > $O/bench_solo1.json 2> $O/bench_solo1.err
BENCH_ENGINE_OPTS=solo=0 timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu > $O/bench_solo0.json 2> $O/bench_solo0.err
cut -c1-200 $O/bench_solo1.json $O/bench_solo0.json
