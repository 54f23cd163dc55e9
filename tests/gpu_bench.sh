#!/bin/bash
# bench + launch list + one full ncu capture of the render kernel
set -x
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"
tail -3 gpurun_out/bench_n1.err
cat gpurun_out/bench_n1.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?"
cat gpurun_out/bench_ref.json; tail -3 gpurun_out/bench_ref.err
