#!/bin/bash
# r01 v12: the triangle configurations after the batched triangle walk (BASELINE configs[3]: 50k triangles, 4K and 8K).
mkdir -p gpurun_out/v12
O=gpurun_out/v12
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --workload mesh50k_4k > $O/bench_mesh50k_4k.json 2> $O/bench_mesh50k_4k.err
cut -c1-160 $O/bench_mesh50k_4k.json
timeout 500 python bench.py --steps 1 --warmup 3 --no-cpu --workload mesh50k_8k > $O/bench_mesh50k_8k.json 2> $O/bench_mesh50k_8k.err
cut -c1-160 $O/bench_mesh50k_8k.json; tail -2 $O/bench_mesh50k_8k.err
