#!/bin/bash
# r02 multi-GPU: the bench line (with other_configs) under torchrun, BENCH_DEBUG spans per rank
O=gpurun_out/r02n$1
mkdir -p $O
BENCH_DEBUG=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $1 --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err
grep "rank" $O/bench.err | grep -v Warn | cut -c1-400
tail -3 $O/bench.err | cut -c1-300
python - <<PY
import json
d=json.load(open("$O/bench.json"))
print({k:d[k] for k in ("value","ms_per_step","n_gpus","gpu_launches")}, d["e2e"]["ms_per_step"], d["config"]["frame_equals_1gpu"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_slowest_rank"])
for k,v in d.get("other_configs",{}).items(): print(k, v.get("ms_per_step"), v.get("frame_equals_1gpu"), v.get("kernel_ms_slowest_rank"), v.get("exchange"))
PY
