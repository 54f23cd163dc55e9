#!/bin/bash
# schedule look-ahead A/B at N GPUs (default 1): bench lines with "lookahead" = 1 (default) and 0
N=${1:-1}
O=gpurun_out/look$N
mkdir -p $O
run() { if [ "$N" = 1 ]; then python bench.py "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"; fi; }
for rep in ${REPS:-1 2}; do
for la in 1 0; do
  BENCH_ENGINE_OPTS=lookahead=$la run --steps 40 --warmup 5 --no-cpu --other-configs off > $O/la${la}_$rep.json 2> $O/la${la}_$rep.err
  python - <<PY
import json
d=json.load(open("$O/la${la}_$rep.json"))
print("N=$N lookahead=$la rep $rep: ms/step %.4f e2e %.4f kernel(slowest) %.4f equal1gpu %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_slowest_rank"], d["config"]["frame_equals_1gpu"]))
PY
done; done
