#!/bin/bash
N=$1
for b in 4 8 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 40 --warmup 5 --no-cpu --other-configs off --band-rows $b > /tmp/b$b.json 2> /tmp/b$b.err
python - <<PY
import json
d=json.load(open("/tmp/b$b.json"))
print("N=$N band_rows=$b: ms/step %.4f e2e %.4f kernel(slowest) %.4f mean-kernel %.4f equal1gpu %s" % (d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["kernel_ms_slowest_rank"], d["roofline"]["kernel_ms"], d["config"]["frame_equals_1gpu"]))
PY
done
