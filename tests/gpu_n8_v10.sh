#!/bin/bash
# N=8 (and N=4 on the same box) after the WIDE variant: render.map and the 1,024-sphere grid
mkdir -p gpurun_out/v10; O=gpurun_out/v10
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $1 "${@:3}"; }
run 8 29621 --steps 20 --warmup 3 --no-cpu > $O/bench_n8.json 2> $O/bench_n8.err; echo "n8 rc=$?"
run 8 29622 --steps 5 --warmup 3 --no-cpu --workload spheres1024_4k > $O/bench_spheres1024_n8.json 2> $O/bench_spheres1024_n8.err; echo "grid n8 rc=$?"
run 4 29623 --steps 5 --warmup 3 --no-cpu --workload spheres1024_4k > $O/bench_spheres1024_n4.json 2> $O/bench_spheres1024_n4.err; echo "grid n4 rc=$?"
run 4 29624 --steps 20 --warmup 3 --no-cpu > $O/bench_n4.json 2> $O/bench_n4.err; echo "n4 rc=$?"
for f in bench_n8 bench_n4 bench_spheres1024_n8 bench_spheres1024_n4; do python - $O/$f.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], d.get('ms_per_step'), d.get('value'), (d.get('config') or {}).get('frame_equals_1gpu'), (d.get('e2e') or {}).get('value'))
PY
done
