#!/bin/bash
# N=2 revalidation after the WIDE variant: both bench arms as the driver launches them, frame_equals_1gpu checked by bench.py
mkdir -p gpurun_out/v10; O=gpurun_out/v10
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 "${@:2}"; }
run 29611 --impl reference --steps 1 --warmup 0 > $O/ref_n2.json 2> $O/ref_n2.err; echo "ref rc=$?"
run 29612 --steps 20 --warmup 3 --no-cpu > $O/bench_n2.json 2> $O/bench_n2.err; echo "n2 rc=$?"
run 29613 --steps 5 --warmup 3 --no-cpu --workload spheres1024_4k > $O/bench_spheres1024_n2.json 2> $O/bench_spheres1024_n2.err; echo "grid n2 rc=$?"
run 29614 --steps 3 --warmup 3 --no-cpu --workload spheres1024_4k --exchange nccl > $O/bench_spheres1024_n2_nccl.json 2> $O/bench_spheres1024_n2_nccl.err; echo "grid nccl rc=$?"
for f in ref_n2 bench_n2 bench_spheres1024_n2 bench_spheres1024_n2_nccl; do python - $O/$f.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], d.get('ms_per_step'), d.get('value'), (d.get('config') or {}).get('frame_equals_1gpu'), (d.get('e2e') or {}).get('value'))
PY
done
tail -3 $O/*.err
