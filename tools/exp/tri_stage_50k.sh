python -m pytest tests/test_triangles_gpu.py -x -q | tail -1
for o in "tri_tree_min=0" "tri_tree_min=1024"; do
BENCH_ENGINE_OPTS=$o python bench.py --steps 2 --warmup 3 --no-cpu --other-configs off --workload mesh50k_4k | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('mesh50k_4k $o', round(d['ms_per_step'],1), d['config']['launch']['tree_grid'], round(d['roofline']['achieved'],2))"
done
