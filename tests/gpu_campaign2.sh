python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 20 --warmup 3 > gpurun_out/final_n1.json 2> gpurun_out/final_n1.err
python bench.py --steps 5 --warmup 3 --no-cpu --workload spheres1024_4k > gpurun_out/final_spheres1024.json 2> gpurun_out/final_spheres1024.err
python bench.py --steps 3 --warmup 3 --no-cpu --workload spheres255_8k > gpurun_out/final_spheres255_8k.json 2> gpurun_out/final_spheres255_8k.err
python bench.py --steps 2 --warmup 3 --no-cpu --workload mesh50k_4k > gpurun_out/final_mesh50k_4k.json 2> gpurun_out/final_mesh50k_4k.err
drivers/_bin/orbit_bench 2560 1440 240 > gpurun_out/final_orbit_1440p.json 2>&1
bash tests/gpu_profile.sh > gpurun_out/profile.log 2>&1
