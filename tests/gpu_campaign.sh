#!/bin/bash
# end-of-round measurement campaign (1 GPU): the other BASELINE.json configs + profile of the current kernel
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/final_n1.json 2> gpurun_out/final_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err
python bench.py --steps 5 --warmup 3 --no-cpu --workload spheres1024_4k > gpurun_out/final_spheres1024.json 2> gpurun_out/final_spheres1024.err
python bench.py --steps 3 --warmup 3 --no-cpu --workload spheres255_8k > gpurun_out/final_spheres255_8k.json 2> gpurun_out/final_spheres255_8k.err
python bench.py --steps 2 --warmup 3 --no-cpu --workload mesh50k_4k > gpurun_out/final_mesh50k_4k.json 2> gpurun_out/final_mesh50k_4k.err
drivers/_bin/orbit_bench 2560 1440 240 > gpurun_out/final_orbit_1440p.json 2>&1
RTB_FRAMES=240 drivers/_bin/rayinteractive > gpurun_out/final_rayinteractive_800x600.txt 2>&1
drivers/_bin/raypng > gpurun_out/final_raypng.txt 2>&1
bash tests/gpu_profile.sh > gpurun_out/profile.log 2>&1
ls -la gpurun_out | tail -20
