#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02ak_leafrule.log
: > $L
python scripts/bvh_scene_times.py head >> $L 2>&1
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
bash scripts/gpu_r02_all_configs.sh r02ak >> $L 2>&1
timeout 1200 python bench.py --steps 5 --warmup 3 --workload mesh_and_spheres_4k > gpurun_out/r02ak_bench_mesh_and_spheres_4k.json 2> gpurun_out/r02ak_bench_mesh.err
tail -c 1500 gpurun_out/r02ak_bench_mesh_and_spheres_4k.json >> $L
cat $L
