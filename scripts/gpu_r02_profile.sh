# usage: bash scripts/gpu_r02_profile.sh <tag>  — round-2 profile set on one B200 (each ncu run only after its plain run exited 0)
tag=$1
set -x
# 1. headline: plain bench line, launch list, one --set full capture of the render kernel (no L2 fill ahead of it: clean traffic)
python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench_reflect_refract_1080p.json 2> gpurun_out/${tag}_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${tag}_launches_reflect_refract_1080p.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-l2-flush"
$B > gpurun_out/${tag}_plain_noflush.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 2 \
  -o gpurun_out/${tag}_prof_render_kernel $B > gpurun_out/${tag}_ncu_full.log 2>&1
python scripts/ncu_summary.py gpurun_out/${tag}_prof_render_kernel.ncu-rep > gpurun_out/${tag}_ncu_render_kernel_summary.md 2>&1
# 2. config 5 at its BASELINE size: accelerated == faithful on the full frame, oracle rows, bench line with the cpu leg, ncu
python scripts/verify_config5.py > gpurun_out/${tag}_verify_config5.json 2> gpurun_out/${tag}_verify_config5.err
python bench.py --workload mesh_and_spheres_4k --steps 5 --warmup 3 --cpu-budget-s 40 > gpurun_out/${tag}_bench_mesh_and_spheres_4k.json 2> gpurun_out/${tag}_bench_mesh.err
M="python bench.py --workload mesh_and_spheres_4k --steps 1 --warmup 3 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${tag}_launches_mesh_and_spheres_4k.csv $M > gpurun_out/${tag}_ncu_launches_mesh.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:wave_forward_kernel -s 19 -c 3 -o gpurun_out/${tag}_prof_mesh $M > gpurun_out/${tag}_ncu_full_mesh.log 2>&1
python scripts/ncu_summary.py gpurun_out/${tag}_prof_mesh.ncu-rep > gpurun_out/${tag}_ncu_mesh_summary.md 2>&1
ls -la gpurun_out/ | tail -20
