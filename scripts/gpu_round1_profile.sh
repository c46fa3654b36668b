set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r1_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_launches.csv $B > gpurun_out/r1_ncu_launches.log 2>&1
$B > gpurun_out/r1_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 2 -o gpurun_out/r1_prof $B > gpurun_out/r1_ncu_full.log 2>&1
for w in chapter7_800x400 cover_1080p primitives_csg_1080p teapot_standin_1080p; do
  timeout 300 python bench.py --steps 3 --warmup 3 --workload $w > gpurun_out/r1_bench_$w.log 2>&1; tail -c 1500 gpurun_out/r1_bench_$w.log
done
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload mesh_and_spheres_4k > gpurun_out/r1_bench_mesh4k.log 2>&1; tail -c 1500 gpurun_out/r1_bench_mesh4k.log
