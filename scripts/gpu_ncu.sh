# usage: bash scripts/gpu_ncu.sh <tag> <workload>  — plain run, then one ncu --set full capture of the render kernel
tag=$1; wl=$2
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload $wl"
$B > gpurun_out/${tag}_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 -o gpurun_out/${tag}_prof $B > gpurun_out/${tag}_ncu.log 2>&1
tail -2 gpurun_out/${tag}_ncu.log
