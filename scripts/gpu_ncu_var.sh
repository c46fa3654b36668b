# usage: bash scripts/gpu_ncu_var.sh <tag> <workload> <lib.so>
tag=$1; wl=$2; export RTC_CUDA_LIB=$PWD/$3
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload $wl"
$B > gpurun_out/${tag}_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 -o gpurun_out/${tag}_prof $B > gpurun_out/${tag}_ncu.log 2>&1
tail -1 gpurun_out/${tag}_ncu.log
