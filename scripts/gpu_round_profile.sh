# usage: bash scripts/gpu_round_profile.sh <tag> — default bench line, ncu launch list, one ncu --set full capture
tag=$1
python bench.py > gpurun_out/${tag}_bench.log 2>&1; tail -c 400 gpurun_out/${tag}_bench.log
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/${tag}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_launches.log 2>&1
$B > gpurun_out/${tag}_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 2 -o gpurun_out/${tag}_prof $B > gpurun_out/${tag}_ncu_full.log 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.log 2>&1; tail -c 600 gpurun_out/${tag}_bench_reference.log
