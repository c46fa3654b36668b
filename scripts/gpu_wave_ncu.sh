# usage: bash scripts/gpu_wave_ncu.sh <tag> [workload] — launch list and one full capture of the wavefront passes (pure queue mode)
tag=$1; wl=${2:-reflect_refract_1080p}
export RTC_WAVEFRONT=1 RTC_WAVE_INLINE_MIN=${RTC_WAVE_INLINE_MIN:-33}
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload $wl"
$B > gpurun_out/${tag}_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:wave -c 66 --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_launches.log 2>&1
$B > gpurun_out/${tag}_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:wave_forward -s 18 -c 6 -o gpurun_out/${tag}_prof $B > gpurun_out/${tag}_ncu_full.log 2>&1
tail -3 gpurun_out/${tag}_ncu_full.log
