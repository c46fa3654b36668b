#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02bf_minb3.log
: > $L
python scripts/bvh_scene_times.py head >> $L 2>&1
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
echo "lean kernel with 3 CTAs/SM (variant):" >> $L
RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_mb3.so python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | grep -o '"ms_per_step": [0-9.]*' | head -1 >> $L
bash scripts/gpu_r02_all_configs.sh r02bf >> $L 2>&1
cat $L
