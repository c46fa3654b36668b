#!/bin/bash
# usage: gpu_ool.sh <logname> <variant>...   — HEAD and each build/var variant: BVH scenes + headline frame
mkdir -p gpurun_out
L=gpurun_out/$1; shift
: > $L
echo "HEAD" >> $L
python scripts/bvh_scene_times.py head teapot_standin cover primitives_csg 2>&1 | cut -c1-120 >> $L
python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | grep -o '"ms_per_step": [0-9.]*' | head -1 >> $L
for v in "$@"; do
  echo "variant $v" >> $L
  RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_$v.so python scripts/bvh_scene_times.py $v teapot_standin cover primitives_csg 2>&1 | cut -c1-120 >> $L
  RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_$v.so python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | grep -o '"ms_per_step": [0-9.]*' | head -1 >> $L
done
cat $L
