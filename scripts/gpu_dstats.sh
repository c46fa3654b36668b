#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02al_dev_stats.log
: > $L
for wl in teapot_standin_1080p mesh_and_spheres_4k cover_1080p; do
  RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_dstats.so python scripts/dev_stats_frame.py $wl >> $L 2>&1
done
cat $L
