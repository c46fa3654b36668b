#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02ao_big_minb.log
: > $L
for mb in 10 12; do
  echo "RTK_BIG_MINB=$mb" >> $L
  RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_mb$mb.so python scripts/bvh_scene_times.py mb$mb mesh_and_spheres >> $L 2>&1
done
cat $L
