#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02ai_sah_stop_wave.log
: > $L
for kt in 0.5 1.0 2.0; do
  echo "RTC_SAH_KT=$kt" >> $L
  RTC_SAH_KT=$kt python scripts/bvh_scene_times.py kt$kt teapot_standin cover primitives_csg >> $L 2>&1
done
echo "RTC_LEAF_MAX=3" >> $L
RTC_LEAF_MAX=3 python scripts/bvh_scene_times.py lm3 teapot_standin cover >> $L 2>&1
cat $L
