#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02aj_wide.log
: > $L
for kt in 0.5 1.0 2.0; do
  echo "binary nodes, RTC_SAH_KT=$kt" >> $L
  RTC_SAH_KT=$kt RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_bvh2.so python scripts/bvh_scene_times.py b2kt$kt teapot_standin cover mesh_and_spheres >> $L 2>&1
  echo "four-child nodes, RTC_SAH_KT=$kt" >> $L
  RTC_SAH_KT=$kt python scripts/bvh_scene_times.py b4kt$kt teapot_standin cover mesh_and_spheres >> $L 2>&1
done
python - >> $L 2>&1 <<'PY'
import numpy as np
for n in ["teapot_standin", "cover", "mesh_and_spheres"]:
    a, b = np.load("/tmp/il_%s_b2kt2.0.npy" % n), np.load("/tmp/il_%s_b4kt1.0.npy" % n)
    print(n, "identical", np.array_equal(a, b))
PY
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
cat $L
