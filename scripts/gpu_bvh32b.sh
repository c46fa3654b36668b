#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02ae_bvh32.log
: > $L
RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_f64.so python scripts/bvh_scene_times.py f64 >> $L 2>&1
python scripts/bvh_scene_times.py f32 >> $L 2>&1
python - >> $L 2>&1 <<'PY'
import numpy as np
for n in ["teapot_standin", "cover", "primitives_csg", "mesh_and_spheres"]:
    a, b = np.load("/tmp/il_%s_f64.npy" % n), np.load("/tmp/il_%s_f32.npy" % n)
    print(n, "identical", np.array_equal(a, b))
PY
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
cat $L
