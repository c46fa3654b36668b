#!/bin/bash
# single-precision BVH nodes (HEAD) vs the double-precision nodes (variant f64): times, identical frames, then the GPU tests
mkdir -p gpurun_out
L=gpurun_out/r02ad_bvh32.log
: > $L
RTC_CUDA_LIB=$PWD/build/var/librtc_cuda_f64.so python scripts/bvh_scene_times.py f64 >> $L 2>&1
python scripts/bvh_scene_times.py f32 >> $L 2>&1
python - >> $L 2>&1 <<'PY'
import numpy as np
for n in ["teapot_standin", "cover", "primitives_csg", "mesh_and_spheres"]:
    a, b = np.load("/tmp/il_%s_f64.npy" % n), np.load("/tmp/il_%s_f32.npy" % n)
    print(n, "identical", np.array_equal(a, b))
PY
python -m pytest tests -x -q -m gpu >> $L 2>&1
tail -5 $L
python bench.py --steps 20 --warmup 3 >> $L 2>&1
python bench.py --workload mesh_and_spheres_4k --steps 5 --warmup 3 >> $L 2>&1
python bench.py --workload teapot_standin_1080p --steps 10 --warmup 3 >> $L 2>&1
python bench.py --workload cover_1080p --steps 10 --warmup 3 >> $L 2>&1
grep -v "^{" $L | head -40
grep -o '"workload": "[a-z_0-9]*"\|"ms_per_step": [0-9.]*\|"parity": {[^}]*}' $L
