# usage: bash scripts/gpu_devbvh.sh <tag>  — the GPU-built BVH: parity on the small mesh cases (builder forced), config 5 at scale
tag=$1
RTC_DEVICE_BVH_MIN=16 RTC_TIMING=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_edge_cases.py -m gpu -q -x -k "teapot or mesh or solids or big_accel" > gpurun_out/${tag}_devbvh_tests.log 2>&1
echo "rc=$?" >> gpurun_out/${tag}_devbvh_tests.log
grep -c "device BVH build" gpurun_out/${tag}_devbvh_tests.log
RTC_TIMING=1 python scripts/verify_config5.py > gpurun_out/${tag}_verify_config5_devbvh.json 2> gpurun_out/${tag}_verify_config5_devbvh.err
RTC_TIMING=1 python bench.py --workload mesh_and_spheres_4k --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_mesh_devbvh.json 2> gpurun_out/${tag}_bench_mesh_devbvh.err
RTC_DEVICE_BVH_MIN=1000 RTC_TIMING=1 python bench.py --workload teapot_standin_1080p --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_teapot_devbvh.json 2> gpurun_out/${tag}_bench_teapot_devbvh.err
python bench.py --workload teapot_standin_1080p --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_teapot_hostbvh.json 2>/dev/null
tail -4 gpurun_out/${tag}_devbvh_tests.log; cat gpurun_out/${tag}_verify_config5_devbvh.json; grep -h "device BVH\|build_accel\|rtc_scene_create" gpurun_out/${tag}_verify_config5_devbvh.err gpurun_out/${tag}_bench_teapot_devbvh.err | tail -6
for f in mesh_devbvh teapot_devbvh teapot_hostbvh; do python - gpurun_out/${tag}_bench_$f.json <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1], "dev %.3f ms e2e %.2f ms" % (d["ms_per_step"], d["e2e"]["ms_per_step"]), d["e2e"]["rank0_ms_per_step"])
PY
done
