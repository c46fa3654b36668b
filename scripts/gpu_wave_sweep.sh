# usage: bash scripts/gpu_wave_sweep.sh <tag> — RTC_WAVE_REFILL 0/1: N=1 frame times and the one-GPU shard probe
tag=$1
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/${tag}_pytest.log; tail -3 gpurun_out/${tag}_pytest.log
for T in 0 1; do
for w in reflect_refract_1080p teapot_standin_1080p cover_1080p mesh_and_spheres_4k; do
  RTC_WAVE_REFILL=$T RTC_WAVEFRONT=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload $w > gpurun_out/${tag}_T${T}_$w.log 2>&1
  python - "refill=$T $w" gpurun_out/${tag}_T${T}_$w.log <<'PY'
import json,sys
for l in open(sys.argv[2]):
    if l.startswith('{"metric'):
        d=json.loads(l); print(sys.argv[1], "ms=%.3f"%(d["ms_per_step"]))
        break
else: print(sys.argv[1],"FAILED"); print(open(sys.argv[2]).read()[-800:])
PY
done
RTC_WAVE_REFILL=$T RTC_WAVEFRONT=1 timeout 300 python scripts/shard_probe.py 2>&1 | grep "tile_rows=8\|full" | sed "s/^/refill=$T /"
done
