# usage: bash scripts/gpu_e2e_ab.sh <tag> — GPU tests, then device-time and end-to-end (pinned host canvas) per build/var/*.so
tag=$1
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/${tag}_pytest.log; tail -3 gpurun_out/${tag}_pytest.log
for so in build/var/*.so; do
  n=$(basename $so .so)
  for w in reflect_refract_1080p cover_1080p chapter7_800x400; do
    RTC_CUDA_LIB=$PWD/$so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload $w > gpurun_out/${tag}_${n}_$w.log 2>&1
    python - "$n $w" gpurun_out/${tag}_${n}_$w.log <<'PY'
import json,sys
for l in open(sys.argv[2]):
    if l.startswith('{"metric'):
        d=json.loads(l); print(sys.argv[1], "ms=%.3f e2e_ms=%.3f"%(d["ms_per_step"],d["e2e"]["ms_per_step"]))
        break
else: print(sys.argv[1],"FAILED"); print(open(sys.argv[2]).read()[-600:])
PY
  done
  RTC_CUDA_LIB=$PWD/$so timeout 120 python scripts/e2e_breakdown.py 2>&1 | tail -2 | sed "s/^/$n /"
done
