# usage: bash scripts/gpu_variants.sh <tag> [workload]   — bench every build/var/*.so through RTC_CUDA_LIB
tag=${1:-var}; wl=${2:-reflect_refract_1080p}
for so in build/var/*.so; do
  n=$(basename $so .so)
  RTC_CUDA_LIB=$PWD/$so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload $wl > gpurun_out/${tag}_$n.log 2>&1
  python - "$n" gpurun_out/${tag}_$n.log <<'PY'
import json,sys
for l in open(sys.argv[2]):
    if l.startswith('{"metric'):
        d=json.loads(l); print(sys.argv[1], "ms=%.3f Mrays/s=%.0f e2e_ms=%.2f"%(d["ms_per_step"],d["value"],d["e2e"]["ms_per_step"]))
        break
else: print(sys.argv[1],"FAILED"); print(open(sys.argv[2]).read()[-800:])
PY
done
