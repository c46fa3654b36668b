# usage: bash scripts/gpu_variants.sh <tag> <workload>...   — bench every build/var/*.so through RTC_CUDA_LIB
tag=$1; shift
for wl in "$@"; do
for so in build/var/*.so; do
  n=$(basename $so .so)
  RTC_CUDA_LIB=$PWD/$so timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload $wl > gpurun_out/${tag}_${n}_$wl.log 2>&1
  python - "$n $wl" gpurun_out/${tag}_${n}_$wl.log <<'PY'
import json,sys
for l in open(sys.argv[2]):
    if l.startswith('{"metric'):
        d=json.loads(l); print(sys.argv[1], "ms=%.3f Mrays/s=%.0f"%(d["ms_per_step"],d["value"]))
        break
else: print(sys.argv[1],"FAILED"); print(open(sys.argv[2]).read()[-600:])
PY
done; done
