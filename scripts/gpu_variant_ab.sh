# usage: bash scripts/gpu_variant_ab.sh <tag> "<variant names, '-' = the in-tree library>" "<workloads>"  — device ms per frame
tag=$1
for v in $2; do
  for w in $3; do
    if [ "$v" = "-" ]; then lib=""; else lib="build/var/librtc_cuda_$v.so"; fi
    RTC_CUDA_LIB=$lib python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload $w 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('%-10s %-24s dev %.3f ms  e2e %.2f ms' % ('$v', '$w', d['ms_per_step'], d['e2e']['ms_per_step']))
"
  done
done
