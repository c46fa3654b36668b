# usage: bash scripts/gpu_wave2.sh <tag> [workloads...] — GPU tests, recursive vs wavefront on the given workloads, shard probe (wavefront)
tag=$1; shift
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15) > gpurun_out/${tag}_pytest.log; tail -4 gpurun_out/${tag}_pytest.log
for w in "$@"; do
for m in 0 1; do
  RTC_WAVEFRONT=$m timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload $w > gpurun_out/${tag}_w${m}_$w.log 2>&1
  python - "$w wave=$m" gpurun_out/${tag}_w${m}_$w.log <<'PY'
import json,sys
for l in open(sys.argv[2]):
    if l.startswith('{"metric'):
        d=json.loads(l); print(sys.argv[1], "ms=%.3f Mrays/s=%.0f e2e_ms=%.2f rays=%d"%(d["ms_per_step"],d["value"],d["e2e"]["ms_per_step"],d["config"]["rays_per_frame"]))
        break
else: print(sys.argv[1],"FAILED"); print(open(sys.argv[2]).read()[-800:])
PY
done; done
RTC_WAVEFRONT=1 timeout 300 python scripts/shard_probe.py > gpurun_out/${tag}_shard_probe_wave.log 2>&1; cat gpurun_out/${tag}_shard_probe_wave.log
