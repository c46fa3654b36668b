#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/r02aq_build.log
: > $L
python scripts/bvh_scene_times.py head >> $L 2>&1
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
for w in teapot_standin_1080p mesh_and_spheres_4k; do
  timeout 900 python bench.py --steps 5 --warmup 3 --workload $w > gpurun_out/r02aq_bench_$w.json 2> gpurun_out/r02aq_bench_$w.err
  python - $w >> $L <<'PY'
import json, sys
d = json.loads([l for l in open("gpurun_out/r02aq_bench_%s.json" % sys.argv[1]) if l.startswith("{")][-1])
c = d["cpu_baseline"]
print(sys.argv[1], "dev", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["e2e"]["rank0_ms_per_step"], "parity", c.get("parity_identical_8bit"), c.get("parity_max_linear_diff"), "frac", d["roofline"]["frac"])
PY
done
cat $L
