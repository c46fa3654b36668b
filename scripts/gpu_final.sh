#!/bin/bash
# usage: bash scripts/gpu_final.sh <tag> — the round's closing run on one B200: GPU tests, smoke, the driver's two bench arms, the mesh line
tag=$1
mkdir -p gpurun_out
L=gpurun_out/${tag}_final.log
: > $L
python -m pytest tests -x -q -m gpu 2>&1 | tail -3 >> $L
python -c "import __graft_entry__ as g; g.smoke()" >> $L 2>&1
python bench.py > gpurun_out/${tag}_bench_default.json 2> gpurun_out/${tag}_bench_default.err; echo "bench default rc=$?" >> $L
python bench.py --impl reference > gpurun_out/${tag}_bench_reference_arm.json 2> gpurun_out/${tag}_bench_reference_arm.err; echo "bench reference rc=$?" >> $L
python bench.py --workload mesh_and_spheres_4k --steps 5 --warmup 3 --cpu-budget-s 40 > gpurun_out/${tag}_bench_mesh_and_spheres_4k.json 2> gpurun_out/${tag}_bench_mesh.err; echo "bench mesh rc=$?" >> $L
python - $tag >> $L <<'PY'
import json, sys
tag = sys.argv[1]
for name in ["bench_default", "bench_reference_arm", "bench_mesh_and_spheres_4k"]:
    try:
        d = json.loads([l for l in open("gpurun_out/%s_%s.json" % (tag, name)) if l.startswith("{")][-1])
        c = d.get("cpu_baseline") or {}
        print(name, d.get("impl", "cuda"), d["config"]["workload"], "ms/step %.3f" % d["ms_per_step"], "value %.1f %s" % (d["value"], d["unit"]),
              "e2e %.3f ms" % d["e2e"].get("ms_per_step", -1) if "ms_per_step" in d["e2e"] else d["e2e"], "frac", (d.get("roofline") or {}).get("frac"),
              "parity", c.get("parity_identical_8bit"), c.get("parity_max_linear_diff"), "launches", d.get("gpu_launches"), "clocks", (d.get("clocks") or {}).get("sm_mhz"))
    except Exception as e:
        print(name, "ERR", e)
PY
cat $L
