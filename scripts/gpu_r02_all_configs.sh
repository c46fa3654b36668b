# usage: bash scripts/gpu_r02_all_configs.sh <tag> — every config at its BASELINE size on one B200 WITH the oracle leg (bounded rows):
# bench line (device, e2e, roofline, cpu_baseline all-cores + serial, parity of the e2e frame against the oracle rows)
tag=${1:-r02}
for w in chapter7_800x400 reflect_refract_1080p cover_1080p primitives_csg_1080p teapot_standin_1080p; do
  timeout 900 python bench.py --steps 10 --warmup 3 --workload $w > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err
  echo "$w rc=$?"
done
python - $tag <<'PY'
import json, sys, glob
tag = sys.argv[1]
for f in sorted(glob.glob("gpurun_out/%s_bench_*.json" % tag)):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        c = d["cpu_baseline"] or {}
        print("%-28s dev %.3f ms  %.0f Mrays/s  e2e %.2f ms  cpu %.0f ms (%s thr) serial %s ms  parity %s max8 %s lin %.1e  frac %s" % (
            d["config"]["workload"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], c.get("frame_ms_extrapolated", -1), c.get("cores"),
            round(c.get("serial", {}).get("frame_ms_extrapolated", -1)), c.get("parity_identical_8bit"), c.get("parity_max_8bit_diff"),
            c.get("parity_max_linear_diff", -1), d["roofline"]["frac"]))
    except Exception as e:
        print(f, "ERR", e)
PY
