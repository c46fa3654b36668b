#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --workload primitives_csg_1080p --steps 2 --warmup 3 --no-cpu-baseline --no-l2-flush"
$B > gpurun_out/r02bi_plain_csg.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 \
  -o gpurun_out/r02bi_prof_csg $B > gpurun_out/r02bi_ncu_csg.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02bi_prof_csg.ncu-rep > gpurun_out/r02bi_ncu_csg_summary.md 2>&1
B="python bench.py --workload cover_1080p --steps 2 --warmup 3 --no-cpu-baseline --no-l2-flush"
$B > gpurun_out/r02bi_plain_cover.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 \
  -o gpurun_out/r02bi_prof_cover $B > gpurun_out/r02bi_ncu_cover.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02bi_prof_cover.ncu-rep > gpurun_out/r02bi_ncu_cover_summary.md 2>&1
B="python bench.py --workload teapot_standin_1080p --steps 2 --warmup 3 --no-cpu-baseline --no-l2-flush"
$B > gpurun_out/r02bi_plain_teapot.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 3 -c 1 \
  -o gpurun_out/r02bi_prof_teapot $B > gpurun_out/r02bi_ncu_teapot.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02bi_prof_teapot.ncu-rep > gpurun_out/r02bi_ncu_teapot_summary.md 2>&1
cat gpurun_out/r02bi_ncu_csg_summary.md
