# usage: bash scripts/gpu_push_ab.sh <tag> <N> — bit-identity check, then bench.py at N GPUs with and without rtc_canvas_push_tiles
tag=$1; N=$2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $TR scripts/check_multigpu.py > gpurun_out/${tag}_check_n$N.log 2>&1; tail -3 gpurun_out/${tag}_check_n$N.log
for f in "" "--no-push-tiles"; do
  timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 3 $f > gpurun_out/${tag}_bench_n${N}${f}.log 2>&1
  echo "push=[$f] $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/${tag}_bench_n${N}${f}.log | head -2 | tr '\n' ' ')"
done
timeout 300 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload mesh_and_spheres_4k > gpurun_out/${tag}_bench_mesh4k_n$N.log 2>&1
echo "mesh4k $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/${tag}_bench_mesh4k_n$N.log | head -2 | tr '\n' ' ')"
