# usage: bash scripts/gpu_scale.sh <tag> <N> — multi-GPU bit-identity check + bench.py under torchrun at N GPUs (headline + 4K mesh)
tag=$1; N=$2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 300 $TR scripts/check_multigpu.py > gpurun_out/${tag}_check_n$N.log 2>&1; tail -6 gpurun_out/${tag}_check_n$N.log
timeout 300 $TR bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/${tag}_bench_n$N.log 2>&1; grep '^{"metric' gpurun_out/${tag}_bench_n$N.log | cut -c1-700 || tail -20 gpurun_out/${tag}_bench_n$N.log
timeout 300 $TR bench.py --gpus $N --steps 5 --warmup 3 --workload mesh_and_spheres_4k > gpurun_out/${tag}_bench_mesh4k_n$N.log 2>&1; grep '^{"metric' gpurun_out/${tag}_bench_mesh4k_n$N.log | cut -c1-700 || tail -20 gpurun_out/${tag}_bench_mesh4k_n$N.log
