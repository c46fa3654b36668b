# usage: bash scripts/gpu_nN.sh <tag> <N> [check]  — torchrun bench (and optionally check_multigpu) on N GPUs of this box
tag=$1; N=$2
if [ "$3" = "check" ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/check_multigpu.py > gpurun_out/${tag}_check_n$N.log 2>&1
echo "check rc=$?" >> gpurun_out/${tag}_check_n$N.log
fi
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${tag}_bench_reflect_refract_1080p_n$N.json 2> gpurun_out/${tag}_bench_n$N.err
echo "bench rc=$?" >> gpurun_out/${tag}_bench_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 --workload mesh_and_spheres_4k --no-cpu-baseline > gpurun_out/${tag}_bench_mesh_and_spheres_4k_n$N.json 2> gpurun_out/${tag}_bench_mesh_n$N.err
tail -12 gpurun_out/${tag}_check_n$N.log 2>/dev/null; cat gpurun_out/${tag}_bench_reflect_refract_1080p_n$N.json; tail -3 gpurun_out/${tag}_bench_n$N.err; cat gpurun_out/${tag}_bench_mesh_and_spheres_4k_n$N.json
