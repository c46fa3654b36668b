# usage: bash scripts/gpu_nN_mesh.sh <tag> <N> — the million-triangle scene on N GPUs: sharded frame == single-GPU frame (ipc, nccl, host canvas), bench line with the oracle leg
tag=$1; N=$2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/check_multigpu.py mesh_and_spheres > gpurun_out/${tag}_check_mesh_n$N.log 2>&1
echo "check rc=$?" >> gpurun_out/${tag}_check_mesh_n$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 --workload mesh_and_spheres_4k --cpu-budget-s 20 > gpurun_out/${tag}_bench_mesh_and_spheres_4k_n$N.json 2> gpurun_out/${tag}_bench_mesh_n$N.err
echo "bench rc=$?" >> gpurun_out/${tag}_bench_mesh_n$N.err
grep "identical\|MULTIGPU\|rc=" gpurun_out/${tag}_check_mesh_n$N.log; tail -2 gpurun_out/${tag}_bench_mesh_n$N.err; cat gpurun_out/${tag}_bench_mesh_and_spheres_4k_n$N.json | cut -c1-900
