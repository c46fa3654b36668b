# usage: bash scripts/gpu_nN_scene.sh <tag> <N> <scene> <workload> — one scene on N GPUs: sharded frame == single-GPU frame, bench line with the oracle leg
tag=$1; N=$2; scene=$3; wl=$4
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/check_multigpu.py $scene > gpurun_out/${tag}_check_${scene}_n$N.log 2>&1
echo "check rc=$?" >> gpurun_out/${tag}_check_${scene}_n$N.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --workload $wl --cpu-budget-s 15 > gpurun_out/${tag}_bench_${wl}_n$N.json 2> gpurun_out/${tag}_bench_${wl}_n$N.err
echo "bench rc=$?" >> gpurun_out/${tag}_bench_${wl}_n$N.err
grep "identical\|MULTIGPU\|rc=" gpurun_out/${tag}_check_${scene}_n$N.log; tail -2 gpurun_out/${tag}_bench_${wl}_n$N.err; cut -c1-400 gpurun_out/${tag}_bench_${wl}_n$N.json
