set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/check_multigpu.py > gpurun_out/r02k_check_n2.log 2>&1
echo "check rc=$?" >> gpurun_out/r02k_check_n2.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r02k_bench_n2.json 2> gpurun_out/r02k_bench_n2.err
echo "bench rc=$?" >> gpurun_out/r02k_bench_n2.err
tail -12 gpurun_out/r02k_check_n2.log; cat gpurun_out/r02k_bench_n2.json; tail -5 gpurun_out/r02k_bench_n2.err
