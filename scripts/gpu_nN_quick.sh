# usage: bash scripts/gpu_nN_quick.sh <tag> <N>  — torchrun bench of the headline on N GPUs, with the library's timing prints
tag=$1; N=$2
RTC_TIMING=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/${tag}_bench_reflect_refract_1080p_n$N.json 2> gpurun_out/${tag}_bench_n$N.err
echo "bench rc=$?" >> gpurun_out/${tag}_bench_n$N.err
cat gpurun_out/${tag}_bench_reflect_refract_1080p_n$N.json; grep "staged" gpurun_out/${tag}_bench_n$N.err | tail -8
