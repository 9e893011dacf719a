# Multi-GPU bench lines the way the driver launches them (on an N-GPU box):
#   /usr/local/graft/bin/gpurun --gpus 8 --timeout 1500 -- 'bash tools/gpu_scale.sh "8 4 2" r02 v11'
set -x
mkdir -p gpurun_out
nvidia-smi -L | head -8
for n in $1; do
  timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/${2}_bench_c3_n${n}_${3}.json 2> gpurun_out/bench_n${n}.err
  tail -c 600 gpurun_out/${2}_bench_c3_n${n}_${3}.json; tail -3 gpurun_out/bench_n${n}.err
done
