# C5 at N GPUs with one lane and with two lanes per rank:
#   /usr/local/graft/bin/gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_c5n8.sh 8'
set -x
mkdir -p gpurun_out
for l in 2 1; do
  timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29560 + l)) bench.py --gpus $1 --workload c5 --steps 4 --warmup 2 --lanes $l > gpurun_out/r02_bench_c5_n$1_lanes$l.json 2> gpurun_out/c5.err; tail -3 gpurun_out/c5.err
done
