set -x
mkdir -p gpurun_out
N=${NGPU:-8}
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/bench_c3_n${N}_v9.json 2> gpurun_out/bench_c3_n${N}_v9.err; tail -3 gpurun_out/bench_c3_n${N}_v9.err
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload c5 --steps 3 --warmup 1 > gpurun_out/bench_c5_n${N}_v9.json 2> gpurun_out/bench_c5_n${N}_v9.err; tail -3 gpurun_out/bench_c5_n${N}_v9.err
