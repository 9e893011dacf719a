set -x
mkdir -p gpurun_out
N=${NGPU:-2}
for N in 2 4; do
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N > gpurun_out/bench_c3_n${N}_v8.json 2> gpurun_out/bench_c3_n${N}_v8.err; tail -c 200 gpurun_out/bench_c3_n${N}_v8.json; tail -3 gpurun_out/bench_c3_n${N}_v8.err
done
