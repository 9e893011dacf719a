set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout -k 5 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -25 gpurun_out/pytest_gpu.log
./profiles/microbench/pair_rate_probe > gpurun_out/pair_rate_probe.jsonl 2>&1; cat gpurun_out/pair_rate_probe.jsonl
timeout -k 5 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c3_n1.json 2> gpurun_out/bench_c3_n1.err; tail -c 3000 gpurun_out/bench_c3_n1.json; tail -5 gpurun_out/bench_c3_n1.err
timeout -k 5 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_c3_n2.json 2> gpurun_out/bench_c3_n2.err; tail -c 3000 gpurun_out/bench_c3_n2.json; tail -15 gpurun_out/bench_c3_n2.err
