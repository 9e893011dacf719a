set -x
mkdir -p gpurun_out
timeout -k 5 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/bench_c3_n2_v7.json 2> gpurun_out/bench_c3_n2_v7.err; tail -c 1500 gpurun_out/bench_c3_n2_v7.json; tail -5 gpurun_out/bench_c3_n2_v7.err
timeout -k 5 300 python -m pytest tests/test_gpu_peer_frame.py -x -q 2>&1 | tail -2
