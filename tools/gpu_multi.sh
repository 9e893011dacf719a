# The round's multi-GPU measurements (on an 8-GPU box, from the repo root):
#   /usr/local/graft/bin/gpurun --gpus 8 --timeout 1800 -- 'bash tools/gpu_multi.sh'
# bench lines at 8 / 4 / 2 GPUs the way the driver launches them, C5 at 8 and 1 GPUs, and the single-process
# multi-device context (rtb_init with n_devices = 1, 2, 4, 8).  A/B of the frame fence: tools/gpu_ab2.sh; per-rank
# breakdown of the host-frame call: tools/gpu_bd.sh; host topology: tools/numa_probe.py.
set -x
mkdir -p gpurun_out
bash tools/gpu_scale.sh "8 4 2" r02 v12
timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 8 --workload c5 --steps 4 --warmup 2 > gpurun_out/r02_bench_c5_n8_v12.json 2> gpurun_out/c5n8.err; tail -c 400 gpurun_out/r02_bench_c5_n8_v12.json; tail -2 gpurun_out/c5n8.err
timeout -k 5 300 python bench.py --workload c5 --steps 4 --warmup 2 > gpurun_out/r02_bench_c5_n1_v12.json 2> gpurun_out/c5n1.err; tail -2 gpurun_out/c5n1.err
timeout -k 5 400 python tools/local_group_time.py > gpurun_out/r02_local_group_time.jsonl 2> gpurun_out/lg.err; cat gpurun_out/r02_local_group_time.jsonl; tail -3 gpurun_out/lg.err
