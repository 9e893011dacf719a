# What a round's GPU check runs (on the GPU box, from the repo root):
#   /usr/local/graft/bin/gpurun --timeout 1200 -- 'bash tools/gpu_call.sh'
set -x
mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout -k 5 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log | cut -c1-120
timeout -k 5 400 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; tail -c 150 gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
