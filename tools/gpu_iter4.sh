set -x
mkdir -p gpurun_out
timeout -k 5 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -8 gpurun_out/pytest_gpu.log
timeout -k 5 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c3_n1.json 2> gpurun_out/bench_c3_n1.err; tail -3 gpurun_out/bench_c3_n1.err
