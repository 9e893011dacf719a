set -x
mkdir -p gpurun_out
timeout -k 5 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log
timeout -k 5 600 python tools/dense_sweep.py --workloads c3 --worlds 1,8 --tilings=-1 --libs ch2,ch4,ch8s2,ch8c256s2 > gpurun_out/dense_sweep_v10b.jsonl 2>&1; cat gpurun_out/dense_sweep_v10b.jsonl | cut -c1-260
