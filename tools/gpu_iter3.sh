set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout -k 5 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_noalloc_alone.log; tail -25 gpurun_out/pytest_noalloc_alone.log
timeout -k 5 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -25 gpurun_out/pytest_gpu.log
for m in "dense defer" "binned defer" "dense" ; do timeout -k 5 120 python tools/e2e_breakdown.py $m >> gpurun_out/e2e_breakdown.jsonl 2>&1; done; cat gpurun_out/e2e_breakdown.jsonl
