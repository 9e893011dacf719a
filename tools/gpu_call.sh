set -x
mkdir -p gpurun_out
timeout -k 5 200 python -m pytest tests/test_gpu_tile_cull.py -x -q 2>&1 | tail -2
(timeout -k 5 100 python tools/e2e_breakdown.py; timeout -k 5 100 python tools/e2e_breakdown.py binned; timeout -k 5 100 python tools/e2e_breakdown.py c4 binned) > gpurun_out/e2e_breakdown.jsonl 2>&1; cat gpurun_out/e2e_breakdown.jsonl
