mkdir -p gpurun_out
timeout -k 5 300 python -m pytest tests/test_gpu_tile_cull.py -x -q 2>&1 | tail -12
