mkdir -p gpurun_out
for r in 1 0 1 0; do echo "ramp $r"; RTB_CHUNK_RAMP=$r timeout -k 5 100 python tools/e2e_breakdown.py 2>&1 | head -1; done > gpurun_out/e2e_ramp.txt; cat gpurun_out/e2e_ramp.txt
timeout -k 5 200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_tile_cull.py tests/test_gpu_bands_and_fullsize.py -x -q 2>&1 | tail -2
