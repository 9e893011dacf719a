mkdir -p gpurun_out
timeout -k 5 300 python -m pytest tests/test_gpu_tile_cull.py tests/test_gpu_parity.py -x -q 2>&1 | tail -2
for wl in c3 c4; do timeout -k 5 200 python tools/variant_sweep.py --workload $wl --variants cull,binned --flush; done 2>&1 | cut -c1-125
