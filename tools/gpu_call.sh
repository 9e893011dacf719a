set -x
mkdir -p gpurun_out
timeout -k 5 300 python -m pytest tests/test_gpu_prepack.py tests/test_gpu_tile_cull.py tests/test_gpu_bands_and_fullsize.py -x -q 2>&1 | tail -15
timeout -k 5 300 python bench.py --workload c5 --steps 3 --warmup 1 > gpurun_out/bench_c5_n1_v7.json 2> gpurun_out/bench_c5_n1_v7.err; tail -c 1800 gpurun_out/bench_c5_n1_v7.json; tail -3 gpurun_out/bench_c5_n1_v7.err
