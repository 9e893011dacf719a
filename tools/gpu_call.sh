set -x
mkdir -p gpurun_out
for v in 0 1 2 3 4 5; do RTB_BINNED_VARIANT=$v timeout -k 5 100 python -m pytest tests/test_gpu_tile_cull.py -x -q 2>&1 | tail -1; done > gpurun_out/pytest_variants.log 2>&1; cat gpurun_out/pytest_variants.log
for wl in c2 c3 c4; do timeout -k 5 200 python tools/variant_sweep.py --workload $wl --variants cull,binned:0,binned:1,binned:2,binned:3,binned:4,binned:5 --flush; done > gpurun_out/sweep_binned.jsonl 2>&1; cut -c1-120 gpurun_out/sweep_binned.jsonl
