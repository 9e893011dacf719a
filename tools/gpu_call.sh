set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tile_cull.py tests/test_gpu_parity.py -x -q > gpurun_out/pytest_cull.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_cull.log
tail -15 gpurun_out/pytest_cull.log
for wl in c2 c3 c4; do timeout 300 python tools/variant_sweep.py --workload $wl --variants 0,cull,binned --flush; done > gpurun_out/sweep_binned.jsonl 2>&1; cat gpurun_out/sweep_binned.jsonl
RTB_SWEEP_TILE_CULL=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"binned|bin_tiles" -c 2 -o gpurun_out/binned_full python tools/variant_sweep.py --child --workload c3 --reps 2 --flush > gpurun_out/ncu_binned.log 2>&1; tail -3 gpurun_out/ncu_binned.log
