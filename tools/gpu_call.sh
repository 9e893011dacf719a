set -x
mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout -k 5 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout -k 5 400 python bench.py > gpurun_out/bench_c3_v7.json 2> gpurun_out/bench_c3_v7.err; tail -c 600 gpurun_out/bench_c3_v7.json
timeout -k 5 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_c3_v7.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; tail -2 gpurun_out/ncu_launches.log | cut -c1-300
RTB_SWEEP_TILE_CULL=2 timeout -k 5 300 ncu --set full --clock-control none --import-source on -k regex:"binned|bin_tiles" -c 2 -o gpurun_out/binned_full_v4 python tools/variant_sweep.py --child --workload c3 --reps 2 --flush > gpurun_out/ncu_binned.log 2>&1; tail -2 gpurun_out/ncu_binned.log
