set -x
mkdir -p gpurun_out
timeout -k 5 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log | cut -c1-150
timeout -k 5 400 python bench.py > gpurun_out/bench_c3_v9.json 2> gpurun_out/bench_c3_v9.err; tail -c 200 gpurun_out/bench_c3_v9.json; tail -3 gpurun_out/bench_c3_v9.err
timeout -k 5 300 python bench.py --workload c5 --steps 3 --warmup 1 > gpurun_out/bench_c5_n1_v9.json 2> gpurun_out/bench_c5_n1_v9.err
timeout -k 5 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_c3_v9.json 2> gpurun_out/bench_ref_c3_v9.err; tail -c 300 gpurun_out/bench_ref_c3_v9.json
timeout -k 5 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_c3_v9.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
timeout -k 5 300 ncu --set full --clock-control none --import-source on -k regex:raycast_grid -c 1 --launch-skip 3 -o gpurun_out/grid_full_v9 python tools/variant_sweep.py --child --workload c3 --reps 2 --flush > gpurun_out/ncu_grid.log 2>&1; tail -1 gpurun_out/ncu_grid.log
timeout -k 5 200 python tools/e2e_breakdown.py > gpurun_out/e2e_breakdown_v9.jsonl 2>&1; cat gpurun_out/e2e_breakdown_v9.jsonl
timeout -k 5 200 python tools/band_kernel_time.py --variants 0 > gpurun_out/band_kernel_v9.jsonl 2>&1; cut -c1-100 gpurun_out/band_kernel_v9.jsonl
