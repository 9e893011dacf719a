mkdir -p gpurun_out
for wl in c2 c1r; do timeout -k 5 200 python tools/variant_sweep.py --workload $wl --variants 0,10,15,20 --flush --reps 50; done > gpurun_out/sweep_small.jsonl 2>&1; cut -c1-140 gpurun_out/sweep_small.jsonl
