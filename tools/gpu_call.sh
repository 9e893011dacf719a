mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in c3 c4 c2; do timeout -k 5 200 python tools/variant_sweep.py --workload $wl --variants 0,20,cull,binned --flush; done > gpurun_out/sweep_v9.jsonl 2>&1; cut -c1-125 gpurun_out/sweep_v9.jsonl
