mkdir -p gpurun_out
timeout -k 5 200 python tools/variant_sweep.py --workload c3 --variants 0,0 --flush 2>&1 | cut -c1-125
timeout -k 5 200 python tools/variant_sweep.py --workload c4 --variants 0 --flush 2>&1 | cut -c1-125
