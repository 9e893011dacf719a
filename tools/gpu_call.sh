mkdir -p gpurun_out
timeout -k 5 200 python tools/variant_sweep.py --workload c3 --variants 0,50 --flush 2>&1 | cut -c1-160
timeout -k 5 200 python tools/variant_sweep.py --workload c4 --variants 0,50 --flush 2>&1 | cut -c1-160
