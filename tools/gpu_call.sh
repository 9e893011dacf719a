mkdir -p gpurun_out
timeout -k 5 200 python tools/variant_sweep.py --workload c3 --variants 0,32,33,31 --flush 2>&1 | cut -c1-125
