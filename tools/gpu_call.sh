mkdir -p gpurun_out
for wl in c3 c4 c2; do timeout -k 5 200 python tools/variant_sweep.py --workload $wl --variants 0,15,20 --flush; done 2>&1 | cut -c1-125
