mkdir -p gpurun_out
for v in 3 4; do RTB_BINNED_VARIANT=$v timeout -k 5 150 python -m pytest tests/test_gpu_tile_cull.py -x -q 2>&1 | tail -1; done
for wl in c2 c3 c4; do timeout -k 5 200 python tools/variant_sweep.py --workload $wl --variants binned:0,binned:3,binned:4,binned:0,binned:3 --flush; done > gpurun_out/sweep_wlist.jsonl 2>&1; cut -c1-125 gpurun_out/sweep_wlist.jsonl
