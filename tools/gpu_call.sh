mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in c3 c4 c2; do timeout -k 5 200 python tools/variant_sweep.py --workload $wl --variants 0,cull,binned --flush; done > gpurun_out/sweep_v8b.jsonl 2>&1; cut -c1-125 gpurun_out/sweep_v8b.jsonl
timeout -k 5 400 python bench.py > gpurun_out/bench_c3_v8.json 2> gpurun_out/bench_c3_v8.err; python -c "
import json; j=json.loads(open('gpurun_out/bench_c3_v8.json').read().strip().splitlines()[-1]); print(j['value'], j['ms_per_step'], j['e2e']['value'], j['e2e']['ms_per_step'], j['roofline']['frac'], j['roofline']['dispatch']['frac'], j['binned']['value'], j['binned']['e2e']['value'], j['binned']['same_frame_as_dense'])"; tail -3 gpurun_out/bench_c3_v8.err
timeout -k 5 200 python tools/band_kernel_time.py --variants 0 > gpurun_out/band_kernel_v8.jsonl 2>&1; cut -c1-110 gpurun_out/band_kernel_v8.jsonl
