# What a round's GPU check runs (on the GPU box, from the repo root):
#   /usr/local/graft/bin/gpurun --timeout 2400 -- 'bash tools/gpu_call.sh r02 v11'
# (compute-sanitizer: tools/sanitize.sh -- closed on this pool, profiles/r02_sanitizer_closed_on_pool.txt;
#  ncu: tools/gpu_ncu.sh; multi-GPU lines: tools/gpu_scale.sh on a --gpus N box)
set -x
mkdir -p gpurun_out
timeout -k 5 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout -k 5 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log | cut -c1-160
timeout -k 5 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${1:-r02}_bench_c3_n1_${2:-v11}.json 2> gpurun_out/bench_c3.err; tail -c 300 gpurun_out/${1:-r02}_bench_c3_n1_${2:-v11}.json; tail -3 gpurun_out/bench_c3.err
timeout -k 5 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${1:-r02}_bench_ref_c3_${2:-v11}.json 2> gpurun_out/bench_ref.err; tail -c 300 gpurun_out/${1:-r02}_bench_ref_c3_${2:-v11}.json; tail -3 gpurun_out/bench_ref.err
for wl in c1 c2; do timeout -k 5 300 python bench.py --workload $wl --steps 20 --warmup 5 > gpurun_out/${1:-r02}_bench_${wl}_n1_${2:-v11}.json 2> gpurun_out/bench_$wl.err; tail -c 200 gpurun_out/${1:-r02}_bench_${wl}_n1_${2:-v11}.json; tail -2 gpurun_out/bench_$wl.err; done
