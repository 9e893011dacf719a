# ncu evidence of the round's kernels (on the GPU box, from the repo root):
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_ncu.sh r02 v10'
# 1. the launch list of the default bench command; 2. --set full captures of the fixed sequence of tools/ncu_target.py.
tag=${1:-r02}; ver=${2:-v10}
set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-c4 --no-cpu-baseline > gpurun_out/ncu_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_bench_c3_${ver}.csv \
    python bench.py --steps 3 --warmup 3 --no-c4 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python tools/ncu_target.py > gpurun_out/ncu_plain_target.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'raycast_dense_kernel|bin_tiles_kernel|raycast_binned_kernel' -c 10 \
    -f -o gpurun_out/${tag}_prof_${ver} python tools/ncu_target.py > gpurun_out/ncu_target.log 2>&1
tail -3 gpurun_out/ncu_target.log
ls -la gpurun_out/*.ncu-rep
