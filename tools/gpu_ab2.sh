# A/B of the multi-GPU frame fence on an N-GPU box: RTB_GROUP_STREAM_SIGNAL = 0 (in-kernel signal) / 1 (stream write, the default)
#   /usr/local/graft/bin/gpurun --gpus 2 --timeout 600 -- 'bash tools/gpu_ab2.sh 2'
set -x
mkdir -p gpurun_out
for v in 0 1; do
  RTB_GROUP_STREAM_SIGNAL=$v timeout -k 5 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $((29700 + v)) \
    bench.py --gpus $1 --steps 20 --warmup 5 --no-c4 > gpurun_out/ab_signal_${v}_n$1.json 2> gpurun_out/ab.err
  tail -2 gpurun_out/ab.err
done
