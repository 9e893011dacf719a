# Per-rank breakdown of the host-frame call at N GPUs (tools/e2e_breakdown_group.py), dense and binned:
#   /usr/local/graft/bin/gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_bd.sh "8 2"'
set -x
mkdir -p gpurun_out
for n in $1; do for m in dense binned; do
  timeout -k 5 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) tools/e2e_breakdown_group.py c3 $m 2>gpurun_out/bd.err | grep "^{" >> gpurun_out/e2e_breakdown_group.jsonl
done; done
cat gpurun_out/e2e_breakdown_group.jsonl; tail -3 gpurun_out/bd.err
