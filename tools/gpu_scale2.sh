set -x
mkdir -p gpurun_out
timeout -k 5 600 python -m pytest tests/test_gpu_group.py tests/test_gpu_peer_frame.py -m gpu -x -q > gpurun_out/pytest_group.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_group.log; tail -4 gpurun_out/pytest_group.log
bash tools/gpu_scale.sh "$1" $2 $3
