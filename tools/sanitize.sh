# compute-sanitizer over the small GPU tests (on the GPU box, from the repo root):
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/sanitize.sh r02'
# Logs: gpurun_out/<tag>_sanitizer_<tool>.txt (copy into profiles/).  RTB_SANITIZE=1 makes the tests pick small frames.
tag=${1:-r02}
mkdir -p gpurun_out
export RTB_SANITIZE=1
for tool in memcheck racecheck synccheck; do
  timeout -k 5 ${SAN_TIMEOUT:-420} compute-sanitizer --tool $tool --target-processes all --print-limit 20 \
    python -m pytest tests/test_gpu_parity.py tests/test_gpu_tile_cull.py tests/test_gpu_peer_frame.py -m gpu -q -k "not much_faster_at_c3" \
    > gpurun_out/${tag}_sanitizer_${tool}.txt 2>&1
  echo "rc=$?" >> gpurun_out/${tag}_sanitizer_${tool}.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|rc=" gpurun_out/${tag}_sanitizer_${tool}.txt | tail -5
done
