# Short tail of the dense launch on / off for one rank's rows of C3 / C4 at world 1, 2, 4, 8 (one GPU):
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash tools/tail_sweep.sh'
for t in 0 1; do echo "RTB_DENSE_TAIL=$t"; RTB_DENSE_TAIL=$t python tools/dense_sweep.py --child --workloads c3,c4 --worlds 1,2,4,8 | cut -c1-300; done
