# Occupancy experiment on the dense kernel (experiments build: RTB_BUILD_EXP=1 opencl_ray_tracer_b200/build.sh):
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash tools/occ_sweep.sh'
export RTB_LIB=$PWD/opencl_ray_tracer_b200/librtb_exp.so
for x in 0 26000 36000 55000 110000; do echo "extra_smem $x"; RTB_DENSE_EXTRA_SMEM=$x python tools/dense_sweep.py --child --workloads c3 --worlds 1,4,8 | cut -c1-220; done
unset RTB_LIB
echo product; python tools/dense_sweep.py --child --workloads c3,c4 --worlds 1,8 | cut -c1-220
