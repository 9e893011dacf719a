#!/bin/bash
# In-tree build of the product binaries (they travel to the GPU box with the snapshot):
#   librtb.so       C ABI (include/rtb.h): sm_100a kernels + host packer
#   librtb_host.so  C++ host mirror (Cube, Scene, Random, PNG) with a small C surface for Python
#   raycast_cli     headless driver (RenderState + Mode::CUDA)
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
CXXFLAGS="-std=c++17 -O2 -ffp-contract=off -fPIC -Wall"
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17 \
  -diag-suppress 177 -Xcompiler -fPIC,-O2,-ffp-contract=off -shared \
  -lrt -lpthread -o librtb.so csrc/rtb_api.cu csrc/rtb_group.cu csrc/rtb_noalloc.cu csrc/raycast_kernels.cu csrc/scene_pack_device.cu csrc/scene_pack.cpp "$@"
if [ -n "$RTB_BUILD_EXP" ]; then  # A/B build for tools/dense_sweep.py: the round-1 dense kernel behind RTB_GRID_VARIANT
  $NVCC -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -std=c++17 -DRTB_EXPERIMENTS \
    -diag-suppress 177 -Xcompiler -fPIC,-O2,-ffp-contract=off -shared -lrt -lpthread \
    -o librtb_exp.so csrc/rtb_api.cu csrc/rtb_group.cu csrc/rtb_noalloc.cu csrc/raycast_kernels.cu csrc/scene_pack_device.cu csrc/scene_pack.cpp
fi
g++ $CXXFLAGS -shared -o librtb_host.so host/host_capi.cpp host/Scene.cpp host/Cube.cpp host/PngWriter.cpp \
  csrc/scene_pack.cpp
g++ $CXXFLAGS -o raycast_cli host/raycast_cli.cpp host/RenderState.cpp host/Scene.cpp host/Cube.cpp \
  host/PngWriter.cpp -L. -lrtb -Wl,-rpath,'$ORIGIN'
echo "built librtb.so librtb_host.so raycast_cli in $(pwd)"
