#!/bin/bash
# Builds the standalone probe binary (tools/b2_probe.cu) against the csrc objects.
set -e
cd "$(dirname "$0")/.."
python "video-cycle_gan-upscaling_b200/upscaler/_build.py" > /dev/null
O="video-cycle_gan-upscaling_b200/csrc/build"
mkdir -p build
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -c tools/b2_probe.cu -o build/b2_probe.o
nvcc -gencode arch=compute_100a,code=sm_100a build/b2_probe.o $O/b2_host.o $O/b2_conv_direct.o $O/b2_conv_igemm.o $O/b2_conv_halo.o $O/b2_conv_stack.o -o build/b2_probe
echo built build/b2_probe
