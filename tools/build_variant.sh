#!/bin/bash
# build an A/B variant of the library: tools/build_variant.sh NAME -DMACRO=V ...   -> variants/NAME.so (use with DRONERL_B200_LIB)
name=$1; shift
mkdir -p variants
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --fmad=false -shared -Xcompiler -fPIC "$@" -o variants/$name.so dronerl_b200/csrc/dronerl_b200.cu 2>&1 | grep -E "error" ; ls -la variants/$name.so | awk '{print $5, $9}'
