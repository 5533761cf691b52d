#!/bin/bash
# Builds compile-time variants of the Hessian kernel HERE (no GPU needed) as separate libraries under
# mast3r_slam_ysh_b200/variants/ (git-ignored, they travel with gpurun); time them on the box with
#   M3R_LIB_PATH=mast3r_slam_ysh_b200/variants/libm3r_<name>.so python tools/time_phases.py 32 8
# usage: tools/build_variants.sh name1 "<flags1>" name2 "<flags2>" ...
set -e
cd "$(dirname "$0")/../mast3r_slam_ysh_b200/csrc"
make -s -j8 > /dev/null
mkdir -p ../variants _var
FL="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -I../../include"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( nvcc $FL $flags -c gn_hessian.cu -o _var/gn_hessian_$name.o &&
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libm3r_$name.so _var/gn_hessian_$name.o \
      $(ls _obj/*.o | grep -v gn_hessian.o) -lcudart && echo "built $name [$flags]" ) &
done
wait
