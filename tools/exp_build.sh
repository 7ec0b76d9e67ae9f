#!/bin/bash
# ablation builds of the forward kernel: build/exp/libnerf_b200_exp<N>.so with -DNB_EXP=<N> (see mlp_tc.cu)
set -e
cd "$(dirname "$0")/../nerf_shared_b200/csrc"
mkdir -p ../../build/exp
for n in "$@"; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
     --expt-relaxed-constexpr -DNB_EXP=$n -c mlp_tc.cu -o ../../build/exp/mlp_tc_exp$n.o
  objs=$(ls ../../build/obj/*.o | grep -v mlp_tc.o)
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build/exp/libnerf_b200_exp$n.so $objs ../../build/exp/mlp_tc_exp$n.o
done
