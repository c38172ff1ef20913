#!/bin/bash
# dev build with cycle counters in the rollout kernels: ndivplanning_b200/libndivplan_prof.so (needs the normal build first)
set -e
cd "$(dirname "$0")/../ndivplanning_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -DNDP_TC_PROF $K1_EXTRA -c k1_rollout_tc.cu -o ../_build/k1_rollout_tc_prof.o
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -DNDP_TC_PROF -c k2_ndiv_tc.cu -o ../_build/k2_ndiv_tc_prof.o
OBJS="../_build/k2_ndiv_tc_prof.o"
for f in abi k1_rollout_f32 k2_ndiv k3_select k4_gan_train k5_encoder k5_encoder_tc comm; do OBJS="$OBJS ../_build/$f.o"; done
$NVCC -shared -o ../${PROF_OUT:-libndivplan_prof.so} ../_build/k1_rollout_tc_prof.o $OBJS -lcudart -ldl
echo built ../${PROF_OUT:-libndivplan_prof.so}
