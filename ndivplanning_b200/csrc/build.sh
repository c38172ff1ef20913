#!/bin/bash
# Builds libndivplan.so in-tree for sm_100a. Usage: build.sh [extra nvcc flags]
set -e
cd "$(dirname "$0")"
OUT=../libndivplan.so
SRCS="abi.cu k1_rollout_f32.cu k1_rollout_tc.cu k2_ndiv.cu k2_ndiv_tc.cu k3_select.cu k4_gan_train.cu k5_encoder.cu k5_encoder_tc.cu comm.cu"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
mkdir -p ../_build
OBJS=""
pids=""
for f in $SRCS; do
  o=../_build/${f%.cu}.o
  OBJS="$OBJS $o"
  if [ ! -f $o ] || [ $f -nt $o ] || [ common.cuh -nt $o ] || [ planner.cuh -nt $o ] || [ tc_common.cuh -nt $o ] || [ peer.cuh -nt $o ] || [ encoder_tc.cuh -nt $o ] || [ ../../include/ndivplan.h -nt $o ]; then
    $NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC \
      -Xptxas -v "$@" -c $f -o $o > ../_build/${f%.cu}.ptxas.log 2>&1 &
    pids="$pids $!"
  fi
done
for p in $pids; do wait $p || { cat ../_build/*.ptxas.log | grep -i -B2 -A8 error; exit 1; }; done
$NVCC -shared -o $OUT $OBJS -lcudart -ldl
echo built $OUT
