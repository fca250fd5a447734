#!/bin/bash
# Builds liburlg_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
ARCH="-gencode arch=compute_100a,code=sm_100a"
COMMON="-O3 -std=c++17 -lineinfo -Xcompiler -fPIC $ARCH"
mkdir -p build
# env kernels: mask-deciding expressions are written with never-contracted __f*_rn intrinsics (see env_kernels.cu)
$NVCC $COMMON -c env_kernels.cu -o build/env_kernels.o "$@"
for f in tc_gemm storage_kernels policy_kernels ppo_kernels lstm_kernels lstm_tc mlp_fused peer_xchg; do
  if [ -f $f.cu ]; then $NVCC $COMMON -c $f.cu -o build/$f.o "$@"; fi
done
$NVCC -shared $ARCH -o liburlg_b200.so build/*.o -lcudart
echo "built $(pwd)/liburlg_b200.so"
