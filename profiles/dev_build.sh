#!/bin/bash
# development builds of the bench shape only: gpurun_in/libdev.so (parity) and gpurun_in/libtiming.so (phase clocks)
set -e
cd "$(dirname "$0")/../g1_locomotion_b200/csrc"
F="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O2 -ccbin /usr/bin/g++ -DG1MPC_DEV_SHAPE -shared"
mkdir -p ../../gpurun_in
/usr/local/cuda/bin/nvcc $F -Xptxas -v -o ../../gpurun_in/libdev.so g1mpc_kernels.cu 2>&1 | grep -E "error|warning|Used 2[0-9][0-9]|spill stores" | grep -v " 0 bytes spill" || true
/usr/local/cuda/bin/nvcc $F -DG1MPC_PHASE_CLOCKS -o ../../gpurun_in/libtiming.so g1mpc_kernels.cu
/usr/local/cuda/bin/nvcc $F -DG1MPC_PHASE_CLOCKS -DG1MPC_FACTOR_CLOCKS -o ../../gpurun_in/libftiming.so g1mpc_kernels.cu
