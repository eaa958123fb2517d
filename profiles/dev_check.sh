#!/bin/bash
# development loop on the GPU box: parity on the bench shape with libdev.so, phase clocks with libtiming.so
G1MPC_B200_LIB=gpurun_in/libdev.so python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "walking_1024 or determinism or 65536 or cpp_node-over0" 2>&1 | tail -3
G1MPC_B200_LIB=gpurun_in/libtiming.so python profiles/phase_clocks.py 1 2>&1 | cat
G1MPC_B200_LIB=gpurun_in/libtiming.so python profiles/phase_clocks.py 1024 2>&1 | sed -n 1,9p
G1MPC_B200_LIB=gpurun_in/libftiming.so python profiles/phase_clocks.py 1 2>&1 | grep factor
G1MPC_B200_LIB=gpurun_in/libdev.so python profiles/dev_time.py 2>&1 | tail -2
