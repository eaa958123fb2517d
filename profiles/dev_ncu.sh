#!/bin/bash
# source-level ncu capture of the development build (gpurun_in/libdev.so) -> gpurun_out/devncu/
export G1MPC_B200_LIB=gpurun_in/libdev.so
out=gpurun_out/devncu; mkdir -p $out
python profiles/run_solve.py 1024 2 walking > $out/run.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:srbd_mpc_solve -c 1 -o $out/dev -f python profiles/run_solve.py 1024 2 walking > $out/ncu.log 2>&1
ncu -i $out/dev.ncu-rep --page source --print-source cuda,sass --csv > $out/dev_source.csv 2>/dev/null
rm -f $out/dev.ncu-rep
ls -la $out
