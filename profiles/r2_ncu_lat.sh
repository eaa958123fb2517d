#!/bin/bash
# ncu --set full of the latency-mode kernel (one problem, two warps), with the source page
out=gpurun_out/r2ncul; mkdir -p $out
python profiles/run_solve.py 1 3 walking > $out/run.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:srbd_mpc_solve -s 2 -c 1 -o $out/lat -f python profiles/run_solve.py 1 3 walking > $out/ncu.log 2>&1
ncu -i $out/lat.ncu-rep --page raw --csv > $out/lat_raw.csv 2>/dev/null
ncu -i $out/lat.ncu-rep --page source --print-source cuda,sass --csv > $out/lat_source.csv 2>/dev/null
rm -f $out/lat.ncu-rep; tail -1 $out/run.log; ls -la $out
