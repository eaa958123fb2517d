#!/bin/bash
# ncu evidence of the final kernel set of round 2 (each capture only after its own command has exited 0 without ncu):
#   launch list of the bench command; --set full of the bench-shape kernel at 1,024 walking problems, of the hybrid-workspace
#   kernel at N = 20, nc = 8, B = 8,192 and of the two-per-warp hybrid kernel at N = 10, nc = 8, B = 4,096 (config 4)
out=gpurun_out/r2ncu_final; mkdir -p $out
python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extras > $out/bench_plain.json 2> $out/err.log && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extras > $out/ncu_launch.log 2>&1
cap() {  # name, command...
  name=$1; shift
  "$@" > $out/$name.run.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:srbd_mpc_solve -s 1 -c 1 -o $out/$name -f "$@" > $out/$name.ncu.log 2>&1
  ncu -i $out/$name.ncu-rep --page raw --csv > $out/${name}_raw.csv 2>/dev/null
  ncu -i $out/$name.ncu-rep --page source --print-source cuda,sass --csv > $out/${name}_source.csv 2>/dev/null
  rm -f $out/$name.ncu-rep
  tail -1 $out/$name.run.log
}
cap solve_1024 python profiles/run_solve.py 1024 2 walking
cap solve_n20nc8_hybrid python profiles/run_solve_shape.py 20 8 8192 2
cap solve_n10nc8_hybrid2 python profiles/run_solve_shape.py 10 8 4096 2
ls -la $out | head -30
