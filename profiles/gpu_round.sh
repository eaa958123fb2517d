#!/bin/bash
# One GPU-box pass: parity tests, bench (both arms), ncu launch list of the bench command, one --set full capture
# of the solve kernel (with source).  usage: bash profiles/gpu_round.sh <tag>
tag=${1:-run}
out=gpurun_out/$tag
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest.log
tail -3 $out/pytest.log
python bench.py > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"
cat $out/bench.json
python bench.py --impl reference --steps 5 --warmup 1 > $out/bench_ref.json 2>> $out/bench.err; cat $out/bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extras > $out/ncu_launch.log 2>&1
python profiles/run_solve.py 1024 2 walking > $out/run_solve.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:srbd_mpc_solve -c 1 -o $out/solve_full -f \
    python profiles/run_solve.py 1024 2 walking > $out/ncu_full.log 2>&1
ncu -i $out/solve_full.ncu-rep --page raw --csv > $out/solve_full_raw.csv 2>/dev/null
ncu -i $out/solve_full.ncu-rep --page source --print-source cuda,sass --csv > $out/solve_full_source.csv 2>/dev/null
ls -la $out
