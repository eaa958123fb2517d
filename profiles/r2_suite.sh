#!/bin/bash
# full GPU suite + bench (both arms) -> gpurun_out/<tag>/
tag=${1:-suite}; out=gpurun_out/$tag; mkdir -p $out
( time python -m pytest tests -m gpu -q --durations=8 ) 2>&1 | grep -E "^E  |FAILED|passed|failed|Error|^real|s call" | cut -c1-300 | head -80 > $out/pytest.log; cat $out/pytest.log
python bench.py --steps 20 --warmup 5 > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -5 $out/bench.err
python bench.py --impl reference --steps 5 --warmup 1 > $out/bench_ref.json 2>> $out/bench.err
python - <<PY
import json
d = json.load(open("$out/bench.json"))
ex = d.pop("extras", {})
print("value", d["value"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "cpu", d.get("cpu_baseline", {}).get("value"))
print("e2e", d["e2e"]["value"], d["e2e"]["sync"]["value"], d["e2e"]["per_rank"])
for k, v in ex.items():
    if k.startswith("single") or k.startswith("mpc_"):
        print(k, {a: b for a, b in v.items() if a in ("device_p50", "p50", "wall_p50", "device_p99", "wall_p99", "warm_start_attribute")})
    elif k == "python_default_N15_32768_random":
        print(k, round(v["problems_per_s"]), v["status_counts"], v["kernel"]["mode"])
    elif k == "config4_horizon_sweep":
        for a, b in v.items(): print("  ", a, round(b["problems_per_s"]), b["status_counts"], b["kernel"])
    elif k == "next_rows":
        for a, b in v.items(): print("  ", a, {c: e for c, e in b.items() if c in ("problems_per_s", "walker_ticks_per_s", "ipm_iters_mean_per_tick", "status_any_values")})
    else:
        print(k, json.dumps(v)[:300])
PY
