#!/bin/bash
# full GPU suite + bench (both arms) -> gpurun_out/<tag>/
tag=${1:-suite}; out=gpurun_out/$tag; mkdir -p $out
python -m pytest tests -m gpu -q 2>&1 | grep -E "^E  |FAILED|passed|failed|Error" | cut -c1-300 | head -80 > $out/pytest.log; cat $out/pytest.log
python bench.py --steps 20 --warmup 5 > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -5 $out/bench.err
python bench.py --impl reference --steps 5 --warmup 1 > $out/bench_ref.json 2>> $out/bench.err; cat $out/bench_ref.json | cut -c1-600
python - <<PY
import json
d = json.load(open("$out/bench.json"))
ex = d.pop("extras", {})
print(json.dumps({k: d[k] for k in ("value", "ms_per_step", "roofline", "solver", "cpu_baseline") if k in d})[:1800])
print("e2e", d["e2e"]["value"], d["e2e"]["sync"]["value"], d["e2e"]["per_rank"])
for k, v in ex.items():
    print(k, json.dumps(v)[:900])
PY
