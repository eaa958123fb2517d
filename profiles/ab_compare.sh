#!/bin/bash
# gpurun_in/libbase.so (before) against gpurun_in/libdev.so (after): outputs bit for bit, timings side by side
out=gpurun_out/ab; mkdir -p $out
for v in base dev; do echo "== $v"; G1MPC_B200_LIB=gpurun_in/lib$v.so python profiles/ab_dump.py $out/$v.npz 2>&1 | tail -4; done
python - <<'PY'
import numpy as np
a, b = np.load("gpurun_out/ab/base.npz"), np.load("gpurun_out/ab/dev.npz")
bad = [k for k in a.files if not np.array_equal(a[k], b[k], equal_nan=True)]
print("keys", len(a.files), "differing:", bad)
for k in bad:
    if a[k].dtype.kind == "f":
        print(" ", k, "max abs diff %.3e" % np.nanmax(np.abs(a[k] - b[k])))
    else:
        print(" ", k, "n diff", int((a[k] != b[k]).sum()))
PY
if [ -f gpurun_in/libtiming.so ]; then G1MPC_B200_LIB=gpurun_in/libtiming.so python profiles/phase_clocks.py 1 2>&1 | cat; G1MPC_B200_LIB=gpurun_in/libtiming.so python profiles/phase_clocks.py 1024 2>&1 | cat; fi
