"""CPU baseline (A) of BASELINE.md section 3: the reference's OWN g1_mpc/scripts/mpc.py, unmodified, driven through
MPC.update(contact, c_horizon, p_com_horizon, x_current, one_rollout) (mpc.py:360-377), `osqp` supplied by
oracle/osqp_shim.py, one Python process per host core (OMP/MKL/OPENBLAS threads = 1), on the 1,024 walking problems of
BASELINE.json configs[1] with the cpp_node constants applied to the Python object (cost scale 1: the class has no x2).

    python -m oracle.time_reference_py [n_problems] [n_procs]  ->  profiles/r2_reference_py_cpu.json

TEST INFRASTRUCTURE: runs in the build container only -- /root/reference does not exist on the GPU box, and reference
sources are never copied into this repo, so bench.py cannot run this arm there; it quotes the committed JSON
(`cpu_baseline.reference_py`, labelled with the host it was measured on) beside the port it does run."""
import json
import os
import sys
import time

for v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
    os.environ[v] = "1"

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _worker(args):
    lo, hi, n_total = args
    from g1_locomotion_b200 import presets, inputs
    from oracle import ref_loader
    preset = presets.get_preset("cpp_node", cost_scale=1.0)
    batch = inputs.walking_batch(preset, n_total)
    ref = ref_loader.make_reference_mpc(preset)
    ref_loader.reference_update(ref, *[a[lo] for a in batch], x_current=batch[0][lo])      # warm-up (imports, BLAS)
    per = []
    t0 = time.perf_counter()
    for b in range(lo, hi):
        t = time.perf_counter()
        ref_loader.reference_update(ref, *[a[b] for a in batch], x_current=batch[0][b])
        per.append(time.perf_counter() - t)
    return time.perf_counter() - t0, per


def main():
    import multiprocessing as mp
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    procs = int(sys.argv[2]) if len(sys.argv) > 2 else len(os.sched_getaffinity(0))
    cuts = [(i * n // procs, (i + 1) * n // procs, n) for i in range(procs)]
    with mp.get_context("spawn").Pool(procs) as pool:
        t0 = time.perf_counter()
        res = pool.map(_worker, cuts)
        wall = time.perf_counter() - t0
    busy = max(r[0] for r in res)
    per = np.array(sum((r[1] for r in res), []))
    cpu = ""
    try:
        cpu = [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except Exception:
        pass
    out = {"kind": "reference", "impl": "g1_mpc/scripts/mpc.py MPC.update() unmodified + oracle/osqp_shim.py (OSQP restatement, defaults, cold start)",
           "workload": "configs[1]: %d walking problems, cpp_node constants on the Python class (cost scale 1)" % n,
           "host": "build container (no GPU): %d logical cores, %s" % (procs, cpu), "cores": procs,
           "value": n / busy, "unit": "problems/s", "ms_per_solve_single_core_p50": float(1e3 * np.median(per)),
           "ms_per_solve_single_core_p99": float(1e3 * np.percentile(per, 99)), "wall_s_incl_process_start": wall}
    path = os.path.join(ROOT, "profiles", "r2_reference_py_cpu.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
