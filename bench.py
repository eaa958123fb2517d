#!/usr/bin/env python
"""bench.py -- headline benchmark of the batched SRBD-MPC hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one pass of the hot path (MPC::updateMPC, mpc.cpp:448-473: build + QP solve + rollout) over
one batch of 1,024 synthetic walking-gait problems (BASELINE.json configs[1]) per GPU, one fused kernel
launch.  N>1: one rank per GPU under torchrun, every rank owns an independent batch (weak scaling), no
data-path collective; one NCCL all_gather of the first-step results after the timed region is timed
separately (`gather_ms`).  Prints ONE JSON line on rank 0.

`--impl reference` times the CPU restatement of the reference path (oracle/srbd_mpc_oracle.c: build as
coded + OSQP-style ADMM with the reference's default settings, constraint matrix in CSR, vectorisable
Cholesky) on all host threads over the SAME 1,024 problems per step; the reference's own C++ (needs
Eigen/OSQP/OsqpEigen) cannot be built on this image and /root/reference does not exist on the GPU box,
so kind = "port".  The reference's own Python class (mpc.py + the osqp shim) was timed in the build
container by oracle/time_reference_py.py; its committed result is quoted as `cpu_baseline.reference_py`.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "batched MPC build+solve/sec (N=10, whole box, device-timed)"
UNIT = "problems/s"
BATCH = 1024
# roofline.traffic = dram__bytes_read.sum + dram__bytes_write.sum of one solve launch at B = 1,024, READ at run time from the
# newest committed `ncu --set full` summary of that launch (profiles/r*_ncu_full_summary.csv, written by
# profiles/ncu_summary.py); null when no summary is found or its kernel / grid does not match the launch being timed.
WORKLOAD = "configs[1]: 1,024 independent G1 SRBD MPC problems per GPU (cpp_node preset, N=10, nc=4, walking gait schedule)"


def bench_config():
    """`config` of the JSON line -- the same dict on both arms, so that the driver sees one configuration."""
    return {"workload": WORKLOAD, "batch_per_gpu": BATCH, "preset": "cpp_node", "horizon": 10, "num_contacts": 4,
            "problems_per_step": "1,024 per GPU (weak scaling)"}


def ncu_traffic(kernel_substr, grid):
    """(bytes, source file) from the newest committed ncu summary whose kernel name and grid match, else (None, reason)."""
    import csv
    import glob
    import re
    best = None
    for f in glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_summary*.csv")):
        m = re.match(r"r(\d+)_v?(\d+)", os.path.basename(f))
        key = (int(m.group(1)), int(m.group(2))) if m else (0, 0)
        try:
            rows = {r[0]: r for r in csv.reader(open(f)) if len(r) >= 3}
        except Exception:
            continue
        if kernel_substr not in rows.get("Kernel Name", ["", "", ""])[2] or rows.get("Grid Size", ["", "", ""])[2].replace(" ", "") != grid:
            continue
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        try:
            tot = sum(float(rows[k][2]) * unit[rows[k][1]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        except Exception:
            continue
        if best is None or key > best[0]:
            best = (key, tot, os.path.relpath(f, ROOT))
    return (best[1], best[2]) if best else (None, "no committed ncu summary matches this kernel and grid")


def iters_stats(iters):
    """p50 / p99 / max and the histogram of the interior-point iteration counts of one batch (SURVEY section 5 'Metrics')."""
    it = np.asarray(iters.cpu().numpy() if hasattr(iters, "cpu") else iters).astype(np.int64)
    h = np.bincount(it)
    return {"mean": float(it.mean()), "p50": float(np.percentile(it, 50)), "p99": float(np.percentile(it, 99)), "max": int(it.max()),
            "histogram": {str(k): int(v) for k, v in enumerate(h) if v}}


def algorithmic_flops(N, nc, iters_mean, stance_mean):
    """USEFUL FP64 flops (FMA = 2) of the stage-wise interior point per problem -- each quantity counted once,
    not once per lane that recomputes it; term by term in DESIGN.md section 5.
    stance_mean = mean number of stance contacts per problem (swing contacts are eliminated)."""
    per_stage_factor = 320 + 170 + 42 + 18 * 140 + 12 * 144 + 300     # closed-form P22^-1, LDL' of Pi+S, 18-column pipeline, S./P12. products, P update
    per_stage_solve = 330 + 290 + 190 + 130                            # pre, two sweeps, b_k, multiplier output (per solve)
    per_stage_grad = 100 + 9 * nc                                      # wrench, rollout, adjoint
    per_contact = 120 + 90 + 100 + 60 + 130 + 30                       # gradient/D-state, S_k share, predictor, corrector rhs, step, update
    setup = 70 * nc * N + 40 * N
    per_iter = N * (per_stage_factor + 2 * per_stage_solve + per_stage_grad + 150) + stance_mean * per_contact
    return setup + iters_mean * per_iter + N * per_stage_grad + stance_mean * 120


def algorithmic_bytes(N, nc):
    """Compulsory HBM bytes per problem: inputs + x_current in, u_opt/x_opt/status/iters/resid out."""
    return 8 * (13 + 13 * N + nc * N + 3 * nc * N + 3 * N + 13) + 8 * (3 * nc * N + 13 * N) + 8 + 32


class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU with NVML while the timed region runs."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop = [], set(), None, threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                try:
                    r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append(mhz)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self.ok:
            self.t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.ok:
            self.t.join(timeout=1.0)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_pass(preset, batch, nthreads):
    """One pass of the CPU restatement over `batch` -> seconds."""
    from oracle import oracle
    p = oracle.params_from_dict(preset)
    t0 = time.perf_counter()
    oracle.update_batch(p, *batch, x_current=batch[0], nthreads=nthreads)
    return time.perf_counter() - t0


def reference_py_record():
    """The reference's own Python class timed in the build container (oracle/time_reference_py.py) -- it cannot run on the GPU box."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r2_reference_py_cpu.json")))
        d["note"] = "measured in the build container, NOT on this box (/root/reference does not travel); quoted for scale only"
        return d
    except Exception:
        return None


def run_reference(args, rank, world):
    if rank != 0:
        return
    from g1_locomotion_b200 import presets, inputs
    from oracle import oracle
    oracle.build_lib()
    preset = presets.get_preset("cpp_node")
    batch = list(inputs.walking_batch(preset, BATCH))                       # the same 1,024 problems the GPU arm's rank 0 solves
    cores = host_threads()         # torchrun exports OMP_NUM_THREADS=1: the reference arm uses every host thread anyway
    for _ in range(max(min(args.warmup, 2), 1)):
        cpu_reference_pass(preset, batch, cores)
    times = [cpu_reference_pass(preset, batch, cores) for _ in range(args.steps)]
    total = float(np.sum(times))
    value = BATCH * args.steps / total
    t1 = cpu_reference_pass(preset, [a[:64] for a in batch], 1) / 64
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": bench_config(),
            "cpu_baseline": cpu_baseline_record(value, cores, f"all {BATCH} walking problems per step, {args.steps} steps, OpenMP over {cores} threads", t1),
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline_record(value, cores, sample, t1):
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
            "impl": "oracle/srbd_mpc_oracle.c: mpc.cpp build as coded + OSQP restatement (defaults: eps 1e-3, adaptive rho every 100 "
                    "iterations, cold start), CSR constraint matrix, one Cholesky per rho update",
            "single_core_ms_per_solve": 1e3 * t1,
            "accuracy_note": "stops at OSQP's eps 1e-3 (forces O(100 N) from the minimiser); the GPU arm solves to a gap of 1e-10",
            "reference_py": reference_py_record()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip configs 3/4/5, the latency legs and the next-row adaptors")
    ap.add_argument("--config5", action="store_true",
                    help="BASELINE.json configs[4] alone: 2^20 random problems cut into contiguous shards over the ranks (strong "
                         "scaling), generated on the device, one result gather; prints its own JSON line")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from g1_locomotion_b200 import BatchedMPC, presets, inputs, _lib

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL writes its version banner to stdout when the communicator comes up (NCCL_DEBUG=VERSION on the GPU boxes):
        # send file descriptor 1 to stderr until it has, so that rank 0's stdout carries the one JSON line only
        sys.stdout.flush()
        saved_out = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.all_reduce(torch.zeros(1, device=dev))
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_out, 1)
            os.close(saved_out)
    preset = presets.get_preset("cpp_node")
    N, nc = preset["horizon"], preset["num_contacts"]
    solver = BatchedMPC(preset, device=local_rank)
    if args.config5:
        line = run_config5(args, solver, preset, dev, rank, world)
        if rank == 0:
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    host_batch = inputs.walking_batch(preset, BATCH, seed0=1000 + 100000 * rank)
    ins = [torch.as_tensor(a, dtype=torch.float64).to(dev) for a in host_batch]
    u_opt = torch.empty(BATCH, N, 3 * nc, device=dev, dtype=torch.float64)
    x_opt = torch.zeros(BATCH, N, 13, device=dev, dtype=torch.float64)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def step():
        return solver.update_batch(*ins, x_current=ins[0], u_opt=u_opt, x_opt=x_opt)

    def barrier():
        if world > 1:
            dist.barrier()

    def allmax(vals):
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def allgather(val):
        t = torch.tensor([val], device=dev, dtype=torch.float64)
        if world == 1:
            return [float(val)]
        out = torch.empty(world, device=dev, dtype=torch.float64)
        dist.all_gather_into_tensor(out, t)
        return [float(v) for v in out.tolist()]

    for _ in range(max(args.warmup, 3)):
        _, _, info = step()
    torch.cuda.synchronize()
    assert int((info["status"] != 0).sum()) == 0, "unsolved problems in the bench batch"
    iters_mean = float(info["iters"].double().mean())
    it_stats = iters_stats(info["iters"])

    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    with ClockSampler(local_rank) as clk:
        barrier(); torch.cuda.synchronize()
        t_wall0 = time.perf_counter()
        for k in range(args.steps):
            flush.fill_(k & 0xff)              # evict L2 between timed iterations (outside the event pair)
            starts[k].record()
            step()
            ends[k].record()
        torch.cuda.synchronize(); barrier()
        t_wall = time.perf_counter() - t_wall0
    dev_ms = float(sum(s.elapsed_time(e) for s, e in zip(starts, ends)))
    per_launch_ms = dev_ms / args.steps

    # end to end through the reference-facing host API (numpy buffers; H2D + kernel + D2H inside the call)
    # (a) one synchronous call per step, pageable numpy buffers (staged through the handle's pinned block)
    for _ in range(3):
        solver.update_host(*host_batch, x_current=host_batch[0])
    barrier(); torch.cuda.synchronize()
    e2e_steps = min(args.steps, 50)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        uh, xh, _ = solver.update_host(*host_batch, x_current=host_batch[0])
    e2e_sync_s = time.perf_counter() - t0
    barrier()
    # (b) the same call in its two halves (g1mpc_update_host_begin/_end), HOST_SLOTS batches in flight, caller buffers in
    #     page-locked memory: every step still copies its inputs host->device and its results device->host
    depth = solver.HOST_SLOTS
    hin = []
    for a in host_batch:
        b = solver.pinned(*a.shape); b[...] = a; hin.append(b)
    outs = [solver.out_buffers(BATCH) for _ in range(depth)]

    def pipelined(nsteps):
        chk = 0.0
        for k in range(nsteps):
            sl = k % depth
            if k >= depth:
                u_k, x_k, inf_k = solver.update_host_end(sl)
                chk += float(u_k[0, 0, 2])                       # the step's result is read on the host
            solver.update_host_begin(sl, *hin, x_current=hin[0], out=outs[sl])
        for k in range(max(nsteps - depth, 0), nsteps):
            u_k, x_k, inf_k = solver.update_host_end(k % depth)
            chk += float(u_k[0, 0, 2])
        return chk, u_k, inf_k
    pipelined(2 * depth)
    barrier(); torch.cuda.synchronize()
    e2e_pipe_steps = 4 * e2e_steps
    t0 = time.perf_counter()
    _, u_p, inf_p = pipelined(e2e_pipe_steps)
    e2e_s = time.perf_counter() - t0
    barrier()
    assert int((inf_p["status"] != 0).sum()) == 0 and np.array_equal(u_p, uh), "pipelined host API differs from the synchronous call"
    # (c) where a step's time goes on THIS rank: the copies alone (page-locked buffers, CUDA events on the copy stream), the kernel
    #     alone (above), and what is left of a blocking step = host work (ctypes, staging memcpy, stream synchronise)
    cp = torch.cuda.Stream(device=dev)
    d_in = [torch.empty_like(t) for t in ins]
    h_in = [torch.as_tensor(a) for a in hin]
    h_out = [torch.as_tensor(outs[0]["u_opt"]), torch.as_tensor(outs[0]["x_opt"])]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    h2d_ms = d2h_ms = 0.0
    with torch.cuda.stream(cp):
        for rep in range(12):
            ev[0].record(cp)
            for d, h in zip(d_in, h_in):
                d.copy_(h, non_blocking=True)
            ev[1].record(cp)
            h_out[0].copy_(u_opt, non_blocking=True); h_out[1].copy_(x_opt, non_blocking=True)
            ev[2].record(cp)
            cp.synchronize()
            if rep >= 2:
                h2d_ms += ev[0].elapsed_time(ev[1]) / 10; d2h_ms += ev[1].elapsed_time(ev[2]) / 10
    barrier()

    # result collection over NCCL (first-step GRFs + x_opt[1], what the WBID layer consumes)
    gather_ms = None
    if world > 1:
        from g1_locomotion_b200 import sharding
        payload = sharding.wbid_payload(u_opt, x_opt)
        sharding.gather_results(payload, world * BATCH)
        torch.cuda.synchronize(); barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(); gathered = sharding.gather_results(payload, world * BATCH); g1.record(); torch.cuda.synchronize()
        assert gathered.shape[0] == world * BATCH
        gather_ms = float(g0.elapsed_time(g1))

    dev_ms_max, e2e_s_max, gather_ms_max, e2e_sync_s_max = allmax([dev_ms, e2e_s, gather_ms or 0.0, e2e_sync_s])
    per_rank = {"kernel_ms_per_step": allgather(per_launch_ms), "e2e_pipelined_ms_per_step": allgather(1e3 * e2e_s / e2e_pipe_steps),
                "e2e_blocking_ms_per_step": allgather(1e3 * e2e_sync_s / e2e_steps), "h2d_ms_per_step": allgather(h2d_ms),
                "d2h_ms_per_step": allgather(d2h_ms), "ipm_iters_max": allgather(it_stats["max"])}
    value = world * BATCH * args.steps / (dev_ms_max * 1e-3)
    e2e_value = world * BATCH * e2e_pipe_steps / e2e_s_max
    e2e_sync_value = world * BATCH * e2e_steps / e2e_sync_s_max

    # legs every rank takes part in: the 65,536-per-rank weak-scaling figure and BASELINE.json configs[4] (2^20 problems, strong scaling)
    collective = None
    if not args.no_extras:
        collective = {"weak_65536_per_gpu": run_weak_65536(solver, preset, dev, rank, world),
                      "config5": run_config5(args, solver, preset, dev, rank, world)}

    if rank == 0:
        stance_mean = float((ins[2] > 0).double().sum() / BATCH)
        fl = algorithmic_flops(N, nc, iters_mean, stance_mean) * BATCH
        peak_tf = _fp64_peak(local_rank)
        ach_tf = fl / (per_launch_ms * 1e-3) / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        by = algorithmic_bytes(N, nc) * BATCH
        kinfo = solver.kernel_info(BATCH)
        ppw = kinfo["problems_per_warp"]
        kname = "srbd_mpc_solve_warp_kernel<v2::WLay<10, 4>, %d>" % (32 // ppw)
        traffic, traffic_src = ncu_traffic("srbd_mpc_solve_warp_kernel", "(%d,1,1)" % ((BATCH + ppw - 1) // ppw))
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": bench_config(),
                "timing": {"l2": "flushed between timed steps (256 MiB fill, outside the event pair)", "clock": "CUDA events on the launching stream, max over ranks"},
                "solver": {"ipm_iters": it_stats, "kernel": kinfo, "gap_tol": 1e-10, "dual_tol": 1e-6},
                "roofline": {"bound": "fp64", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf if peak_tf else None,
                             "traffic": traffic, "traffic_source": traffic_src, "kernel": kname, "flops_per_problem": fl / BATCH,
                             "peak_source": "g1mpc_fp64_peak_tflops DFMA microbenchmark, measured in this run (MEASURED_PEAKS.json has no FP64 entry)",
                             "hbm": {"bound": "hbm", "achieved": by / (per_launch_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                     "frac": by / (per_launch_ms * 1e-3) / 1e9 / hbm_peak, "bytes_per_problem": by / BATCH,
                                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s"}},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 8 * BATCH * (13 + 13 * N + nc * N + 3 * nc * N + 3 * N + 13),
                        "d2h_bytes_per_step": BATCH * (8 * (3 * nc * N + 13 * N + 4) + 8), "steps": e2e_pipe_steps,
                        "api": "BatchedMPC.update_host_begin/_end -> g1mpc_update_host_begin/_end: %d batches in flight, caller "
                               "buffers page-locked (inputs copied host->device and results device->host every step)" % depth,
                        "sync": {"value": e2e_sync_value, "steps": e2e_steps,
                                 "api": "BatchedMPC.update_host -> g1mpc_update_host, one blocking call per step, pageable numpy buffers"},
                        "per_rank": per_rank,
                        "breakdown_note": "per rank and step: copies alone (page-locked, CUDA events on a copy stream), kernel alone, one blocking "
                                          "step, one pipelined step; blocking - (h2d + kernel + d2h) = host work of the call (ctypes, staging "
                                          "memcpy of 5 MB, synchronise), which is what the ranks of one box compete for"},
                "gpu_launches": args.steps, "clocks": clk.summary(), "wall_s_timed_region": t_wall,
                "gather_ms": gather_ms_max if world > 1 else None}
        if not args.no_extras:
            line["extras"] = extras(solver, preset, dev)
            line["extras"].update(collective)
            if world == 1:
                line["extras"]["config3_65536_random"] = collective["weak_65536_per_gpu"]      # configs[2]: the same launch
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(preset, host_batch)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_weak_65536(solver, preset, dev, rank, world):
    """65,536 random problems PER RANK (config-3 distribution, seed 7 + rank), one launch, max over ranks: the weak-scaling
    figure of the throughput regime beside the one-wave 1,024-problem headline."""
    import torch
    import torch.distributed as dist
    from g1_locomotion_b200 import inputs
    B = 65536
    ins = list(inputs.random_batch(preset, B, seed=7 + rank, device=dev))
    for _ in range(2):
        _, _, info = solver.update_batch(*ins, x_current=ins[0], want_resid=False)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        solver.update_batch(*ins, x_current=ins[0], want_resid=False)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 3, float((info["status"] != 0).sum())], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out = {"problems_per_s": world * B / (float(t[0]) * 1e-3), "ms_per_launch": float(t[0]), "n_gpus": world, "unsolved_max_per_rank": int(t[1]),
           "note": "inputs (315 MB per rank) exceed the 126 MB L2"}
    if rank == 0:
        out["ipm_iters"] = iters_stats(info["iters"])
        out["kernel"] = solver.kernel_info(B)
    return out


def run_config5(args, solver, preset, dev, rank, world):
    """Config 5: 2^20 problems of the config-3 distribution, contiguous shard per rank generated on the
    device with the shard's seed (13 + rank) over sharding.shard_range, one solve launch per rank, then the 200 B/problem
    WBID payload gathered over NCCL (sharding.gather_results).  Strong scaling: the total is fixed, value = 2^20 / max over ranks."""
    import torch
    import torch.distributed as dist
    from g1_locomotion_b200 import inputs, sharding
    total = 1 << 20
    lo, hi = sharding.shard_range(total, rank, world)
    ins = list(inputs.random_batch(preset, hi - lo, seed=13 + rank, device=dev))
    N, nc = preset["horizon"], preset["num_contacts"]
    u_opt = torch.empty(hi - lo, N, 3 * nc, device=dev, dtype=torch.float64)
    x_opt = torch.empty(hi - lo, N, 13, device=dev, dtype=torch.float64)
    steps, warm = 3, 1
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(steps)]
    for k in range(warm + steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        if k >= warm:
            ev[k - warm][0].record()
        _, _, info = solver.update_batch(*ins, x_current=ins[0], u_opt=u_opt, x_opt=x_opt, want_resid=False)
        if k >= warm:
            ev[k - warm][1].record()
        gathered = sharding.gather_results(sharding.wbid_payload(u_opt, x_opt), total) if world > 1 else sharding.wbid_payload(u_opt, x_opt)
        if k >= warm:
            ev[k - warm][2].record()
    torch.cuda.synchronize()
    assert gathered.shape[0] == total
    solve_ms = sum(a.elapsed_time(b) for a, b, _ in ev) / steps
    gather_ms = sum(b.elapsed_time(c) for _, b, c in ev) / steps
    t = torch.tensor([solve_ms, gather_ms, float((info["status"] != 0).sum()), float(info["iters"].double().sum())], device=dev, dtype=torch.float64)
    mx = t.clone()
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    del ins, u_opt, x_opt, gathered
    torch.cuda.empty_cache()
    return {"metric": "config 5: 2^20-problem batch sharded across the GPUs of one box (build+solve/sec, device-timed)",
            "value": total / (float(mx[0]) * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": float(mx[0]), "gather_ms": float(mx[1]), "value_with_gather": total / ((float(mx[0]) + float(mx[1])) * 1e-3),
            "higher_is_better": True, "scaling": "strong", "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[4]: 1,048,576 random problems (config-3 distribution, seed 13 + rank per shard), "
                                   "contiguous shards, inputs generated on the device (far larger than L2)",
                       "problems_per_gpu": hi - lo, "gather_bytes": total * 200, "kernel": solver.kernel_info(hi - lo)},
            "unsolved": int(t[2]), "ipm_iters_mean": float(t[3]) / total}


def _fp64_peak(index):
    import ctypes as C
    from g1_locomotion_b200 import _lib
    v = C.c_double()
    rc = _lib.load().g1mpc_fp64_peak_tflops(index, C.byref(v))
    return float(v.value) if rc == 0 else None


def extras(solver, preset, dev):
    """Rank 0 only: config 4 (horizon sweep, nc = 8), single-problem latency (one random problem; the configs[0] replay cold and
    warm; the drop-in MPC.update() with host arrays), and the next-row adaptors."""
    import torch
    from g1_locomotion_b200 import inputs, BatchedMPC, MPC, presets as _presets
    out = {}
    # config 4: horizon sweep with four corner contacts per foot -- SURVEY 8d's batch (B = 4,096 / 1,024 / 256, seed 11) and a
    # machine-filling one (8 problems per SM-resident slot would not fit memory at N = 40: 16,384 / 8,192 / 4,096)
    sweep = {}
    for n_h, b_h, b_fill in ((10, 4096, 16384), (20, 1024, 8192), (40, 256, 4096)):
        pr = _presets.get_preset("cpp_node", horizon=n_h, num_contacts=8)
        sv = BatchedMPC(pr, device=dev.index)
        for tag, bb in (("B%d" % b_h, b_h), ("fill_B%d" % b_fill, b_fill)):
            ins_h = list(inputs.random_batch(pr, bb, seed=11, device=dev))
            for _ in range(2):
                _, _, info_h = sv.update_batch(*ins_h, x_current=ins_h[0], want_resid=False)
            torch.cuda.synchronize()
            h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            h0.record()
            for _ in range(3):
                sv.update_batch(*ins_h, x_current=ins_h[0], want_resid=False)
            h1.record(); torch.cuda.synchronize()
            ms_h = h0.elapsed_time(h1) / 3
            st = torch.bincount(info_h["status"], minlength=5).tolist()
            ki = sv.kernel_info(bb)
            sweep["N%d_nc8_%s" % (n_h, tag)] = {"problems_per_s": bb / (ms_h * 1e-3), "ms_per_launch": ms_h, "status_counts": st,
                                                "ipm_iters_mean": float(info_h["iters"].double().mean()), "kernel": ki}
            if ki["smem_bytes"] == 0:
                # global-workspace kernel: HBM traffic is the interesting number here (SURVEY 8d "report both") -- from the committed
                # ncu capture of this very launch (kernel name and grid must match), against the algorithmic bytes of the API arrays
                tr, src = ncu_traffic("WLay<%d, 8>" % n_h, "(%d,1,1)" % min(bb, 148 * ki["ctas_per_sm"]))
                alg = algorithmic_bytes(n_h, 8) * bb
                sweep["N%d_nc8_%s" % (n_h, tag)]["hbm"] = {
                    "traffic": tr, "traffic_source": src, "algorithmic_bytes": alg,
                    "achieved_GBps_on_traffic": None if tr is None else tr / (ms_h * 1e-3) / 1e9,
                    "note": "the per-warp workspace (stage factors + per-contact state) streams through L2/HBM: traffic >> algorithmic by design "
                            "of this variant; it buys eight resident warps per SM instead of one or two"}
            del ins_h
        sv.close()
    out["config4_horizon_sweep"] = sweep
    # the reference's OTHER preset (Python class defaults: N = 15, dt 0.04, mu 0.3, full inertia, cost scale 1; mpc.py:12-51) at a
    # machine-filling batch of the config-3 distribution -- what a caller who swaps the import of g1_mpc/scripts/mpc.py gets in bulk
    pr = _presets.get_preset("python_default")
    sv = BatchedMPC(pr, device=dev.index)
    ins_h = list(inputs.random_batch(pr, 32768, seed=11, device=dev))
    for _ in range(2):
        _, _, info_h = sv.update_batch(*ins_h, x_current=ins_h[0], want_resid=False)
    torch.cuda.synchronize()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record()
    for _ in range(3):
        sv.update_batch(*ins_h, x_current=ins_h[0], want_resid=False)
    h1.record(); torch.cuda.synchronize()
    ms_h = h0.elapsed_time(h1) / 3
    out["python_default_N15_32768_random"] = {"problems_per_s": 32768 / (ms_h * 1e-3), "ms_per_launch": ms_h,
                                              "status_counts": torch.bincount(info_h["status"], minlength=5).tolist(),
                                              "ipm_iters_mean": float(info_h["iters"].double().mean()), "kernel": sv.kernel_info(32768)}
    del ins_h
    sv.close()

    def timed_call(fn, spin):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if spin:
            torch.cuda._sleep(300000)       # host ahead of the device: the event pair brackets the kernel alone
        w0 = time.perf_counter()
        a.record(); r = fn(); b.record(); b.synchronize()
        return a.elapsed_time(b) * 1e3, (time.perf_counter() - w0) * 1e6, r

    pct = lambda v, q: float(np.percentile(v, q))
    # one fixed random problem (as in round 1)
    ins = list(inputs.random_batch(preset, 8, seed=7, device=dev))
    one = [t[:1].contiguous() for t in ins]
    lat, dlat = [], []
    for i in range(320):
        d, _, _ = timed_call(lambda: solver.solve_one(*one, x_current=one[0]), True)
        e, _, r = timed_call(lambda: solver.solve_one(*one, x_current=one[0]), False)
        if i >= 20:
            dlat.append(d); lat.append(e)
    out["single_solve_us"] = {"p50": pct(lat, 50), "p99": pct(lat, 99), "device_p50": pct(dlat, 50), "device_p99": pct(dlat, 99), "launches": len(lat),
                              "ipm_iters": int(r[2]["it"][0]),
                              "timing": "device_*: CUDA events, launch queued behind a spin kernel (the kernel alone); p50/p99: events around the bare "
                                        "BatchedMPC.solve_one call (Python + launch overhead inside)"}
    # configs[0]: the closed-loop replay of SURVEY 8d (tests/golden/config1_replay.npz, written by oracle/gen_config1_replay.py):
    # 200 consecutive ticks, one B = 1 call per tick as the node makes them (mpc_to_wbid_node.cpp:494)
    z = np.load(os.path.join(ROOT, "tests", "golden", "config1_replay.npz"))
    ticks = [torch.as_tensor(z[k], dtype=torch.float64, device=dev) for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")]
    T = ticks[0].shape[0]
    ustar = torch.as_tensor(z["u_star"], dtype=torch.float64, device=dev)
    for mode in ("cold", "warm"):
        rk, rh, rw, rit, err = [], [], [], [], 0.0
        prev = torch.zeros(1, preset["horizon"], 12, dtype=torch.float64, device=dev)
        for rep in range(4):
            for t in range(T):
                tick = [a[t:t + 1] for a in ticks]
                uw = prev if (mode == "warm" and t > 0) else None
                call = lambda: solver.solve_one(*tick, x_current=tick[0], u_warm=uw, warm_shift=1)
                keep = prev.clone()
                d, _, r = timed_call(call, True)
                if rep == 3:
                    err = max(err, float((r[0][0] - ustar[t]).abs().max()))
                prev.copy_(keep)            # the second, bare call starts from the same previous solution
                e, w, r = timed_call(call, False)
                prev.copy_(r[0])
                if rep > 0:
                    rk.append(d); rh.append(e); rw.append(w); rit.append(int(r[2]["it"][0]))
        out["single_solve_walking_replay_us" + ("" if mode == "cold" else "_warm")] = {
            "workload": "configs[0]: 200 consecutive ticks of the straight-line walking demo closed on the oracle's solution (SURVEY 8d generator, "
                        "seed 20251104), one problem per call, 3 timed passes after one warm-up pass" +
                        ("" if mode == "cold" else "; every tick after the first starts from the previous tick's solution shifted by one stage "
                                                   "(g1mpc_solve_batch_warm) -- the warm start the reference asks its solver for and never gets"),
            "device_p50": pct(rk, 50), "device_p99": pct(rk, 99), "device_min": float(np.min(rk)), "p50": pct(rh, 50), "p99": pct(rh, 99),
            "wall_p50": pct(rw, 50), "wall_p99": pct(rw, 99), "solves": len(rk), "max_force_error_vs_exact_minimiser_N": err,
            "ipm_iters": iters_stats(np.array(rit)),
            "timing": "device_*: CUDA events, launch queued behind a spin kernel (kernel alone); p50/p99: CUDA events around the bare "
                      "BatchedMPC.solve_one call; wall_*: host clock from the call to the end of the event synchronize (launch and sync included)"}
    # the kept API itself: MPC.update() of the drop-in class, numpy arrays in, numpy arrays out (H2D, kernel, D2H inside)
    m = MPC.cpp_node(device=dev.index)
    m.init_matrices()
    hz = {k: z[k] for k in ("x0", "x_ref", "contact", "c_horizon", "p_com")}
    upd = {}
    for mode in ("cold", "warm"):
        m.warm_start = mode == "warm"
        w = []
        for rep in range(3):
            for t in range(T):
                m.x0 = hz["x0"][t].reshape(13, 1)
                m.x_ref_hor = hz["x_ref"][t]
                w0 = time.perf_counter()
                m.update(hz["contact"][t], hz["c_horizon"][t], hz["p_com"][t], x_current=hz["x0"][t], one_rollout=False)
                w1 = time.perf_counter()
                if rep > 0:
                    w.append((w1 - w0) * 1e6)
        upd[mode] = {"p50": pct(w, 50), "p99": pct(w, 99), "calls": len(w)}
    m.warm_start = False
    out["mpc_update_host_us"] = {"api": "g1_locomotion_b200.MPC.update(contact, c_horizon, p_com_horizon, x_current) -- the reference's Python "
                                        "signature (mpc.py:360), host arrays in and out, wall clock around the call, configs[0] replay",
                                 "p50": upd["cold"]["p50"], "p99": upd["cold"]["p99"], "calls": upd["cold"]["calls"],
                                 "warm_start_attribute": upd["warm"]}
    out["next_rows"] = next_rows(solver, preset, dev)
    return out


def next_rows(solver, preset, dev):
    """SURVEY 8f rows on the device, each against the HBM roofline of its own algorithmic bytes: the input generator
    (g1mpc_walk_inputs_batch), the WBID output adaptor (g1mpc_wbid_refs_batch), and the closed loop of many walkers
    (g1mpc_closed_loop_batch: generator -> solve -> advance per tick)."""
    import torch
    B, N = 65536, preset["horizon"]
    f64 = dict(device=dev, dtype=torch.float64)
    gen = torch.Generator(device=dev); gen.manual_seed(5)

    def fleet(b, spread=512):
        x0 = torch.zeros(b, 13, **f64)
        x0[:, 3] = 0.05 + 0.35 * 0.05 * (torch.arange(b, device=dev) % spread)
        x0[:, 4], x0[:, 5], x0[:, 9], x0[:, 12] = 7.44339342e-05, 5.97983255e-01, 0.35, float(np.float32(-9.80665))
        x0[:, 0:12] += 0.003 * torch.randn(b, 12, generator=gen, **f64)
        cur = torch.zeros(b, 4, 3, **f64)
        cur[:, :, 0] = x0[:, 3:4] + torch.tensor([-0.0293, 0.1406, -0.0293, 0.1406], **f64)
        cur[:, :, 1] = torch.tensor([0.118508, 0.118508, -0.118508, -0.118508], **f64)
        t = ((torch.arange(b, device=dev) % 40).double() + 0.5) * preset["dt"]
        xp = torch.zeros(b, N, 13, **f64)
        xp[:, :, 3] = x0[:, 3:4] + 0.35 * preset["dt"] * torch.arange(N, device=dev)
        xp[:, :, 4], xp[:, :, 9] = 7.44339342e-05, 0.35
        return x0, cur.reshape(b, 12).contiguous(), t, xp

    def timed(fn, reps=20):
        for _ in range(3):
            r = fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            r = fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps, r

    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0)
    except Exception:
        peak = 6650.0
    out = {}
    x0, cur, t, xp = fleet(B)
    x_ref = torch.zeros(B, N, 13, **f64)
    ms, (x_ref, contact, c_hor, p_com) = timed(lambda: solver.walk_inputs(x0, cur, t, xp, x_ref=x_ref))
    by = 8 * (13 + 12 + 1 + 13 * N) + 8 * (7 * N + 4 * N + 12 * N + 3 * N)      # x0, contact points, time, previous x_opt in; 26 N doubles out
    out["walk_inputs"] = {"problems": B, "ms_per_launch": ms, "problems_per_s": B / (ms * 1e-3), "bytes_per_problem": by,
                          "hbm": {"achieved": B * by / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": B * by / (ms * 1e-3) / 1e9 / peak},
                          "note": "the call also allocates its three output tensors (torch.empty) inside the event pair"}
    u, x, info = solver.update_batch(x0, x_ref, contact, c_hor, p_com, x_current=x0, want_resid=False)
    feet = cur.reshape(B, 4, 3).contiguous()
    ms, _ = timed(lambda: solver.wbid_refs(u, x, feet, 35.0))
    by = 8 * (12 + 13 + 12) + 8 * 42          # first-step forces, x_opt[1], foot positions in; 42 doubles out
    out["wbid_refs"] = {"problems": B, "ms_per_launch": ms, "problems_per_s": B / (ms * 1e-3), "bytes_per_problem": by,
                        "hbm": {"achieved": B * by / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": B * by / (ms * 1e-3) / 1e9 / peak},
                        "note": "reads two rows of u_opt/x_opt per problem: 32-byte sectors touched are 2,104 B in, 336 B out"}
    Bw, T = 4096, 20
    # the walkers start where the demo starts (the node's first Raibert point is computed from x_opt = 0, i.e. near the world
    # origin); x_ref and x_opt are carried from call to call like the node's members
    x0, cur, t, _ = fleet(Bw, spread=8)
    xr, xo = torch.zeros(Bw, N, 13, **f64), torch.zeros(Bw, N, 13, **f64)
    ms, res = timed(lambda: solver.closed_loop(x0, cur, t, T, x_ref=xr, x_opt=xo, record=False), reps=3)
    out["closed_loop"] = {"walkers": Bw, "ticks": T, "ms_per_call": ms, "walker_ticks_per_s": Bw * T / (ms * 1e-3),
                          "launches_per_tick": 3, "unsolved_walkers": int((res["status_any"] != 0).sum()),
                          "status_any_values": torch.unique(res["status_any"]).tolist(), "total_ticks_simulated": 6 * T,
                          "ipm_iters_mean_per_tick": float(res["iters_sum"].double().mean()) / T,
                          "note": "walk_inputs -> solve (x_current = x0) -> advance per tick, no host round trip; state advances across calls"}
    # the same fleet with warm start across ticks (every tick after the first starts from the previous solution shifted by one stage)
    x0, cur, t, _ = fleet(Bw, spread=8)
    xr, xo = torch.zeros(Bw, N, 13, **f64), torch.zeros(Bw, N, 13, **f64)
    ms, res = timed(lambda: solver.closed_loop(x0, cur, t, T, x_ref=xr, x_opt=xo, record=False, warm_start=True), reps=3)
    out["closed_loop_warm"] = {"walkers": Bw, "ticks": T, "ms_per_call": ms, "walker_ticks_per_s": Bw * T / (ms * 1e-3),
                               "unsolved_walkers": int((res["status_any"] != 0).sum()),
                               "ipm_iters_mean_per_tick": float(res["iters_sum"].double().mean()) / T}
    # main.py's SRBD-only loop (planner -> mapping -> solve -> x0 <- x_opt[2]) on the device, cold and warm
    pl = torch.tensor([[1000.0, 0.25, 0.1, 0.2, 100.0]], **f64).repeat(Bw, 1).contiguous()
    pl[:, 1] += 0.05 * (torch.arange(Bw, device=dev) % 5)
    pl[:, 0] = 0.2 * (torch.arange(Bw, device=dev) % 3)
    for tag, warm in (("main_loop", False), ("main_loop_warm", True)):
        ms, res = timed(lambda: solver.main_loop(pl, T, total_duration=T * preset["dt"], warm_start=warm, record=False), reps=2)
        out[tag] = {"walkers": Bw, "ticks": T, "ms_per_call": ms, "walker_ticks_per_s": Bw * T / (ms * 1e-3), "launches_per_tick": 5,
                    "status_any_values": torch.unique(res["status_any"]).tolist(),
                    "ipm_iters_mean_per_tick": float(res["iters_sum"].double().mean()) / T}
    out["main_loop"]["note"] = ("g1mpc_main_loop_batch: main.py:200-326 as coded (x0 <- x_opt[2], rollout from x_opt[1], reference schedule of "
                                "testing_double_support); consecutive solutions differ by 100-250 N there, so the warm start falls back to cold")
    B2 = 65536
    plB = pl[:1].repeat(B2, 1).contiguous()
    tB = torch.rand(B2, generator=gen, **f64) * 20.0
    ms, _ = timed(lambda: solver.plan_gait(plB, tB))
    by = 8 * 6 + 8 * 16 * N
    out["plan_gait"] = {"problems": B2, "ms_per_launch": ms, "problems_per_s": B2 / (ms * 1e-3), "bytes_per_problem": by,
                        "hbm": {"achieved": B2 * by / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": B2 * by / (ms * 1e-3) / 1e9 / peak}}
    return out


def cpu_baseline(preset, host_batch):
    """The port on this box's host cores over the SAME 1,024 problems (bounded: passes until 3 s have been spent)."""
    from oracle import oracle
    oracle.build_lib()
    cores = host_threads()
    small = [a[:64] for a in host_batch]
    cpu_reference_pass(preset, small, cores)
    t1 = cpu_reference_pass(preset, small, 1) / 64
    reps, t_all = 0, 0.0
    while t_all < 3.0 and reps < 50:
        t_all += cpu_reference_pass(preset, list(host_batch), cores)
        reps += 1
    return cpu_baseline_record(BATCH * reps / t_all, cores, f"all {BATCH} walking problems x {reps} passes, OpenMP over {cores} host threads", t1)


if __name__ == "__main__":
    main()
