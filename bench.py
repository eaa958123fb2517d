#!/usr/bin/env python
"""bench.py -- headline benchmark of the batched SRBD-MPC hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one pass of the hot path (MPC::updateMPC, mpc.cpp:448-473: build + QP solve + rollout) over
one batch of 1,024 synthetic walking-gait problems (BASELINE.json configs[1]) per GPU, one fused kernel
launch.  N>1: one rank per GPU under torchrun, every rank owns an independent batch (weak scaling), no
data-path collective; one NCCL all_gather of the first-step results after the timed region is timed
separately (`gather_ms`).  Prints ONE JSON line on rank 0.

`--impl reference` times the CPU restatement of the reference path (oracle/srbd_mpc_oracle.c: build as
coded + OSQP-style ADMM with the reference's default settings) on all host threads; the reference's own
C++ (needs Eigen/OSQP/OsqpEigen) cannot be built on this image and /root/reference does not exist on
the GPU box, so kind = "port".
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
# dram__bytes_read.sum + dram__bytes_write.sum of one solve launch at B=1024 from the committed `ncu --set full`
# capture (profiles/r1_v8_ncu_full_summary.csv: 2.827520 MB read, 0 B written): inputs are read once, the 2 MB of
# outputs stay in L2.
NCU_TRAFFIC_BYTES = 2827520
WORKLOAD = "configs[1]: 1,024 independent G1 SRBD MPC problems per GPU (cpp_node preset, N=10, nc=4, walking gait schedule)"


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


def run_reference(args, rank, world):
    if rank != 0:
        return
    from g1_locomotion_b200 import presets, inputs
    from oracle import oracle
    oracle.build_lib()
    preset = presets.get_preset("cpp_node")
    sample_B = 256
    batch = [a[:sample_B] for a in inputs.walking_batch(preset, sample_B)]
    cores = host_threads()         # torchrun exports OMP_NUM_THREADS=1: the reference arm uses every host thread anyway
    for _ in range(max(args.warmup, 1)):
        cpu_reference_pass(preset, batch, cores)
    times = [cpu_reference_pass(preset, batch, cores) for _ in range(args.steps)]
    total = float(np.sum(times))
    value = sample_B * args.steps / total
    sample = f"{sample_B} of the 1,024 walking problems per step, {args.steps} steps, OpenMP over {cores} threads"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "cpu_impl": "oracle/srbd_mpc_oracle.c (mpc.cpp restatement + OSQP-style ADMM, defaults eps 1e-3)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the config-3 throughput and B=1 latency extras")
    ap.add_argument("--config5", action="store_true",
                    help="BASELINE.json configs[4] instead of the headline: 2^20 random problems cut into contiguous shards over the "
                         "ranks (strong scaling), generated on the device, one result gather; prints its own JSON line")
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
        run_config5(args, solver, preset, dev, rank, world)
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

    for _ in range(max(args.warmup, 3)):
        _, _, info = step()
    torch.cuda.synchronize()
    assert int((info["status"] != 0).sum()) == 0, "unsolved problems in the bench batch"
    iters_mean = float(info["iters"].double().mean())

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

    tot = torch.tensor([dev_ms, e2e_s, gather_ms or 0.0, e2e_sync_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_s_max, gather_ms_max, e2e_sync_s_max = [float(v) for v in tot.tolist()]
    value = world * BATCH * args.steps / (dev_ms_max * 1e-3)
    e2e_value = world * BATCH * e2e_pipe_steps / e2e_s_max
    e2e_sync_value = world * BATCH * e2e_steps / e2e_sync_s_max

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
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": WORKLOAD, "batch_per_gpu": BATCH, "preset": "cpp_node", "l2": "flushed between timed steps (256 MiB fill)",
                           "ipm_iters_mean": iters_mean, "kernel": solver.kernel_info(BATCH)},
                "roofline": {"bound": "fp64", "achieved": ach_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": ach_tf / peak_tf if peak_tf else None,
                             "traffic": NCU_TRAFFIC_BYTES, "kernel": "v2::srbd_mpc_solve_warp_kernel<10,4,16>", "flops_per_problem": fl / BATCH,
                             "peak_source": "g1mpc_fp64_peak_tflops DFMA microbenchmark, measured in this run (MEASURED_PEAKS.json has no FP64 entry)",
                             "hbm": {"bound": "hbm", "achieved": by / (per_launch_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                     "frac": by / (per_launch_ms * 1e-3) / 1e9 / hbm_peak, "bytes_per_problem": by / BATCH,
                                     "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s"}},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 8 * BATCH * (13 + 13 * N + nc * N + 3 * nc * N + 3 * N + 13),
                        "d2h_bytes_per_step": BATCH * (8 * (3 * nc * N + 13 * N + 4) + 8), "steps": e2e_pipe_steps,
                        "api": "BatchedMPC.update_host_begin/_end -> g1mpc_update_host_begin/_end: %d batches in flight, caller "
                               "buffers page-locked (inputs copied host->device and results device->host every step)" % depth,
                        "sync": {"value": e2e_sync_value, "steps": e2e_steps,
                                 "api": "BatchedMPC.update_host -> g1mpc_update_host, one blocking call per step, pageable numpy buffers"}},
                "gpu_launches": args.steps, "clocks": clk.summary(), "wall_s_timed_region": t_wall,
                "gather_ms": gather_ms_max if world > 1 else None}
        if not args.no_extras:
            line["extras"] = extras(solver, preset, dev)
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(preset, host_batch)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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
    steps, warm = min(args.steps, 5), 2
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
    if rank == 0:
        print(json.dumps({"metric": "config 5: 2^20-problem batch sharded across the GPUs of one box (build+solve/sec, device-timed)",
                          "value": total / (float(mx[0]) * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
                          "ms_per_step": float(mx[0]), "gather_ms": float(mx[1]), "value_with_gather": total / ((float(mx[0]) + float(mx[1])) * 1e-3),
                          "higher_is_better": True, "scaling": "strong", "dtype": "f64", "data": "synthetic",
                          "config": {"workload": "configs[4]: 1,048,576 random problems (config-3 distribution, seed 13 + rank per shard), "
                                                 "contiguous shards, inputs generated on the device (far larger than L2)",
                                     "problems_per_gpu": hi - lo, "gather_bytes": total * 200, "kernel": solver.kernel_info(hi - lo)},
                          "unsolved": int(t[2]), "ipm_iters_mean": float(t[3]) / total}), flush=True)


def _fp64_peak(index):
    import ctypes as C
    from g1_locomotion_b200 import _lib
    v = C.c_double()
    rc = _lib.load().g1mpc_fp64_peak_tflops(index, C.byref(v))
    return float(v.value) if rc == 0 else None


def extras(solver, preset, dev):
    """Config-3 throughput regime (65,536 random problems) and single-problem latency (p50 over 500 B=1 launches)."""
    import torch
    from g1_locomotion_b200 import inputs
    out = {}
    ins = list(inputs.random_batch(preset, 65536, seed=7, device=dev))
    for _ in range(2):
        _, _, info = solver.update_batch(*ins, x_current=ins[0], want_resid=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        solver.update_batch(*ins, x_current=ins[0], want_resid=False)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    out["config3_65536_random"] = {"problems_per_s": 65536 / (ms * 1e-3), "ms_per_launch": ms, "unsolved": int((info["status"] != 0).sum()),
                                   "ipm_iters_mean": float(info["iters"].double().mean())}
    # config 4: horizon sweep with four corner contacts per foot (B = 4,096 / 1,024 / 256, seed 11)
    from g1_locomotion_b200 import BatchedMPC, presets as _presets
    sweep = {}
    for n_h, b_h in ((10, 4096), (20, 1024), (40, 256)):
        pr = _presets.get_preset("cpp_node", horizon=n_h, num_contacts=8)
        sv = BatchedMPC(pr, device=dev.index)
        ins_h = list(inputs.random_batch(pr, b_h, seed=11, device=dev))
        for _ in range(2):
            _, _, info_h = sv.update_batch(*ins_h, x_current=ins_h[0], want_resid=False)
        torch.cuda.synchronize()
        h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0.record()
        for _ in range(3):
            sv.update_batch(*ins_h, x_current=ins_h[0], want_resid=False)
        h1.record(); torch.cuda.synchronize()
        ms_h = h0.elapsed_time(h1) / 3
        sweep["N%d_nc8_B%d" % (n_h, b_h)] = {"problems_per_s": b_h / (ms_h * 1e-3), "ms_per_launch": ms_h, "unsolved": int((info_h["status"] != 0).sum()),
                                            "ipm_iters_mean": float(info_h["iters"].double().mean()), "kernel": sv.kernel_info(b_h)}
        sv.close()
    out["config4_horizon_sweep"] = sweep
    one = [t[:1].contiguous() for t in ins]
    lat = []
    for i in range(520):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); solver.update_batch(*one, x_current=one[0], want_resid=False); b.record(); b.synchronize()
        if i >= 20:
            lat.append(a.elapsed_time(b) * 1e3)
    # the same launch with the host kept ahead of the device (a 150 us spin kernel is queued first), so that the event pair
    # brackets the kernel alone and not the ~25 us of Python/ctypes between the first event and the launch
    dlat = []
    for i in range(220):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(300000)
        a.record(); _, _, info1 = solver.update_batch(*one, x_current=one[0], want_resid=False); b.record(); b.synchronize()
        if i >= 20:
            dlat.append(a.elapsed_time(b) * 1e3)
    out["single_solve_us"] = {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)), "launches": len(lat),
                              "timing": "CUDA events around one B=1 call of BatchedMPC.update_batch (host call overhead inside)",
                              "device_p50": float(np.percentile(dlat, 50)), "device_p99": float(np.percentile(dlat, 99)),
                              "device_timing": "same, launch queued behind a spin kernel: the kernel alone", "ipm_iters": int(info1["iters"][0])}
    out["next_rows"] = next_rows(solver, preset, dev)
    # config 1 (BASELINE.json configs[0]): the states of the straight-line walking demo, one B=1 call per control tick as
    # the node makes them (mpc_to_wbid_node.cpp:494); each tick timed once behind the spin kernel (kernel alone) and once bare
    ticks = [torch.as_tensor(a, dtype=torch.float64, device=dev) for a in inputs.walking_replay(preset, n_ticks=200)]
    rk, rh, rw, rit = [], [], [], []
    for rep in range(6):
        for t in range(ticks[0].shape[0]):
            tick = [a[t:t + 1] for a in ticks]
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda._sleep(300000)
            a.record(); _, _, info_t = solver.update_batch(*tick, x_current=tick[0], want_resid=False); b.record(); b.synchronize()
            c, d = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            w0 = time.perf_counter()
            c.record(); solver.update_batch(*tick, x_current=tick[0], want_resid=False); d.record(); d.synchronize()
            w1 = time.perf_counter()
            if rep > 0:
                rk.append(a.elapsed_time(b) * 1e3); rh.append(c.elapsed_time(d) * 1e3); rw.append((w1 - w0) * 1e6)
                rit.append(int(info_t["iters"][0]))
    out["single_solve_walking_replay_us"] = {
        "workload": "configs[0]: 200 consecutive ticks of the straight-line walking demo, one problem per call, 5 timed passes after one warm-up pass",
        "device_p50": float(np.percentile(rk, 50)), "device_p99": float(np.percentile(rk, 99)), "device_min": float(np.min(rk)),
        "p50": float(np.percentile(rh, 50)), "p99": float(np.percentile(rh, 99)), "solves": len(rk),
        "wall_p50": float(np.percentile(rw, 50)), "wall_p99": float(np.percentile(rw, 99)),
        "timing": "device_*: CUDA events, launch queued behind a spin kernel (kernel alone); p50/p99: CUDA events around the bare "
                  "Python call; wall_*: host clock from the call to the end of the event synchronize (launch and sync included)",
        "ipm_iters_p50": float(np.percentile(rit, 50)), "ipm_iters_max": int(np.max(rit)), "ipm_iters_mean": float(np.mean(rit))}
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
                          "note": "walk_inputs -> solve (x_current = x0) -> advance per tick, no host round trip; state advances across calls"}
    return out


def cpu_baseline(preset, host_batch):
    from oracle import oracle
    oracle.build_lib()
    cores = host_threads()
    sample = [a[:256] for a in host_batch]
    cpu_reference_pass(preset, [a[:32] for a in host_batch], cores)
    t1 = cpu_reference_pass(preset, [a[:32] for a in host_batch], 1) / 32
    reps, t_all = 0, 0.0
    while t_all < 2.0 and reps < 50:
        t_all += cpu_reference_pass(preset, sample, cores)
        reps += 1
    return {"value": 256 * reps / t_all, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"first 256 of the 1,024 walking problems x {reps} passes, OpenMP over {cores} host threads "
                      f"(oracle/srbd_mpc_oracle.c: mpc.cpp build as coded + OSQP-style ADMM, eps 1e-3, cold start)",
            "single_core_ms_per_solve": 1e3 * t1}


if __name__ == "__main__":
    main()
