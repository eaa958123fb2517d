"""e2e pipelined throughput through the host API for several pipeline depths (experiment build with G1MPC_HOST_SLOTS = 8)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
BATCH = 1024
preset = presets.get_preset("cpp_node")
solver = BatchedMPC(preset, device=0)
host_batch = [np.ascontiguousarray(a) for a in inputs.walking_batch(preset, BATCH)]
hin = []
for a in host_batch:
    b = solver.pinned(*a.shape); b[...] = a; hin.append(b)
for depth in (1, 2, 3, 4, 5, 6, 8):
    outs = [solver.out_buffers(BATCH) for _ in range(depth)]
    def pipelined(nsteps):
        for k in range(nsteps):
            sl = k % depth
            if k >= depth:
                u_k, x_k, inf_k = solver.update_host_end(sl); float(u_k[0, 0, 2])
            solver.update_host_begin(sl, *hin, x_current=hin[0], out=outs[sl])
        for k in range(max(nsteps - depth, 0), nsteps):
            u_k, x_k, inf_k = solver.update_host_end(k % depth); float(u_k[0, 0, 2])
    pipelined(2 * depth); torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        t0 = time.perf_counter(); pipelined(240); best = min(best, time.perf_counter() - t0)
    print("depth %d: %.4f ms per step, %.3f M problems/s" % (depth, best / 240 * 1e3, BATCH * 240 / best / 1e6))
