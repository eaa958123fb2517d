"""Kernel time of the (uninstrumented) development build: B=1 latency and B=1,024 walking batch, CUDA events.
    G1MPC_B200_LIB=gpurun_in/libdev.so python profiles/dev_time.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from g1_locomotion_b200 import BatchedMPC, presets, inputs
preset = presets.get_preset("cpp_node")
s = BatchedMPC(preset, device=0)
full = [torch.as_tensor(a, dtype=torch.float64).cuda() for a in inputs.walking_batch(preset, 1024)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for B, reps in ((1, 300), (1024, 100)):
    ins = [t[:B].contiguous() for t in full]
    for _ in range(10):
        s.update_batch(*ins, x_current=ins[0])
    ts = []
    for k in range(reps):
        if B > 1:
            flush.fill_(k & 0xff)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(300000)                   # host ahead of the device: the events bracket the kernel alone
        e0.record(); s.update_batch(*ins, x_current=ins[0]); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    print("B = %4d  p50 %.1f us  mean %.1f us  -> %.3f M problems/s" % (B, np.median(ts), np.mean(ts), B / np.median(ts)))
