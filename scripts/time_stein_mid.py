"""Stein force at the weak-scaling particle counts: CUDA-core mid path against the tcgen05 path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from smi_experiments_b200 import ops
g = torch.Generator(device="cuda").manual_seed(0)
for m, d in ((40, 6604), (80, 6604), (160, 6604), (256, 6604), (160, 100000)):
    X = torch.randn(m, d, device="cuda", generator=g); G = torch.randn(m, d, device="cuda", generator=g)
    for name, path in (("small/mid", ops.PATH_SMALL), ("tc", ops.PATH_TC)):
        for _ in range(3):
            ops.stein_force(X, G, path=path)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20):
            F, loss, h = ops.stein_force(X, G, path=path)
        b.record(); torch.cuda.synchronize()
        print("m=%4d d=%6d %-9s %8.1f us  loss %.6g h %.6g" % (m, d, name, 1e3 * a.elapsed_time(b) / 20, float(loss), float(h)))
import smi_experiments_b200 as smi
from bench import profile_report
lib = smi._lib.load()
for m, d in ((80, 6604), (160, 6604), (256, 6604)):
    X = torch.randn(m, d, device="cuda", generator=g); G = torch.randn(m, d, device="cuda", generator=g)
    lib.smi_profile(1)
    for _ in range(5):
        ops.stein_force(X, G, path=ops.PATH_SMALL)
    torch.cuda.synchronize()
    lib.smi_profile(0)
    prof = profile_report(lib)
    print(m, d, {k: round(1e3 * t / 5, 1) for k, (c, t) in prof.items()})
