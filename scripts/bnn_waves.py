"""Latency vs throughput of the per-draw BNN kernel: time it with 1, 2, 3, 4 CTAs per SM worth of draws."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import smi_experiments_b200 as smi
from bench import make_uci, profile_report
lib = smi._lib.load()
dev = torch.device("cuda")
xt, yt = make_uci()
xd, yd = torch.from_numpy(xt).to(dev), torch.from_numpy(yt).to(dev)
for m, S in ((4, 37), (8, 37), (12, 37), (16, 37), (20, 100)):
    model = smi.uci_bnn(13)
    eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                      smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S, use_cuda_graph=False)
    st = eng.init(smi.PRNGKey(1), x=xd, y=yd, subsample=100)
    for _ in range(3):
        st, loss = eng.update(st, x=xd, y=yd, subsample=100)
    torch.cuda.synchronize()
    lib.smi_profile(1)
    for _ in range(10):
        st, loss = eng.update(st, x=xd, y=yd, subsample=100)
    torch.cuda.synchronize()
    lib.smi_profile(0)
    prof = profile_report(lib)
    print("draws %5d: %s" % (m * S, "  ".join("%s %.1fus" % (k.replace("_kernel", ""), 1e3 * v[1] / v[0]) for k, v in prof.items())))
