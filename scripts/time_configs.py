"""Time one SMI/SVGD step of every BASELINE config (C1-C4) with the per-kernel breakdown of libsmi's
profiler:  python scripts/time_configs.py [c1 c2 c3 c4] [--reps N]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import smi_experiments_b200 as smi
from bench import make_uci, profile_report


def timed(name, eng, state, args, kwargs, reps=20):
    lib = smi._lib.load()
    for _ in range(3):
        state, loss = eng.update(state, *args, **kwargs)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        state, loss = eng.update(state, *args, **kwargs)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    # eager pass with per-kernel events
    eng2 = eng
    g = getattr(eng2, "use_cuda_graph", None)
    eng2.use_cuda_graph = False
    lib.smi_profile(1)
    for _ in range(5):
        state, loss = eng2.update(state, *args, **kwargs)
    torch.cuda.synchronize()
    lib.smi_profile(0)
    prof = profile_report(lib)
    eng2.use_cuda_graph = g
    print("%s: %.3f ms/step (graph replay, back to back)  loss %.4g" % (name, ms, float(loss)))
    tot = sum(v[1] for v in prof.values())
    for k, (cnt, t) in sorted(prof.items(), key=lambda kv: -kv[1][1]):
        print("    %-28s x%-3d %8.3f ms/step  %5.1f%%" % (k, cnt // 5, t / 5, 100 * t / tot))
    sys.stdout.flush()


def main():
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["c1", "c2", "c3", "c4"]
    dev = torch.device("cuda")
    rng = np.random.default_rng(0)
    if "c1" in which:
        for h in (5, 50, 100):
            model = smi.reg_1d_bnn(h)
            eng = smi.SteinVI(model, smi.AutoNormal(model), smi.Adam(1e-3), smi.RBFKernel(), num_stein_particles=5,
                              num_elbo_particles=100)
            x = torch.linspace(-1.5, 1.7, 40, device=dev)
            y = torch.sin(x)
            st = eng.init(smi.PRNGKey(0), x, y)
            timed("C1 sine BNN h=%d m=5 S=100" % h, eng, st, (x, y), {})
    if "c2" in which:
        xt, yt = make_uci()
        xd, yd = torch.from_numpy(xt).to(dev), torch.from_numpy(yt).to(dev)
        for m in (5, 20):
            model = smi.uci_bnn(13)
            eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                              smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=100)
            st = eng.init(smi.PRNGKey(1), x=xd, y=yd, subsample=100)
            timed("C2 UCI BNN m=%d S=100 B=100" % m, eng, st, (), dict(x=xd, y=yd, subsample=100))
    if "c3" in which:
        for d in (10, 100, 1000):
            model = smi.GaussianTarget(d)
            eng = smi.SVGD(model, smi.Adam(0.05), smi.RBFKernel(), num_stein_particles=20, repulsion_temperature=1.0,
                           guide_kwargs={"init_loc_fn": smi.init_to_uniform(radius=2.0)})
            st = eng.init(smi.PRNGKey(2))
            timed("C3 SVGD gauss d=%d m=20" % d, eng, st, (), {})
            eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=2.0)),
                              smi.Adagrad(0.05), smi.RBFKernel(), num_stein_particles=20, num_elbo_particles=100)
            st = eng.init(smi.PRNGKey(2))
            timed("C3 SMI gauss d=%d m=20 S=100" % d, eng, st, (), {})
    if "c4" in which:
        model = smi.mnist_mlp()
        B, n = 128, 60000
        for m, S in ((5, 55), (10, 55)):
            eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.05)),
                              smi.Adagrad(0.8), smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S)
            xb = torch.from_numpy(rng.integers(0, 256, (B, 784)).astype(np.float32) / 255.0).to(dev)
            yb = torch.from_numpy(rng.integers(0, 10, B).astype(np.int32)).to(dev)
            st = eng.init(smi.PRNGKey(4), xb, yb, n=n, batch_size=B)
            timed("C4 MNIST MLP m=%d S=%d B=128" % (m, S), eng, st, (xb, yb), dict(n=n, batch_size=B), reps=5)


if __name__ == "__main__":
    main()
