"""Host-side breakdown of an end-to-end C2 step (host x/y -> H2D -> step -> float(loss))."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import smi_experiments_b200 as smi
from bench import make_uci, C2
xt, yt = make_uci()
model = smi.uci_bnn(13)
eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                  smi.RBFKernel(), num_stein_particles=C2["m"], num_elbo_particles=C2["S"])
st = eng.init(smi.PRNGKey(1), x=xt, y=yt, subsample=100)
kw = dict(subsample=100)
for _ in range(10):
    st, loss = eng.update(st, x=xt, y=yt, **kw); float(loss)
N = 300
T = dict(bind=0.0, alloc=0.0, replay=0.0, item=0.0, total=0.0)
torch.cuda.synchronize()
t00 = time.perf_counter()
for _ in range(N):
    t0 = time.perf_counter()
    s2 = eng._own(st); data = eng._bind((), dict(x=xt, y=yt, **kw))
    t1 = time.perf_counter()
    b = eng._alloc(data)
    t2 = time.perf_counter()
    eng._graph.replay()
    t3 = time.perf_counter()
    v = float(b["loss"])
    t4 = time.perf_counter()
    T["bind"] += t1 - t0; T["alloc"] += t2 - t1; T["replay"] += t3 - t2; T["item"] += t4 - t3
T["total"] = time.perf_counter() - t00
print({k: round(1e6 * v / N, 1) for k, v in T.items()}, "us per step")
# public path
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(N):
    st, loss = eng.update(st, x=xt, y=yt, **kw); float(loss)
print("update()+float(loss): %.1f us per step" % (1e6 * (time.perf_counter() - t0) / N))
torch.cuda.synchronize(); t0 = time.perf_counter()
xd, yd = torch.from_numpy(xt).cuda(), torch.from_numpy(yt).cuda()
for _ in range(5):
    st, loss = eng.update(st, x=xd, y=yd, **kw)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(N):
    st, loss = eng.update(st, x=xd, y=yd, **kw)
torch.cuda.synchronize()
print("device-resident, no per-step sync: %.1f us per step" % (1e6 * (time.perf_counter() - t0) / N))
