"""Few eager C2 SMI steps (for ncu): python scripts/prof_c2.py [steps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import smi_experiments_b200 as smi
from bench import make_uci, C2
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
xt, yt = make_uci()
dev = torch.device("cuda")
model = smi.uci_bnn(13)
eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                  smi.RBFKernel(), num_stein_particles=C2["m"], num_elbo_particles=C2["S"], use_cuda_graph=False)
xd, yd = torch.from_numpy(xt).to(dev), torch.from_numpy(yt).to(dev)
st = eng.init(smi.PRNGKey(1), x=xd, y=yd, subsample=100)
for _ in range(steps):
    st, loss = eng.update(st, x=xd, y=yd, subsample=100)
torch.cuda.synchronize()
print("loss", float(loss))
