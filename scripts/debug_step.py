import sys
import numpy as np, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import smi_experiments_b200 as smi
from oracle import models as omodels, step as ostep, elbo as oelbo, stein as ostein
from test_gpu_step import _snap
I, N, B, m, S = 6, 50, 10, 4, 6
rng = np.random.default_rng(0)
x = rng.standard_normal((N, I)).astype(np.float32)
y = rng.standard_normal(N).astype(np.float32)
model = smi.uci_bnn(I)
eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                  smi.RBFKernel(), num_stein_particles=m, num_elbo_particles=S, use_cuda_graph=False)
state = eng.init(smi.PRNGKey(1), x=x, y=y, subsample=B)
om = omodels.uci_bnn(I)
rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / np.linalg.norm(b))
for it in range(5):
    snap = _snap(state)
    state, loss = eng.update(state, x=x, y=y, subsample=B)
    nk, att = ostep.draw_keys(snap["key"])
    eps, idx = ostep.draw_noise(att, m, S, om, "normal", N, B)
    b = eng._bufs
    print("it", it, "eps err", np.abs(b["eps"].cpu().numpy() - eps).max(), "idx eq", (b["idx"].cpu().numpy() == idx).all())
    G64, E64 = oelbo.elbo_grads(om, snap["psi"], b["eps"].cpu().numpy(), idx, x, y, N, 1.0, "normal", torch.float64)
    G32, E32 = oelbo.elbo_grads(om, snap["psi"].astype(np.float32), b["eps"].cpu().numpy(), idx, x, y, N, 1.0, "normal", torch.float32)
    Gg = b["G"].cpu().numpy()
    print("   G rel err gpu", rel(Gg, G64), " fp32 oracle", rel(G32, G64), " per particle", [rel(Gg[i], G64[i]) for i in range(m)])
    F64, l64, h64, _ = ostein.stein_force(snap["psi"], G64, dtype=np.float64)
    Fg, lg, hg, _ = ostein.stein_force(snap["psi"], Gg.astype(np.float64), dtype=np.float64)
    print("   loss gpu", float(loss), "oracle", float(l64), "oracle force on gpu G", float(lg), "h gpu", float(b["h"].item()), "h64", float(h64))
    # which coordinates of G differ most
    dG = np.abs(Gg - G64); i, k = np.unravel_index(dG.argmax(), dG.shape)
    print("   worst G coord", i, k, Gg[i, k], G64[i, k], "elbo", b["elbo"].cpu().numpy()[i], E64[i])
