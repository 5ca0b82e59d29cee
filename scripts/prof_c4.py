"""Few eager C4 SMI steps (for ncu): python scripts/prof_c4.py [steps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import smi_experiments_b200 as smi
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda")
rng = np.random.default_rng(0)
model = smi.mnist_mlp()
eng = smi.SteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.05)), smi.Adagrad(0.8),
                  smi.RBFKernel(), num_stein_particles=10, num_elbo_particles=55, use_cuda_graph=False)
xb = torch.from_numpy(rng.integers(0, 256, (128, 784)).astype(np.float32) / 255.0).to(dev)
yb = torch.from_numpy(rng.integers(0, 10, 128).astype(np.int32)).to(dev)
st = eng.init(smi.PRNGKey(4), xb, yb, n=60000, batch_size=128)
for _ in range(steps):
    st, loss = eng.update(st, xb, yb, n=60000, batch_size=128)
torch.cuda.synchronize()
print("loss", float(loss))
