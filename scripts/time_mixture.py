"""One rank's share of the weak-scaled C2 step on ONE GPU: ELBO gradients of `nrows` particles against all m mixture
components (per-kernel times from libsmi's event profiler).  python scripts/time_mixture.py [m] [nrows] [reps]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import smi_experiments_b200 as smi
from smi_experiments_b200 import models, _lib, ops
from bench import make_uci
m = int(sys.argv[1]) if len(sys.argv) > 1 else 160
nrows = int(sys.argv[2]) if len(sys.argv) > 2 else 20
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
S, B = 100, 100
dev = torch.device("cuda")
md, lay = models.make_desc(smi.uci_bnn(13), _lib.GUIDE_NORMAL)
L, d = lay.latent_dim, lay.particle_dim
rng = np.random.default_rng(0)
psi = rng.uniform(-0.1, 0.1, (m, d)).astype(np.float32)
for r in lay.rows:
    psi[:, r["rho"]:r["rho"] + r["size"]] = np.log(np.expm1(0.1))
xt, yt = make_uci()
psi, x, y = torch.from_numpy(psi).to(dev), torch.from_numpy(xt).to(dev), torch.from_numpy(yt).to(dev)
key, kout = torch.tensor([0, 1], dtype=torch.int32, device=dev), torch.zeros(2, dtype=torch.int32, device=dev)
eps = torch.empty(nrows, S, L, device=dev)
idx = torch.empty(nrows, S, B, dtype=torch.int32, device=dev)
G, E = torch.empty(nrows, d, device=dev), torch.empty(nrows, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = _lib.load()
def step():
    ops.prng_step(md, key, kout, m, S, 455, B, 0, nrows, eps, idx)
    ops.elbo_grad(md, psi, eps, idx, x, y, S, B, 4.55, 1.0, 0, nrows, G, E)
for _ in range(3):
    step()
torch.cuda.synchronize()
evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
for a, b in evs:       # wall time of the launch sequence (kernels on libsmi's internal side stream overlap)
    flush.zero_()
    a.record()
    step()
    b.record()
torch.cuda.synchronize()
wall = sum(a.elapsed_time(b) for a, b in evs) / reps * 1e3
lib.smi_profile(1)
for _ in range(reps):
    flush.zero_()
    step()
torch.cuda.synchronize()
lib.smi_profile(0)
buf = C.create_string_buffer(1 << 16)
lib.smi_profile_report(buf, len(buf))
tot = 0.0
for ln in buf.value.decode().splitlines():
    nm, cnt, ms = ln.split()
    print("%-28s %7.1f us" % (nm, float(ms) / reps * 1e3))
    tot += float(ms) / reps * 1e3
print("%-28s %7.1f us   (m=%d, rows=%d)" % ("sum", tot, m, nrows))
print("%-28s %7.1f us   (CUDA events around the eager launch sequence)" % ("wall", wall))
