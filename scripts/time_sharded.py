"""Per-kernel breakdown of the sharded C2 step under torchrun (weak scaling: 20 particles per GPU):
torchrun --nproc-per-node N scripts/time_sharded.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import smi_experiments_b200 as smi
from smi_experiments_b200.dist import ShardedSteinVI
from bench import make_uci, C2, profile_report
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
dev = torch.device("cuda", local)
xt, yt = make_uci()
xd, yd = torch.from_numpy(xt).to(dev), torch.from_numpy(yt).to(dev)
model = smi.uci_bnn(13)
eng = ShardedSteinVI(model, smi.AutoNormal(model, init_loc_fn=smi.init_to_uniform(radius=0.1)), smi.Adam(5e-3),
                     smi.RBFKernel(), num_stein_particles=C2["m"] * world, num_elbo_particles=C2["S"])
st = eng.init(smi.PRNGKey(1), x=xd, y=yd, subsample=100)
kw = dict(x=xd, y=yd, subsample=100)
for _ in range(5):
    st, loss = eng.update(st, **kw)
torch.cuda.synchronize(); dist.barrier()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50):
    st, loss = eng.update(st, **kw)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 50
lib = smi._lib.load()
g = eng.use_cuda_graph
eng.use_cuda_graph = False
lib.smi_profile(1)
ta, tb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ta.record()
for _ in range(5):
    st, loss = eng.update(st, **kw)
tb.record()
torch.cuda.synchronize()
lib.smi_profile(0)
prof = profile_report(lib)
if rank == 0:
    print("N=%d m=%d: %.3f ms/step (as run by update), eager %.3f ms/step" % (world, C2["m"] * world, ms, ta.elapsed_time(tb) / 5))
    tot = 0.0
    for k, (cnt, t) in sorted(prof.items(), key=lambda kv: -kv[1][1]):
        print("    %-28s x%-3d %8.3f ms/step" % (k, cnt // 5, t / 5)); tot += t / 5
    print("    kernels total %.3f ms/step (rest = NCCL + launch gaps)" % tot)
dist.destroy_process_group()
