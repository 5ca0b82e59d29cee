"""Phase trace of the one-launch Stein kernel (build with SMI_NVCC_EXTRA=-DSMI_STEIN_TRACE):
python scripts/stein_trace.py m d"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import smi_experiments_b200 as smi
from smi_experiments_b200 import ops
m, d = int(sys.argv[1]), int(sys.argv[2])
g = torch.Generator(device="cuda").manual_seed(0)
X = torch.randn(m, d, device="cuda", generator=g)
G = torch.randn(m, d, device="cuda", generator=g)
A = torch.randn(8192, 8192, device="cuda")
for _ in range(60):   # bring the clocks up before tracing
    A @ A
torch.cuda.synchronize()
for _ in range(3):
    out = ops.stein_force(X, G)
    torch.cuda.synchronize()
    print("--")
