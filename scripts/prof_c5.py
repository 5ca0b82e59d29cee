"""One Stein force at the C5 sweep point (m = 16384, d = 10000) after a warm-up call: the target of
`ncu --set full -k regex:gemm3xtf32 --launch-skip 2 -c 2` (profiles/r01i_*)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import smi_experiments_b200 as smi
m = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
dev = torch.device("cuda")
gen = torch.Generator(device=dev).manual_seed(4)
X = torch.randn(m, 10000, device=dev, generator=gen)
G = torch.randn(m, 10000, device=dev, generator=gen)
F = torch.empty_like(X)
for _ in range(2):
    smi.ops.stein_force(X, G, out=F, path=smi.ops.PATH_TC)
torch.cuda.synchronize()
print("done", float(F.double().norm()))
